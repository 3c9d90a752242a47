"""Oracle: latent / noise trajectory construction (torch CPU, fp32), step by step like the reference — full
[T, 18, 512] tensors are smoothed and merged in place, nothing is folded into weights.  TEST INFRASTRUCTURE ONLY.

Restates ssar/random/latent.py:7-80, ssar/random/noise.py:4-140, ssar/random/patch.py:12-133 and
ssar/models/latent_n_noise2.py:134-178.  torchcubicspline (un-vendored, unpinned -> PARITY UNPINNED at this boundary)
is restated from its published algorithm in `natural_cubic_spline` / `spline_evaluate`.
"""
from __future__ import annotations

import numpy as np
import torch

from .post import gaussian_filter

UNITFEATS = ["rms", "drop_strength", "onsets", "spectral_flatness"]
ALLFEATS = ["chromagram", "tonnetz", "mfcc", "spectral_contrast"] + UNITFEATS


# ----------------------------------------------------------------------------------------------
# natural cubic spline (torchcubicspline's parameterisation), fp32 Thomas solve along the knot axis
# ----------------------------------------------------------------------------------------------
def natural_cubic_spline(t, y):
    """t [n] knots, y [n, ...] values -> (a, b, two_c, three_d) each [n-1, ...]."""
    n = len(t)
    shape = y.shape[1:]
    y = y.reshape(n, -1)
    h = t[1:] - t[:-1]
    dy = y[1:] - y[:-1]
    if n == 2:
        z = torch.zeros_like(dy)
        return tuple(v.reshape((1,) + shape) for v in (y[:1], dy / h[:, None], z, z))
    diag = torch.zeros(n)
    diag[:-1] += 2 / h
    diag[1:] += 2 / h
    rhs = torch.zeros_like(y)
    rhs[:-1] += 3 * dy / (h * h)[:, None]
    rhs[1:] += 3 * dy / (h * h)[:, None]
    off = 1 / h
    d = diag.clone()
    r = rhs.clone()
    for i in range(1, n):  # forward elimination
        w = off[i - 1] / d[i - 1]
        d[i] = d[i] - w * off[i - 1]
        r[i] = r[i] - w * r[i - 1]
    k = torch.zeros_like(y)
    k[-1] = r[-1] / d[-1]
    for i in range(n - 2, -1, -1):
        k[i] = (r[i] - off[i] * k[i + 1]) / d[i]
    hh = h[:, None]
    two_c = (6 * dy / hh - 4 * k[:-1] - 2 * k[1:]) / hh
    three_d = (-6 * dy / hh + 3 * (k[:-1] + k[1:])) / (hh * hh)
    return tuple(v.reshape((n - 1,) + shape) for v in (y[:-1], k[:-1], two_c, three_d))


def spline_evaluate(t_knots, coeffs, t):
    a, b, two_c, three_d = coeffs
    idx = (torch.bucketize(t, t_knots) - 1).clamp(0, len(t_knots) - 2)
    f = (t - t_knots[idx]).reshape((-1,) + (1,) * (a.dim() - 1))
    return a[idx] + (b[idx] + (0.5 * two_c[idx] + three_d[idx] * f / 3) * f) * f


def spline_loop_latents(y, size, n_loops=1):
    """random/latent.py:7-13."""
    y = torch.cat((y, y[[0]]))
    t_in = torch.linspace(0, 1, len(y))
    t_out = torch.linspace(0, n_loops, size) % 1
    return spline_evaluate(t_in, natural_cubic_spline(t_in, y), t_out)


# ----------------------------------------------------------------------------------------------
# latent sub-patches
# ----------------------------------------------------------------------------------------------
LAYS = {"low": slice(0, 6), "mid": slice(6, 12), "high": slice(12, 18), "lowmid": slice(0, 12), "midhigh": slice(6, 18),
        "all": slice(0, 18)}


def latent_patch(rng, latents, palette, segmentations, features, tempo, fps, patch_type, segments, loop_bars, seq_feat,
                 seq_feat_weight, mod_feat, mod_feat_weight, merge_type, merge_depth):
    """random/latent.py:16-80 (in place on `latents`)."""
    feature = seq_feat_weight * features[seq_feat]
    seg = segmentations[(seq_feat, segments)]
    perm = torch.randperm(len(palette), generator=rng)
    if patch_type == "segmentation":
        seq = gaussian_filter(palette[perm[:segments][seg]], 5)
    elif patch_type == "feature":
        n = feature.shape[1]
        if n == 1:
            p0, p1 = palette[perm[0]], palette[perm[1]]
            seq = feature[..., None] * p0[None] + (1 - feature[..., None]) * p1[None]
        else:
            seq = torch.einsum("TN,NWL->TWL", feature, palette[perm[:n]])
    else:
        n_loops = len(latents) / fps / 60 / tempo / 4 / loop_bars
        seq = spline_loop_latents(palette[perm[:segments]], len(latents), n_loops)
    seq = gaussian_filter(seq, 1)
    lays = LAYS[merge_depth]
    if merge_type == "average":
        latents[:, lays] = (latents[:, lays] + seq[:, lays]) / 2
    elif merge_type == "modulate":
        m = mod_feat_weight * features[mod_feat][..., None]
        latents[:, lays] = latents[:, lays] * (1 - m) + m * seq[:, lays]
    else:
        latents[:, lays] = seq[:, lays]
    return latents


# ----------------------------------------------------------------------------------------------
# noise modules (lazy, evaluated per batch of frames)
# ----------------------------------------------------------------------------------------------
class Loop:
    def __init__(self, rng, length, size, n_loops=1, sigma=5):
        self.length, self.size, self.sigma = length, size, sigma
        self.noise = torch.randn((3, size[0], size[1]), generator=rng)
        self.idx = torch.linspace(0, n_loops * 2 * torch.pi, length)

    def __call__(self, i, b):
        ph = torch.cos(self.idx[i : i + b, None, None] + self.noise[[0]]) / (self.sigma / 50)
        out = torch.sin(ph + self.noise[[1]]) * self.noise[[2]]
        return out / (out.square().mean(dim=(1, 2), keepdim=True).sqrt() + torch.finfo(out.dtype).eps)


class Blend:
    def __init__(self, rng, length, size, modulator):
        self.length, self.size, self.mod = length, size, modulator
        self.noise = torch.randn((2, modulator.shape[1], size[0], size[1]), generator=rng)

    def __call__(self, i, b):
        m = self.mod[i : i + b].reshape(-1, self.mod.shape[1])
        return torch.einsum("MHW,BM->BHW", self.noise[0], m) + torch.einsum("MHW,BM->BHW", self.noise[1], 1 - m)


class Multiply:
    def __init__(self, rng, length, size, modulator):
        self.length, self.size, self.mod = length, size, modulator
        self.noise = torch.randn((modulator.shape[1], size[0], size[1]), generator=rng)

    def __call__(self, i, b):
        return torch.einsum("MHW,BM->BHW", self.noise, self.mod[i : i + b].reshape(-1, self.mod.shape[1]))


class Average:
    def __init__(self, left, right):
        self.length, self.size, self.left, self.right = left.length, left.size, left, right

    def __call__(self, i, b):
        return (self.left(i, b) + self.right(i, b)) / 2


class Modulate:
    def __init__(self, left, right, modulator):
        self.length, self.size, self.left, self.right = left.length, left.size, left, right
        self.mod = modulator.mean(1)

    def __call__(self, i, b):
        m = self.mod[i : i + b, None, None]
        return self.left(i, b) * m + self.right(i, b) * (1 - m)


class ScaleBias:
    def __init__(self, base, scale, bias):
        self.length, self.size, self.base, self.scale, self.bias = base.length, base.size, base, scale, bias

    def __call__(self, i, b):
        return self.scale * self.base(i, b) + self.bias


NOISE_LAYS = {"low": range(0, 6), "mid": range(6, 12), "high": range(12, 17), "lowmid": range(0, 12),
              "midhigh": range(6, 17), "all": range(0, 17)}


def noise_patch(rng, noise, features, tempo, fps, patch_type, loop_bars, seq_feat, seq_feat_weight, mod_feat,
                mod_feat_weight, merge_type, merge_depth, noise_mean, noise_std):
    """random/noise.py:89-140."""
    feature = seq_feat_weight * features[seq_feat]
    for n in NOISE_LAYS[merge_depth]:
        if patch_type == "blend":
            new = Blend(rng, len(feature), noise[n].size, feature)
        elif patch_type == "multiply":
            new = Multiply(rng, len(feature), noise[n].size, feature)
        else:
            new = Loop(rng, len(feature), noise[n].size, n_loops=len(feature) / fps / 60 / tempo / 4 / loop_bars)
        if merge_type == "average":
            noise[n] = Average(noise[n], new)
        elif merge_type == "modulate":
            noise[n] = Modulate(noise[n], new, mod_feat_weight * features[mod_feat])
        else:
            noise[n] = new
        noise[n] = ScaleBias(noise[n], noise_std, noise_mean)
    return noise


# ----------------------------------------------------------------------------------------------
# the seeded sampler and Patch.forward (random/patch.py:12-133); CPU generator
# ----------------------------------------------------------------------------------------------
def _choice(rng, options, weights=None):
    p = torch.ones(len(options)) / len(options) if weights is None else torch.tensor(weights) / np.sum(weights)
    return options[p.multinomial(num_samples=1, replacement=False, generator=rng)]


def _skewnorm(rng, a, loc, scale):
    u0 = torch.randn((), generator=rng)
    v = torch.randn((), generator=rng)
    d = a / np.sqrt(1 + a ** 2)
    u1 = d * u0 + v * np.sqrt(1 - d ** 2)
    return (loc + scale * torch.where(u0 >= 0, u1, -u1)).item()


class Patch:
    def __init__(self, features, segmentations, tempo, fps=24, seed=42, min_subpatches=5, max_subpatches=20):
        rng = torch.Generator("cpu")
        rng.manual_seed(seed)
        self.rng, self.seed, self.fps, self.tempo = rng, seed, fps, tempo
        self.features, self.segmentations = features, segmentations
        self.length = next(iter(features.values())).shape[0]
        self.num_base_latents = torch.randint(3, 15, size=(), generator=rng).item()
        self.sigma_base_noise = 1 + 9 * torch.rand((), generator=rng).item()
        self.loops_base_noise = _choice(rng, [1, 2, 4, 8, 16, 32, 64])
        self.ks = np.unique([k for (_, k) in segmentations]).tolist()
        depth = (["low", "mid", "high", "lowmid", "midhigh", "all"], [3, 3, 3, 2, 2, 1])

        def common(kinds):
            return dict(
                patch_type=_choice(rng, kinds),
                **({"segments": _choice(rng, self.ks)} if kinds[0] == "segmentation" else {}),
                loop_bars=_choice(rng, [4, 8, 16, 32], [2, 2, 2, 1]),
                seq_feat=_choice(rng, ALLFEATS),
                seq_feat_weight=_skewnorm(rng, 5, 0.666, 0.5),
                mod_feat=_choice(rng, UNITFEATS),
                mod_feat_weight=_skewnorm(rng, 5, 0.666, 0.5),
                merge_type=_choice(rng, ["average", "modulate"], [1, 3]),
                merge_depth=_choice(rng, *depth),
            )

        self.latent_patches = [common(["segmentation", "feature", "loop"])
                               for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng))]
        self.noise_patches = []
        for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng)):
            d = common(["blend", "multiply", "loop"])
            d["noise_mean"] = torch.randn((), generator=rng).item() * 0.5
            d["noise_std"] = _skewnorm(rng, 5, 0.666, 0.5)
            self.noise_patches.append(d)

    def forward(self, palette, downscale_factor=1, aspect_ratio=1):
        self.rng.manual_seed(self.seed)
        base = torch.randperm(len(palette), generator=self.rng)[: self.num_base_latents]
        latents = spline_loop_latents(palette[base], self.length)
        for sp in self.latent_patches:
            latents = latent_patch(self.rng, latents, palette, self.segmentations, self.features, self.tempo, self.fps, **sp)
        noise = [Loop(self.rng, self.length, (round(aspect_ratio * s / downscale_factor), round(s / downscale_factor)),
                      n_loops=self.loops_base_noise, sigma=self.sigma_base_noise)
                 for s in [4, 8, 8, 16, 16, 32, 32, 64, 64, 128, 128, 256, 256, 512, 512, 1024, 1024]]
        for sp in self.noise_patches:
            noise = noise_patch(self.rng, noise, self.features, self.tempo, self.fps, **sp)
        return latents, noise


def fixed_decoder(latents, x, hidden_size, n_latent_split, noise_draws, sigma=5):
    """models/latent_n_noise2.py:146-178 with the standard-normal draws passed in (RNG stays outside)."""
    S, H = n_latent_split, hidden_size
    Wn = latents.shape[1] // S
    outs = []
    for i in range(S):
        env = x[..., i * H : (i + 1) * H]
        env = env / env.sum(dim=-1, keepdim=True)
        outs.append(torch.einsum("BTH,HWL->BTWL", env, latents[i * H : (i + 1) * H, i * Wn : (i + 1) * Wn]))
    lat = torch.cat(outs, dim=2)
    envs = x[..., S * H :]
    B, T, _ = x.shape
    noise = []
    for i in range(envs.shape[-1] // 2):
        mu, sig = envs[..., 2 * i], envs[..., 2 * i + 1]
        z = gaussian_filter(noise_draws[i].transpose(1, 0), sigma).transpose(1, 0)
        noise.append(mu[..., None, None] + sig[..., None, None] * z)
    return lat, noise


class LatentAugmenter:
    """ssar/supervised/latent_augmenter.py:23-105 with the mapped-latent bank `ws` [num, W, L] as an argument (the
    reference maps 16 384 random z through a StyleGAN2 checkpoint, :25-29).  Draws come from numpy's and Python's GLOBAL
    generators in the reference's order (np.random.seed / random.seed reproduce a run).  `normalize` is
    features/processing.py's (the module asks supervised/data.py for it, which does not define one)."""

    FEAT_IDXS = {"chroma": (20, 32), "tonnetz": (32, 38), "onsets": (46, 47), "onsets_low": (47, 48), "onsets_mid": (48, 49),
                 "onsets_high": (49, 50), "volume": (51, 52), "volume_low": (52, 53), "volume_mid": (53, 54),
                 "volume_high": (54, 55), "volume_long": (55, 56), "volume_low_long": (56, 57),
                 "volume_mid_long": (57, 58), "volume_high_long": (58, 59)}

    def __init__(self, ws, n_patches):
        self.ws, self.n_patches = ws, n_patches
        self.num, self.nw = ws.shape[0], ws.shape[1]
        self.feat_idxs = dict(self.FEAT_IDXS)  # the literal's repeated keys collapse (latent_augmenter.py:31-56)
        self.feat_keys = list(self.feat_idxs)
        self.single_dim = -12

    def random_patch(self, feature):
        import random

        from .post import normalize

        n_knots = np.random.randint(3, 12)  # evaluated before the call it is an argument of (:70-71)
        latent = spline_loop_latents(self.ws[np.random.randint(0, self.num, n_knots)].to(feature), len(feature))
        for _ in range(self.n_patches):
            start, stop = self.feat_idxs[random.choice(self.feat_keys)]
            if np.random.rand() > 0.5:
                lay_start = np.random.randint(0, self.nw - 6)
                lay_stop = np.random.randint(lay_start, self.nw)
            else:
                lay_start, lay_stop = 0, self.nw
            if stop - start == 1:
                lat = self.ws[np.random.randint(0, self.num, 1)].to(feature)
                m = normalize(feature[:, start:stop, None])
                latent[:, lay_start:lay_stop] *= 1 - m
                latent[:, lay_start:lay_stop] += m * lat[:, lay_start:lay_stop]
            else:
                lats = self.ws[np.random.randint(0, self.num, stop - start)].to(feature)
                m = normalize(feature[:, start:stop])
                m = m / (m.sum(1, keepdim=True) + 1e-8)
                patch_latent = torch.einsum("Atwl,Atwl->twl", m.permute(1, 0)[..., None, None], lats[:, None])
                if np.random.rand() > 0.666:
                    a, b = self.feat_idxs[random.choice(self.feat_keys[self.single_dim:])]
                    im = normalize(feature[:, a:b, None])
                    latent[:, lay_start:lay_stop] *= 1 - im
                    latent[:, lay_start:lay_stop] += im * patch_latent[:, lay_start:lay_stop]
                else:
                    latent[:, lay_start:lay_stop] = patch_latent[:, lay_start:lay_stop]
        offset = latent.mean(dim=(0, 1), keepdim=True)
        return latent - offset, offset
