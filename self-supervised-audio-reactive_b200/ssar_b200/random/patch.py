"""Mirror of ssar/random/patch.py — the seeded random-patch sampler and its forward pass.

All random draws stay in torch host code in the reference's order (patch.py:41-66, 109-131; noise.py:15,31,46) so a
given seed yields the same patch; the arithmetic runs in the trajectory kernels."""
from __future__ import annotations

import json

import numpy as np
import torch

from . import DEVICE
from .latent import LatentWeights, latent_patch, merge_coefficients, sequence_weights, spline_loop_latents  # noqa: F401
from .mir import ALLFEATS, UNITFEATS
from .noise import Loop, noise_patch
from .. import plan


def random_choice(rng, options, weights=None, n=1, replacement=False):
    """patch.py:12-20."""
    if weights is None:
        probabilities = torch.ones(len(options), device=rng.device) / len(options)
    else:
        probabilities = torch.tensor(weights, device=rng.device) / np.sum(weights)
    idx = probabilities.multinomial(num_samples=n, replacement=replacement, generator=rng)
    return options[idx]


def skewnorm(rng, a, loc, scale, size=()):
    """patch.py:23-32."""
    u0 = torch.randn(size, generator=rng, device=rng.device)
    v = torch.randn(size, generator=rng, device=rng.device)
    d = a / np.sqrt(1 + a ** 2)
    u1 = d * u0 + v * np.sqrt(1 - d ** 2)
    return loc + scale * torch.where(u0 >= 0, u1, -u1)


class Patch:
    """patch.py:35-175.  `device`: where the generator lives (as in the reference); `compute_device`: where the
    trajectory tensors live (defaults to the palette's device at forward time)."""

    def __init__(self, features, segmentations, tempo, fps=24, seed=42, min_subpatches=5, max_subpatches=20,
                 device=DEVICE):
        rng = torch.Generator(device)
        rng.manual_seed(seed)
        self.seed = seed
        self.rng = rng
        self.fps = fps
        self.tempo = tempo
        self.length = features[list(features.keys())[0]].shape[0]
        self.features = features
        self.segmentations = segmentations
        self.num_base_latents = torch.randint(3, 15, size=(), generator=rng, device=rng.device).item()
        self.sigma_base_noise = 1 + 9 * torch.rand((), generator=rng, device=rng.device).item()
        self.loops_base_noise = random_choice(rng, [1, 2, 4, 8, 16, 32, 64])
        self.ks = np.unique([k for (_, k) in segmentations]).tolist()
        self.latent_patches = [
            self.random_latent_patch()
            for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device))
        ]
        self.noise_patches = [
            self.random_noise_patch()
            for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device))
        ]

    def update_intensity(self, val):
        """patch.py:68-75 (including the write of noise_std into latent_patches[p], as in the reference)."""
        for p in range(len(self.latent_patches)):
            self.latent_patches[p]["seq_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
            self.latent_patches[p]["mod_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
        for p in range(len(self.noise_patches)):
            self.noise_patches[p]["seq_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
            self.noise_patches[p]["mod_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
            self.latent_patches[p]["noise_std"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()

    def random_latent_patch(self):
        return dict(
            patch_type=random_choice(self.rng, ["segmentation", "feature", "loop"]),
            segments=random_choice(self.rng, self.ks),
            loop_bars=random_choice(self.rng, [4, 8, 16, 32], weights=[2, 2, 2, 1]),
            seq_feat=random_choice(self.rng, ALLFEATS),
            seq_feat_weight=skewnorm(self.rng, a=5, loc=0.666, scale=0.5).item(),
            mod_feat=random_choice(self.rng, UNITFEATS),
            mod_feat_weight=skewnorm(self.rng, a=5, loc=0.666, scale=0.5).item(),
            merge_type=random_choice(self.rng, ["average", "modulate"], weights=[1, 3]),
            merge_depth=random_choice(
                self.rng, ["low", "mid", "high", "lowmid", "midhigh", "all"], weights=[3, 3, 3, 2, 2, 1]
            ),
        )

    def random_noise_patch(self):
        return dict(
            patch_type=random_choice(self.rng, ["blend", "multiply", "loop"]),
            loop_bars=random_choice(self.rng, [4, 8, 16, 32], weights=[2, 2, 2, 1]),
            seq_feat=random_choice(self.rng, ALLFEATS),
            seq_feat_weight=skewnorm(self.rng, a=5, loc=0.666, scale=0.5).item(),
            mod_feat=random_choice(self.rng, UNITFEATS),
            mod_feat_weight=skewnorm(self.rng, a=5, loc=0.666, scale=0.5).item(),
            merge_type=random_choice(self.rng, ["average", "modulate"], weights=[1, 3]),
            merge_depth=random_choice(
                self.rng, ["low", "mid", "high", "lowmid", "midhigh", "all"], weights=[3, 3, 3, 2, 2, 1]
            ),
            noise_mean=torch.randn((), generator=self.rng, device=self.rng.device).item() * 0.5,
            noise_std=skewnorm(self.rng, a=5, loc=0.666, scale=0.5).item(),
        )

    def forward(self, latent_palette, downscale_factor=1, aspect_ratio=1, latents_out=None):
        """patch.py:108-133 -> (latents [T, 18, 512], list of 17 lazy noise modules).
        latents_out: optional preallocated contiguous [T, 18, 512] tensor to render into (batch pipelines)."""
        self.rng.manual_seed(self.seed)
        dev = latent_palette.device
        T, P = self.length, len(latent_palette)
        feats = {k: v.to(dev) for k, v in self.features.items()}

        base_selection = torch.randperm(P, generator=self.rng, device=self.rng.device)[: self.num_base_latents]
        acc = LatentWeights(T, P, latent_palette.shape[1], 6, dev)
        acc.set_base(base_selection, plan.spline_loop_weights_cpu(len(base_selection), T, 1).to(dev))
        for sp in self.latent_patches:
            sel, S = sequence_weights(self.rng, T, P, self.segmentations, feats, self.tempo, self.fps, sp["patch_type"],
                                      sp["segments"], sp["loop_bars"], sp["seq_feat"], sp["seq_feat_weight"], dev)
            alpha, beta = merge_coefficients(feats, sp["merge_type"], sp["mod_feat"], sp["mod_feat_weight"], T, dev)
            acc.merge(sel, S, alpha, beta, sp["merge_depth"])
        latents = acc.render(latent_palette, out=latents_out)

        noise = [
            Loop(rng=self.rng, length=self.length,
                 size=(round(aspect_ratio * size / downscale_factor), round(size / downscale_factor)),
                 n_loops=self.loops_base_noise, sigma=self.sigma_base_noise, device=dev)
            for size in [4, 8, 8, 16, 16, 32, 32, 64, 64, 128, 128, 256, 256, 512, 512, 1024, 1024]
        ]
        for sp in self.noise_patches:
            noise = noise_patch(self.rng, noise, feats, self.tempo, self.fps, **sp)
        return latents, noise

    __call__ = forward

    def save(self, path):
        with open(path, mode="w") as f:
            f.write(json.dumps(dict(
                seed=self.seed, latent_patches=self.latent_patches, noise_patches=self.noise_patches,
                num_base_latents=self.num_base_latents, sigma_base_noise=self.sigma_base_noise,
                loops_base_noise=self.loops_base_noise,
            )))

    @staticmethod
    def load(path, features, segmentations, tempo, fps, device):
        patch = Patch(features, segmentations, tempo, fps, device=device)
        with open(path, mode="r") as f:
            for key, val in json.loads(f.read()).items():
                setattr(patch, key, val)
        return patch
