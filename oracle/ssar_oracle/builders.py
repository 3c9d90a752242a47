"""Oracle: the latent / noise builders of ssar/features/generate.py:155-230 and the per-song fixed decoder of
ssar/optimize.py:103-142 (torch CPU, fp32).  TEST INFRASTRUCTURE ONLY.

Neither module can be imported here (generate.py loads a StyleGAN2 pickle and decord at
import, optimize.py imports maua + matplotlib), so these are restatements of the cited lines — PINNED since round 2 by
tests/test_oracle_widened.py against fixtures made by compiling and running the reference's unmodified function text
(tests/golden/ref_loader.py::exec_reference_defs).  Random draws and the mapped
latents `ws` are arguments (the reference draws them with torch.randn / G.mapper)."""
from __future__ import annotations

import torch

from .post import clamp_lower_percentile, clamp_peaks_percentile, gaussian_filter, normalize


def _chroma(chroma, fps):
    c = gaussian_filter(chroma, fps / 24)
    c = clamp_peaks_percentile(c, 97)
    c = clamp_lower_percentile(c, 5)
    return c / c.sum(1, keepdim=True)


def _onsets(ons, fps):
    o = gaussian_filter(ons.squeeze(), fps / 24)
    o = clamp_peaks_percentile(o, 97)
    o = clamp_lower_percentile(o, 5)
    return normalize(o).reshape(-1)


def high_chroma_correlation(chroma, ws, fps, noise_draw):
    """generate.py:155-174."""
    c = _chroma(chroma, fps)
    latents = gaussian_filter(torch.einsum("TC,CNL->TNL", c, ws), fps / 24)
    noise = gaussian_filter(noise_draw, fps)
    return latents, noise / noise.std()


def high_onset_correlation(ons, ws, fps, noise_draw):
    """generate.py:177-199."""
    o = _onsets(ons, fps)
    latents = ws[[0]] * o.reshape(-1, 1, 1) + ws[[1]] * (1 - o.reshape(-1, 1, 1))
    latents = gaussian_filter(latents, fps / 24)
    noise = gaussian_filter(noise_draw, fps / 4)
    noise = noise / noise.std(dim=(1, 2, 3), keepdim=True)
    return latents, noise + noise * o.reshape(-1, 1, 1, 1)


def high_both_correlation(chroma, ons, ws, fps, noise_draw):
    """generate.py:202-230."""
    c, o = _chroma(chroma, fps), _onsets(ons, fps)
    cl = torch.einsum("TC,CNL->TNL", c, ws[:12])
    ol = ws[[12]] * o.reshape(-1, 1, 1) + ws[[13]] * (1 - o.reshape(-1, 1, 1))
    latents = gaussian_filter((cl + ol) / 2, fps / 24)
    noise = gaussian_filter(noise_draw, fps / 4)
    noise = noise / noise.std(dim=(1, 2, 3), keepdim=True)
    return latents, noise + noise * o.reshape(-1, 1, 1, 1)


def optimize_fixed_decoder(latents, x, S, G, H, noise_draws):
    """optimize.py:117-142 with the standard-normal draws passed in."""
    Wn = latents.shape[1] // S
    outs = []
    for i in range(S):
        env = x[:, i * (G * H) : (i + 1) * (G * H)].reshape(x.shape[0], G, H).softmax(dim=2)
        env = env / (env.sum(dim=(1, 2), keepdim=True) + 1e-8)
        lat = latents[i * (G * H) : (i + 1) * (G * H), i * Wn : (i + 1) * Wn]
        lat = lat.reshape(G, H, lat.shape[-2], lat.shape[-1])
        outs.append(torch.einsum("TGH,GHWL->TWL", env, lat))
    lat = torch.cat(outs, dim=1)
    envs = x[:, S * G * H :]
    noise = []
    for i in range(envs.shape[-1] // 2):
        mu, sig = envs[:, 2 * i], envs[:, 2 * i + 1]
        n = gaussian_filter(noise_draws[i], 2)
        noise.append(mu[:, None, None] + sig[:, None, None] * n)
    return lat, noise
