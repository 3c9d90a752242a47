"""Deterministic synthetic waveforms shared by the tests, the oracle checks and bench.py.

Mono fp32 in [-1, 1], never exactly silent (the reference divides 0/0 on silent chroma frames,
`ssar/features/rosa/spectral.py:264,280`).  Identical bytes are fed to the oracle and to the CUDA path.
"""
from __future__ import annotations

import math

import torch


def synth_clip(n_samples: int, sr: int, clip_index: int = 0, device="cpu") -> torch.Tensor:
    """AM tone + second tone + exponential chirp + white noise + jittered click train, peak-normalised."""
    g = torch.Generator("cpu").manual_seed(1234 + clip_index)
    t = torch.arange(n_samples, dtype=torch.float64) / sr
    dur = max(n_samples / sr, 1e-3)
    f0 = 110.0 * 2.0 ** ((clip_index % 24) / 12.0) * 2.0  # 220 Hz for clip 0
    f1 = f0 * 2.5199  # 554.37 Hz for clip 0
    click_rate = 2.0 if clip_index == 0 else float(1.5 + 1.5 * torch.rand((), generator=g))
    x = 0.30 * torch.sin(2 * math.pi * f0 * t) * (0.5 + 0.5 * torch.sin(2 * math.pi * 2.0 * t))
    x += 0.20 * torch.sin(2 * math.pi * f1 * t)
    # exponential chirp 55 -> 3520 Hz over the clip: phase = 2*pi*f_a*T/ln(k) * (k^(t/T) - 1)
    k = 3520.0 / 55.0
    x += 0.15 * torch.sin(2 * math.pi * 55.0 * dur / math.log(k) * (k ** (t / dur) - 1.0))
    x = x.to(torch.float32)
    x += 0.05 * torch.randn(n_samples, generator=g)
    n_clicks = int(dur * click_rate)
    if n_clicks > 0:
        pos = (torch.arange(n_clicks, dtype=torch.float64) + 0.25) / click_rate
        pos = pos + 0.02 * (torch.rand(n_clicks, generator=g, dtype=torch.float64) - 0.5)
        idx = (pos * sr).long().clamp(0, n_samples - 1)
        for w, a in enumerate([0.8, -0.5, 0.3, -0.15]):  # short decaying click
            x[(idx + w).clamp(max=n_samples - 1)] += a
    x = x / (1.05 * x.abs().max())
    return x.to(device)


def synth_batch(n_clips: int, n_samples: int, sr: int, first_index: int = 0, device="cpu") -> torch.Tensor:
    return torch.stack([synth_clip(n_samples, sr, first_index + i) for i in range(n_clips)]).to(device)


def synth_batch_device(n_clips: int, n_samples: int, sr: int, device, first_index: int = 0) -> torch.Tensor:
    """Same signal family generated directly on `device` (bench only: avoids a multi-GB host build).

    Not bit-identical to `synth_clip` (device RNG); parity tests always use `synth_clip`.
    """
    g = torch.Generator(device).manual_seed(1234 + first_index)
    t = torch.arange(n_samples, dtype=torch.float32, device=device) / sr
    dur = n_samples / sr
    k = 3520.0 / 55.0
    chirp_phase = 2 * math.pi * 55.0 * dur / math.log(k) * (torch.pow(torch.tensor(k, device=device), t / dur) - 1.0)
    out = torch.empty(n_clips, n_samples, dtype=torch.float32, device=device)
    for i in range(n_clips):
        ci = first_index + i
        f0 = 110.0 * 2.0 ** ((ci % 24) / 12.0) * 2.0
        x = 0.30 * torch.sin(2 * math.pi * f0 * t) * (0.5 + 0.5 * torch.sin(2 * math.pi * 2.0 * t))
        x += 0.20 * torch.sin(2 * math.pi * f0 * 2.5199 * t)
        x += 0.15 * torch.sin(chirp_phase)
        x += 0.05 * torch.randn(n_samples, generator=g, device=device)
        rate = 1.5 + 1.5 * ((ci * 0.6180339887) % 1.0)
        n_clicks = int(dur * rate)
        idx = ((torch.arange(n_clicks, device=device, dtype=torch.float32) + 0.25) / rate * sr).long().clamp(0, n_samples - 1)
        x[idx] += 0.8
        out[i] = x / (1.05 * x.abs().max())
    return out


def synth_features(n_frames: int, n_cols: int, seed: int = 0) -> torch.Tensor:
    """Deterministic [T, C] feature-like matrix (smooth random walk + spikes, columns on different scales)."""
    g = torch.Generator("cpu").manual_seed(9000 + seed)
    walk = torch.randn(n_frames, n_cols, generator=g).cumsum(0) * 0.05
    spikes = (torch.rand(n_frames, n_cols, generator=g) > 0.97).float() * torch.rand(n_frames, n_cols, generator=g) * 3
    jitter = 0.3 * torch.randn(n_frames, n_cols, generator=g)
    scale = torch.linspace(0.5, 20.0, n_cols)
    return ((walk + spikes + jitter) * scale + torch.arange(n_cols) * 0.25).float()


PATCH_FEATURE_DIMS = {"chromagram": 12, "tonnetz": 6, "mfcc": 20, "spectral_contrast": 7, "rms": 1, "drop_strength": 1,
                      "onsets": 1, "spectral_flatness": 1}


def synth_patch_inputs(n_frames: int, seed: int = 0, n_palette: int = 180):
    """Deterministic inputs of `Patch` (SURVEY.md 8d cfg 4): normalised [T, C] features per name, random segmentations
    randint(0, k, [T]) for k in {2,4,6,8,12,16} per feature name (+ "rosa"), and a latent palette [P, 18, 512]."""
    feats = {}
    for j, (name, c) in enumerate(PATCH_FEATURE_DIMS.items()):
        f = synth_features(n_frames, c, 40 + 17 * seed + j)
        f = f - f.min()
        feats[name] = f / (f.max() + 1e-8)
    g = torch.Generator("cpu").manual_seed(7000 + seed)
    segs = {}
    for name in list(PATCH_FEATURE_DIMS) + ["rosa"]:
        for k in (2, 4, 6, 8, 12, 16):
            segs[(name, k)] = torch.randint(0, k, (n_frames,), generator=g)
    palette = torch.randn(n_palette, 18, 512, generator=torch.Generator("cpu").manual_seed(3000 + seed))
    return feats, segs, palette


def projection_matrix(n_in: int = 512, n_out: int = 8) -> torch.Tensor:
    """Fixed random projection used to store compact fingerprints of [T, 18, 512] tensors in the golden fixtures."""
    return torch.randn(n_in, n_out, generator=torch.Generator("cpu").manual_seed(555)) / n_in ** 0.5
