"""Oracle: post-processing of feature matrices (torch CPU / numpy, fp32).  TEST INFRASTRUCTURE ONLY.

Restates ssar/features/processing.py (gaussian_filter :10-48, normalize :52-55, standardize :58-61, spectral_flux
:88-89, onset_envelope :92-98, percentile clamps :101-129, emphasize :132-138), the efficient_quantile wrapper
(ssar/analysis/efficient_quantile/__init__.py:5-6 + efficient_quantile.cpp:86-206, midpoint mode) and the dataset
statistics of ssar/supervised/data.py:317-319.
"""
from __future__ import annotations

import numpy as np
import torch


def gaussian_taps(sigma, radius):
    """processing.py:21-24."""
    k = torch.arange(-radius, radius + 1, dtype=torch.float32)
    k = torch.exp(-0.5 / sigma ** 2 * k ** 2)
    return k / k.sum()


def gaussian_filter(x, sigma, mode="circular"):
    """processing.py:10-48 as an explicit gather: out[t] = sum_k w[k] x[idx(t + k - r)] along dim 0."""
    dim = x.dim()
    T = x.shape[0]
    radius = min(int(sigma * 4), 3 * T)
    w = gaussian_taps(sigma, radius)
    v = torch.arange(T)[:, None] + torch.arange(-radius, radius + 1)[None, :]  # [T, taps] virtual indices
    if radius > T:  # short-sequence branch: mode-pad by T, then replicate-pad the rest (:30-35)
        v = v.clamp(-T, 2 * T - 1)
    if mode == "circular":
        idx = v % T
    elif mode == "reflect":
        idx = torch.where(v < 0, -v, v)
        idx = torch.where(idx >= T, 2 * (T - 1) - idx, idx)
    elif mode == "replicate":
        idx = v.clamp(0, T - 1)
    else:
        raise ValueError(mode)
    flat = x.reshape(T, -1).float()
    out = torch.zeros_like(flat)
    for k in range(2 * radius + 1):  # same accumulation order as a direct convolution
        out += w[k] * flat[idx[:, k]]
    out = out.reshape(x.shape)
    if dim < 3:
        out = out.reshape(tuple(x.shape) + (1,) * (3 - dim)).squeeze()
    return out


def normalize(a):
    """processing.py:52-55."""
    a = a - a.min()
    return a / (a.max() + 1e-8)


def quantile_midpoint(t, q):
    """efficient_quantile mode 3 over the flattened, NaN-free tensor: q is rounded to fp32 first (wrapper builds a
    FloatTensor), rank = q * (n - 1) in fp64, result = mean of the floor / ceil order statistics."""
    v = np.sort(t.detach().cpu().numpy().astype(np.float32).ravel())
    v = v[~np.isnan(v)]
    if v.size == 0:
        return torch.tensor(float("nan"))
    rank = float(np.float32(q)) * (v.size - 1)
    lo, hi = int(rank), int(np.ceil(rank))
    a, b = float(v[lo]), float(v[hi])
    return torch.tensor(np.float32(a + (0.5 if hi > lo else 0.0) * (b - a)))


def quantile_linear(v, q):
    """torch.quantile(v, q) for 1-D fp32 v: rank = fp32(q) * (n - 1) in fp32, lerp of the neighbours in fp32."""
    s = np.sort(v.detach().cpu().numpy().astype(np.float32))
    if s.size == 0:
        raise RuntimeError("quantile() input tensor must be non-empty")
    rank = np.float32(q) * np.float32(s.size - 1)
    below = np.floor(rank)
    w = np.float32(rank - below)
    a, b = s[int(below)], s[int(np.ceil(rank))]
    d = np.float32(b - a)
    return torch.tensor(a + w * d if w < 0.5 else b - d * np.float32(1 - w), dtype=torch.float32)


def standardize(a):
    """processing.py:58-61."""
    return normalize(torch.clamp(a, quantile_midpoint(a, 0.25), quantile_midpoint(a, 0.75) + 1e-10))


def spectral_flux(spec):
    """processing.py:88-89."""
    return torch.cat([spec[1:], torch.zeros(1, spec.shape[1])]) - spec


def onset_envelope(flux):
    """processing.py:92-98."""
    u = (0.5 * (flux + flux.abs())).sum(1)
    u = torch.clamp(u, quantile_midpoint(u, 0.025), quantile_midpoint(u, 0.975))
    u = u - u.min()
    return u / u.max()


def clamp_peaks_percentile(signal, percent):
    """processing.py:101-121: strict interior local maxima only; edges can never be peaks."""
    if signal.dim() < 2:
        signal = signal.unsqueeze(1)
    cols = []
    for sig in signal.unbind(1):
        peak = torch.zeros(len(sig), dtype=torch.bool)
        peak[1:-1] = (sig[1:-1] > sig[2:]) & (sig[1:-1] > sig[:-2])
        cols.append(torch.clamp(sig, max=quantile_linear(sig[peak], percent / 100)))
    return torch.stack(cols, 1)


def clamp_upper_percentile(signal, percentile):
    """processing.py:124-125."""
    s = signal if signal.dim() >= 2 else signal[:, None]
    thr = torch.stack([quantile_linear(c, percentile / 100) for c in s.reshape(s.shape[0], -1).unbind(1)])
    return torch.minimum(signal, thr.reshape(signal.shape[1:]))


def clamp_lower_percentile(signal, percentile):
    """processing.py:128-129."""
    s = signal if signal.dim() >= 2 else signal[:, None]
    thr = torch.stack([quantile_linear(c, percentile / 100) for c in s.reshape(s.shape[0], -1).unbind(1)])
    return torch.maximum(signal, thr.reshape(signal.shape[1:]))


def emphasize(envs, strength, percentile):
    """processing.py:132-138."""
    e = envs if envs.dim() >= 2 else envs[:, None]
    mn = e.min(0).values
    x = e - mn
    mx = x.max(0).values
    x = x / mx
    qv = torch.stack([quantile_linear(c, percentile / 100) for c in x.unbind(1)])
    x = x * (1 + torch.tanh(strength * (x - qv)))
    return (x * mx + mn).reshape(envs.shape)


def feature_stats(feats):
    """supervised/data.py:317-319: mean and population std per feature over every window and frame (fp64 here)."""
    f = feats.reshape(-1, feats.shape[-1]).double().numpy()
    return np.mean(f, 0), np.std(f, 0)


def overlapping_slices(tensor, length):
    """supervised/data.py:255-258: for start in range(0, L, L // 4): the chunks of L rows of tensor[start:], the last chunk
    of every start dropped; stacked in that order."""
    out = []
    for start in range(0, length, length // 4):
        out += list(torch.split(tensor[start:], length)[:-1])
    return torch.stack(out)
