"""Mirror of ssar/features/processing.py (filters, normalisation, percentile clamps) — time-first tensors like the
reference ([T], [T, C], [T, C, L], [T, C, H, W]); the batched `[B, T, C]` forms live in `ssar_b200.ops`."""
from __future__ import annotations

import torch

from .. import ops, plan


def gaussian_filter(x, sigma, mode: str = "circular", causal: float = 1):
    """features/processing.py:10-48.  `causal` is accepted and ignored exactly like the reference (:23 is commented out)."""
    dim = x.dim()
    n_frames = x.shape[0]
    radius = min(int(sigma * 4), 3 * n_frames)
    if radius > n_frames:
        print(
            f"WARNING: Gaussian filter radius ({int(sigma * 4)}) is larger than number of frames ({n_frames}).\n\t Filter size has been lowered to ({radius}). You might want to consider lowering sigma ({sigma})."
        )
    out = ops.gauss1d(x, sigma, mode)
    if dim < 3:  # the reference lifts to 3-D and squeezes every singleton on the way out (:13-14, :45-46)
        out = out.reshape(tuple(x.shape) + (1,) * (3 - dim)).squeeze()
    return out


def normalize(array):
    """features/processing.py:52-55."""
    return ops.normalize(array, 1).view(array.shape)


def quantile(tensor, q):
    """ssar/analysis/efficient_quantile/__init__.py:5-6 (midpoint interpolation over the flattened tensor) as a CUDA
    radix select; returns a 0-d tensor on the input's device."""
    flat = tensor.contiguous().view(1, -1, 1)
    return ops.col_quantile(flat, float(q), "all", "midpoint").view(())


def standardize(array):
    """features/processing.py:58-61."""
    flat = array.contiguous().view(1, -1, 1)
    lo = ops.col_quantile(flat, 0.25, "all", "midpoint")
    hi = ops.col_quantile(flat, 0.75, "all", "midpoint") + 1e-10
    return normalize(ops.clamp_cols(flat, lo, hi).view(array.shape))


def spectral_flux(spec):
    """features/processing.py:88-89."""
    return ops.spectral_flux(spec, 1)


def onset_envelope(flux):
    """features/processing.py:92-98."""
    u = ops.hwr_rowsum(flux.contiguous()).view(1, -1, 1)
    lo = ops.col_quantile(u, 0.025, "all", "midpoint")
    hi = ops.col_quantile(u, 0.975, "all", "midpoint")
    u = ops.clamp_cols(u, lo, hi).view(-1)
    mn, mx = ops.col_minmax(u.view(1, -1, 1))
    u = u - mn.view(())
    return u / (mx - mn).view(())


def _cols(signal):
    s = signal if signal.dim() >= 2 else signal.unsqueeze(1)
    return s.contiguous().view(1, s.shape[0], -1)


def clamp_peaks_percentile(signal, percent):
    """features/processing.py:101-121: clamp each column above the `percent` quantile of its strict interior peaks."""
    s = _cols(signal)
    thr, cnt = ops.col_quantile(s, percent / 100, "peaks", "linear", want_count=True)
    if int(cnt.min()) == 0:  # torch.quantile raises on an empty peak set (:117)
        raise RuntimeError("quantile() input tensor must be non-empty")
    return ops.clamp_cols(s, None, thr).view(s.shape[1], -1)


def clamp_upper_percentile(signal, percentile):
    """features/processing.py:124-125."""
    s = _cols(signal)
    return ops.clamp_cols(s, None, ops.col_quantile(s, percentile / 100)).view(signal.shape)


def clamp_lower_percentile(signal, percentile):
    """features/processing.py:128-129."""
    s = _cols(signal)
    return ops.clamp_cols(s, ops.col_quantile(s, percentile / 100), None).view(signal.shape)


def emphasize(envs, strength, percentile):
    """features/processing.py:132-138."""
    return ops.emphasize(_cols(envs), strength, percentile).view(envs.shape)


def _biquad(audio, kind, sr, fc):
    squeeze = audio.dim() == 1
    y = ops.biquad(audio, plan.biquad_coeffs(kind, sr, fc), clamp=True)
    return y[0] if squeeze else y


def low_pass(audio, sr, fmax=200):
    """features/processing.py:141-142 -> torchaudio lowpass_biquad (Q 0.707, output clamped to [-1, 1])."""
    return _biquad(audio, "low", sr, fmax)


def high_pass(audio, sr, fmin=4000):
    """features/processing.py:149-150."""
    return _biquad(audio, "high", sr, fmin)


def mid_pass(audio, sr, fmin=200, fmax=4000):
    """features/processing.py:145-146: literally low_pass(high_pass(audio, fmax), fmin) — replicated, not "fixed"."""
    return low_pass(high_pass(audio, sr, fmax), sr, fmin)
