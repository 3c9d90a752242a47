"""Mirror of ssar/features/processing.py (filters, normalisation, percentile clamps)."""
from __future__ import annotations

import torch

from .. import ops, plan


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
