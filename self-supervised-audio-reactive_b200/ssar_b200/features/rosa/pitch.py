"""Mirror of ssar/features/rosa/pitch.py (tuning estimation that gates the CQT basis)."""
from __future__ import annotations

from ... import ops


def estimate_tuning(y, sr, n_fft=2048, resolution=0.01, bins_per_octave=12, **kwargs):
    """rosa/pitch.py:9-24 -> 0-d tensor (or [B]) in {-0.5, -0.49, ..., 0.5}."""
    if resolution != 0.01:
        raise NotImplementedError("estimate_tuning: only resolution=0.01 (the reference's only use) is built")
    squeeze = y.dim() == 1
    _, val = ops.estimate_tuning(y, sr, n_fft=n_fft, bins_per_octave=bins_per_octave, **kwargs)
    return val[0] if squeeze else val


def estimate_tuning_index(y, sr, n_fft=2048, bins_per_octave=12):
    """Same estimate as the index into linspace(-0.5, 0.5, 101) (what the CQT basis tables are keyed by)."""
    return ops.estimate_tuning(y, sr, n_fft=n_fft, bins_per_octave=bins_per_octave)[0]
