"""Mirror of ssar/features/rosa/constantq.py — the constant-Q transform used by chroma_cqt.

Only the configuration the reference reaches is built (rosa/spectral.py:302-312): fmin = C1, 7 octaves x 36 bins,
filter_scale 1, sparsity 0.01, gamma 0; `tuning` may be None (estimated, the reference default through chroma_cens),
or a value on the estimator's 0.01 grid."""
from __future__ import annotations

import torch

from ... import ops, plan


def _tuning_index(tuning, n_clips, device):
    if tuning is None:
        return None
    t = float(tuning)
    idx = round((t + 0.5) * 100)
    if idx < 0 or idx > 100 or abs(float(plan.cqt_tuning_values()[idx]) - t) > 1e-6:
        raise NotImplementedError("cqt: tuning must lie on the estimator grid linspace(-0.5, 0.5, 101)")
    return torch.full((n_clips,), idx, dtype=torch.int32, device=device)


def cqt(y, sr, hop_length=1024, fmin=None, n_bins=84, bins_per_octave=12, tuning=0.0, filter_scale=1, sparsity=0.01):
    """rosa/constantq.py:13-26 -> complex64 [n_bins, T] (or [B, n_bins, T])."""
    return vqt(y, sr, hop_length, fmin, n_bins, 0, bins_per_octave, tuning, filter_scale, sparsity)


def vqt(y, sr, hop_length=1024, fmin=None, n_bins=84, gamma=None, bins_per_octave=12, tuning=0.0, filter_scale=1,
        sparsity=0.01):
    """rosa/constantq.py:29-115."""
    if (fmin is not None or n_bins != plan.CQT_N_OCTAVES * plan.CQT_BINS_PER_OCTAVE
            or bins_per_octave != plan.CQT_BINS_PER_OCTAVE or gamma != 0 or filter_scale != 1 or sparsity != 0.01):
        raise NotImplementedError("cqt/vqt: only fmin=C1, n_bins=252, bins_per_octave=36, gamma=0, filter_scale=1, "
                                  "sparsity=0.01 (the reference's chroma configuration) is built")
    squeeze = y.dim() == 1
    B = 1 if squeeze else y.shape[0]
    out = ops.cqt(y, sr, hop_length, _tuning_index(tuning, B, y.device))
    c = torch.view_as_complex(out).transpose(1, 2)
    return c[0] if squeeze else c
