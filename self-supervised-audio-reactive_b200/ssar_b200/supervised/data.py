"""Mirror of ssar/supervised/data.py:32-107 — FEATURE_NAMES and audio2features (waveform -> [T, 59])."""
from __future__ import annotations

import torch

from .. import pipeline

_FN = [
    *[f"mfcc_{i}" for i in range(20)],
    *[f"chroma_{i}" for i in range(12)],
    *[f"tonnetz_{i}" for i in range(6)],
    *[f"contrast_{i}" for i in range(7)],
    "flatness",
    "onsets",
    "onsets_low",
    "onsets_mid",
    "onsets_high",
    "pulse",
    "harmonic_rms",
    "harmonic_rms_low",
    "harmonic_rms_mid",
    "harmonic_rms_high",
    "long_rms",
    "long_rms_low",
    "long_rms_mid",
    "long_rms_high",
]
VELOCITY = False
FEATURE_NAMES = _FN


@torch.inference_mode()
def audio2features(audio, sr, fps, clamp=True, smooth=True, emphasis=False):
    """supervised/data.py:60-107: audio [N] or [channels, N] at `sr` -> features [N' // 1024, 59] at `fps` frames/s."""
    if audio.dim() == 2:
        audio = audio.mean(0)
    return pipeline.audio2features_batch(audio[None], sr, fps, clamp, smooth, emphasis)[0]


def overlapping_slices(tensor, length):
    """supervised/data.py:255-258 (and AudioFeatures.__getitem__, :128-131): the overlapping training windows of one
    track's feature / latent / noise sequence, [T, ...] -> [n_windows, length, ...]."""
    from .. import ops

    return ops.overlapping_slices(tensor, int(length))


def window_mean_std(feature_sequences, length, group=None):
    """supervised/data.py:317-319: np.mean / np.std over ALL windows and frames of the sliced training set — a frame
    counts once per window that holds it.  feature_sequences: iterable of [T_i, 59] CUDA tensors (this rank's tracks);
    the (count, sum, sum of squares) accumulators are all-reduced across ranks (NCCL) when torch.distributed is up."""
    from .. import distributed, ops

    stats = None
    for f in feature_sequences:
        stats = ops.feature_stats(ops.overlapping_slices(f, int(length)), stats)
    return distributed.mean_std_from_stats(distributed.merge_stats(stats, group))
