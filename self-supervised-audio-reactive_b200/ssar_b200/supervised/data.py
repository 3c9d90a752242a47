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
