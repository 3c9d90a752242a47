"""Mirror of the feature part of ssar/random/mir.py (:8-22, :26, :44).  Tempo / beat tracking and the Laplacian
segmentations (:28-42; librosa + scipy + torch_geometric on CPU) are OUT OF SCOPE: they are inputs of `Patch`."""
from __future__ import annotations

from ..features.audio import (chromagram, drop_strength, mfcc, onsets, rms, spectral_contrast, spectral_flatness,
                              tonnetz)
from ..features.processing import gaussian_filter, normalize

AFEATFNS = [chromagram, tonnetz, mfcc, spectral_contrast, spectral_flatness, rms, drop_strength, onsets]
UNITFEATS = ["rms", "drop_strength", "onsets", "spectral_flatness"]
ALLFEATS = ["chromagram", "tonnetz", "mfcc", "spectral_contrast"] + UNITFEATS


def salience_weighted(envelope, short_sigma=5, long_sigma=80):
    """random/mir.py:14-22."""
    if envelope.dim() > 1:
        envelope = envelope.squeeze(1)
    short = gaussian_filter(envelope, short_sigma, mode="reflect", causal=0)
    long = gaussian_filter(envelope, long_sigma, mode="reflect", causal=0)
    weighted = (short / long) ** 2 * envelope
    if weighted.dim() < 2:
        weighted = weighted.unsqueeze(1)
    return weighted


def retrieve_music_features(audio, sr):
    """random/mir.py:26 + :44: the 8 feature functions, each smoothed (sigma 2), salience-weighted and normalised."""
    features = {afn.__name__: afn(audio, sr) for afn in AFEATFNS}
    return {k: normalize(salience_weighted(gaussian_filter(af, sigma=2))) for k, af in features.items()}
