"""Oracle: audio2features (ssar/supervised/data.py:60-107) composed from the oracle's feature functions, call for
call in the reference's order (nothing shared or deduplicated).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import torch

from . import cqt as ocqt
from . import dsp, post


def drop_strength(x, sr):
    """features/audio.py:40-41."""
    return post.emphasize(post.gaussian_filter(dsp.rms(x, sr), 10), 10, 50).unsqueeze(1)


def onsets(x, sr):
    """features/audio.py:27-28."""
    return post.normalize(dsp.onset_strength(dsp.percussive(x), sr).unsqueeze(-1))


def audio2features(audio, sr, fps, clamp=True, smooth=True, emphasis=False):
    if audio.dim() == 2:
        audio = audio.mean(0)
    audio, sr = dsp.resample(audio, sr, fps * 1024), fps * 1024
    harm, perc = dsp.harmonic(audio), dsp.percussive(audio)
    multi = [dsp.mfcc(audio, sr), ocqt.chromagram(harm, sr), ocqt.tonnetz(harm, sr), dsp.spectral_contrast(audio, sr)]
    single = [
        dsp.spectral_flatness(audio, sr),
        dsp.onset_strength(perc, sr),
        dsp.onset_strength(dsp.low_pass(perc, sr), sr),
        dsp.onset_strength(dsp.mid_pass(perc, sr), sr),
        dsp.onset_strength(dsp.high_pass(perc, sr), sr),
        dsp.plp(perc, sr),
        dsp.rms(harm, sr),
        dsp.rms(dsp.low_pass(harm, sr), sr),
        dsp.rms(dsp.mid_pass(harm, sr), sr),
        dsp.rms(dsp.high_pass(harm, sr), sr),
        drop_strength(audio, sr),
        drop_strength(dsp.low_pass(audio, sr), sr),
        drop_strength(dsp.mid_pass(audio, sr), sr),
        drop_strength(dsp.high_pass(audio, sr), sr),
    ]
    feats = torch.cat(multi + [s.reshape(-1, 1) for s in single], dim=1)
    if clamp:
        feats = post.clamp_peaks_percentile(feats, 100 - 2.5)
        feats = post.clamp_lower_percentile(feats, 4 * 2.5)
    if smooth:
        feats = post.gaussian_filter(feats, 0.1 * fps)
    if emphasis:
        feats = post.emphasize(feats, 2, 75)
    return feats
