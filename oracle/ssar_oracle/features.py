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


# ------------------------------------------------------------------------------------------------
# ssar/random/mir.py: the `Patch` inputs (feature part of retrieve_music_information)
# ------------------------------------------------------------------------------------------------
def salience_weighted(envelope, short_sigma=5, long_sigma=80):
    """random/mir.py:14-22."""
    if envelope.dim() > 1:
        envelope = envelope.squeeze(1)
    short = post.gaussian_filter(envelope, short_sigma, mode="reflect")
    long = post.gaussian_filter(envelope, long_sigma, mode="reflect")
    weighted = (short / long) ** 2 * envelope
    if weighted.dim() < 2:
        weighted = weighted.unsqueeze(1)
    return weighted


AFEATFNS = {  # random/mir.py:8, in its order; every function is called on the audio itself (nothing shared)
    "chromagram": lambda a, sr: ocqt.chromagram(a, sr),
    "tonnetz": lambda a, sr: ocqt.tonnetz(a, sr),
    "mfcc": lambda a, sr: dsp.mfcc(a, sr),
    "spectral_contrast": lambda a, sr: dsp.spectral_contrast(a, sr),
    "spectral_flatness": lambda a, sr: dsp.spectral_flatness(a, sr).reshape(-1, 1),
    "rms": lambda a, sr: dsp.rms(a, sr).reshape(-1, 1),
    "drop_strength": lambda a, sr: drop_strength(a, sr),
    "onsets": lambda a, sr: onsets(a, sr),
}


def retrieve_music_features(audio, sr):
    """random/mir.py:26 and :44 (tempo, beats and segmentations, :28-42, are inputs of `Patch` here)."""
    raw = {name: fn(audio, sr) for name, fn in AFEATFNS.items()}
    return {k: post.normalize(salience_weighted(post.gaussian_filter(af, 2))) for k, af in raw.items()}


def audio2features_generate(audio, sr, clamp=True, smooth=True):
    """ssar/features/generate.py:106-136: the generators' variant of audio2features — audio already at sr = 1024 * fps,
    the multi-column features and the flatness pre-smoothed with sigma 10, pulse from the audio itself, clamp to the
    per-column [q10, q97.5] (torch.quantile, linear), then sigma 2."""
    harm, perc = dsp.harmonic(audio), dsp.percussive(audio)
    g10 = lambda f: post.gaussian_filter(f, 10)
    multi = [g10(dsp.mfcc(audio, sr)), g10(ocqt.chromagram(harm, sr)), g10(ocqt.tonnetz(harm, sr)),
             g10(dsp.spectral_contrast(audio, sr))]
    single = [
        g10(dsp.spectral_flatness(audio, sr)),
        dsp.onset_strength(perc, sr),
        dsp.onset_strength(dsp.low_pass(perc, sr), sr),
        dsp.onset_strength(dsp.mid_pass(perc, sr), sr),
        dsp.onset_strength(dsp.high_pass(perc, sr), sr),
        dsp.plp(audio, sr),
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
        feats = feats.clamp(torch.quantile(feats, q=0.1, dim=0), torch.quantile(feats, q=0.975, dim=0))
    if smooth:
        feats = post.gaussian_filter(feats, 2)
    return feats
