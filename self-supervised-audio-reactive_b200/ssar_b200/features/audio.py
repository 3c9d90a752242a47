"""Mirror of ssar/features/audio.py — the public audio features (waveform [N] -> [T, C])."""
from __future__ import annotations

import torch

from .. import ops, plan
from .processing import emphasize, gaussian_filter, normalize
from .rosa.beat import onset_strength, plp  # noqa: F401  (re-exported like the reference)
from .rosa.spectral import chroma_cens


def hpss_waveforms(audio, margin=8.0, want_harm=True, want_perc=True, n_fft=2048, hop_length=1024):
    """stft -> hpss masks -> istft for both components with ONE forward STFT and ONE median pass.

    The reference runs the whole chain twice (harmonic(audio), percussive(audio)); the two results only differ in
    which mask is applied, so sharing the STFT and the medians is exact."""
    squeeze = audio.dim() == 1
    n = audio.shape[-1]
    D = ops.frames_fft(audio, n_fft, hop_length, 1 + n // hop_length, torch.hann_window, want_complex=True)["complex"]
    h, p = ops.hpss(D, float(margin), want_harm, want_perc)
    yh = ops.istft(h, n_fft, hop_length, torch.hann_window, n) if want_harm else None
    yp = ops.istft(p, n_fft, hop_length, torch.hann_window, n) if want_perc else None
    if squeeze:
        yh = yh[0] if yh is not None else None
        yp = yp[0] if yp is not None else None
    return yh, yp


def harmonic(audio, margin=8.0):
    """features/audio.py:13-17."""
    return hpss_waveforms(audio, margin, True, False)[0]


def percussive(audio, margin=8.0):
    """features/audio.py:20-24."""
    return hpss_waveforms(audio, margin, False, True)[1]


def rms(y, sr, frame_length=2048, hop_length=1024, center=True, pad_mode="reflect"):
    """features/audio.py:31-37 -> [T,1]: sqrt(mean(x^2)) over reflect-padded frames, last frame dropped."""
    if not center or pad_mode != "reflect":
        raise NotImplementedError("rms: only center=True, pad_mode='reflect' is built")
    squeeze = y.dim() == 1
    out = ops.frames_rms(y, frame_length, hop_length, y.shape[-1] // hop_length)
    out = out.unsqueeze(-1)
    return out[0] if squeeze else out


def mfcc(y, sr, n_mfcc=20, norm=False, **kwargs):
    """features/audio.py:60-65 -> [T, n_mfcc]."""
    squeeze = y.dim() == 1
    n_fft = kwargs.pop("n_fft", 2048)
    hop = kwargs.pop("hop_length", 1024)
    fmax = kwargs.pop("fmax", None)
    if kwargs:
        raise NotImplementedError(f"mfcc: unsupported kwargs {sorted(kwargs)}")
    bank = plan.mel_bank(float(sr), int(n_fft), 128, 0.0, None if fmax is None else float(fmax), str(y.device))
    out = ops.frames_fft(y, n_fft, hop, y.shape[-1] // hop, torch.hann_window, banks=[(bank, True, True)])
    M = ops.mfcc_from_mel(out["bank0"], out["bank0_max"], plan.dct_matrix(n_mfcc, 128, str(y.device)))
    if norm is True:
        M = M / M.flatten(1).norm(p=2, dim=1)[:, None, None]
    return M[0] if squeeze else M


def spectral_flatness(y, sr, n_fft=2048, hop_length=1024, window=torch.hann_window, center=True, pad_mode="reflect",
                      amin=1e-10, power=2.0):
    """features/audio.py:126-133 -> [T,1]."""
    if amin != 1e-10 or power != 2.0 or not center or pad_mode != "reflect":
        raise NotImplementedError("spectral_flatness: only the reference defaults are built")
    squeeze = y.dim() == 1
    out = ops.frames_fft(y, n_fft, hop_length, y.shape[-1] // hop_length, window, want_flatness=True)["flatness"]
    out = out.unsqueeze(-1)
    return out[0] if squeeze else out


def spectral_contrast(y, sr, n_fft=2048, hop_length=1024, window=torch.hann_window, center=True, pad_mode="reflect",
                      fmin=200.0, n_bands=6, quantile=0.02, linear=False):
    """features/audio.py:72-123 -> [T, n_bands + 1]."""
    if not center or pad_mode != "reflect":
        raise NotImplementedError("spectral_contrast: only center=True, pad_mode='reflect' is built")
    squeeze = y.dim() == 1
    mag = ops.frames_fft(y, n_fft, hop_length, y.shape[-1] // hop_length, window, spec_power=1)["spec"]
    bands = plan.contrast_bands(float(sr), int(n_fft), float(fmin), int(n_bands), float(quantile), str(y.device))
    out = ops.spectral_contrast(mag, bands, linear)
    return out[0] if squeeze else out


def onsets(audio, sr):
    """features/audio.py:27-28 -> [T, 1]."""
    return normalize(onset_strength(percussive(audio), sr).unsqueeze(-1))


def drop_strength(audio, sr):
    """features/audio.py:40-41 -> [T, 1]."""
    return emphasize(gaussian_filter(rms(audio, sr), 10), strength=10, percentile=50).unsqueeze(1)


def chromagram(audio, sr):
    """features/audio.py:44-45 -> [T, 12] (note: HPSS is applied to whatever is passed in, as in the reference)."""
    return chroma_cens(harmonic(audio), sr).transpose(-1, -2)


def tonnetz(y, sr, chroma_fn=None):
    """features/audio.py:48-57 -> [T, 6].  chroma_fn returns [12, T] like the reference's default lambda."""
    chroma_tm = chromagram(y, sr) if chroma_fn is None else chroma_fn(y, sr).transpose(-1, -2)
    return ops.tonnetz(chroma_tm)


def pulse(audio, sr):
    """features/audio.py:68-69 -> [T, 1]."""
    return plp(percussive(audio), sr).unsqueeze(-1)
