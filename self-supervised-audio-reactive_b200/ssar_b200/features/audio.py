"""Mirror of ssar/features/audio.py — the public audio features (waveform [N] -> [T, C])."""
from __future__ import annotations

import torch

from .. import ops, plan
from .rosa.beat import onset_strength  # noqa: F401  (re-exported like the reference)


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
