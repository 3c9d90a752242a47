"""Mirror of ssar/features/rosa/spectral.py — same names, kwargs and output shapes, CUDA kernels underneath.

Shapes: the reference returns frequency-major `[F, T]`; the kernels write frame-major `[T, F]` (every consumer
is per-frame), so the functions here return a transposed VIEW with the reference's shape.  A leading clip
dimension is accepted everywhere (`[B, N]` in -> `[B, F, T]` out) — one launch for the whole batch.
"""
from __future__ import annotations

import torch

from ... import ops, plan
from ..._cabi import SsarError


def _check_stft_args(center, pad_mode):
    if not center or pad_mode != "reflect":
        raise NotImplementedError("only center=True, pad_mode='reflect' (the reference's only use) is built")


def _ft(x, squeeze):
    """frame-major [B,T,F(,2)] -> reference layout [F,T] / [B,F,T] view"""
    x = x.transpose(1, 2)
    return x[0] if squeeze else x


def stft(y, n_fft=2048, hop_length=1024, center=True, window=torch.hann_window, pad_mode="reflect", return_complex=True):
    """rosa/spectral.py:10-21 -> complex64 [1+n_fft/2, 1+N//hop]."""
    _check_stft_args(center, pad_mode)
    squeeze = y.dim() == 1
    n = y.shape[-1]
    out = ops.frames_fft(y, n_fft, hop_length, 1 + n // hop_length, window, want_complex=True)["complex"]
    c = torch.view_as_complex(out)
    c = _ft(c, squeeze)
    return c if return_complex else torch.view_as_real(c)


def spectrogram(y, n_fft=2048, hop_length=1024, power=1, window=torch.hann_window, center=True, pad_mode="reflect"):
    """rosa/spectral.py:59-62: |stft|[:, :-1] ** power -> [1+n_fft/2, N//hop]."""
    _check_stft_args(center, pad_mode)
    squeeze = y.dim() == 1
    if hop_length is None:
        hop_length = n_fft // 4  # torch.stft's default, reached through rosa/pitch.py:40
    n = y.shape[-1]
    if power in (1, 2, 1.0, 2.0):
        s = ops.frames_fft(y, n_fft, hop_length, n // hop_length, window, spec_power=int(power))["spec"]
    else:
        s = ops.frames_fft(y, n_fft, hop_length, n // hop_length, window, spec_power=1)["spec"] ** power
    return _ft(s, squeeze)


def mel(sr, n_fft, n_mels=128, fmin=0.0, fmax=None, htk=False, dtype=torch.float, device="cuda"):
    """rosa/spectral.py:81-110: dense Slaney mel basis [n_mels, 1+n_fft/2] (built once per plan, cached)."""
    if htk:
        raise NotImplementedError("htk mel scale is not used by the reference path")
    return plan.mel_bank(float(sr), int(n_fft), int(n_mels), float(fmin), None if fmax is None else float(fmax),
                         str(torch.device(device))).dense.to(dtype)


def melspectrogram(y, sr, n_fft=2048, hop_length=1024, window=torch.hann_window, center=True, pad_mode="reflect",
                   power=2.0, fmax=None):
    """rosa/spectral.py:65-70: mel_basis @ |stft|^power, fused into the FFT kernel's epilogue -> [128, N//hop]."""
    _check_stft_args(center, pad_mode)
    if power not in (1, 2, 1.0, 2.0):
        raise NotImplementedError("melspectrogram: power must be 1 or 2")
    squeeze = y.dim() == 1
    n = y.shape[-1]
    bank = plan.mel_bank(float(sr), int(n_fft), 128, 0.0, None if fmax is None else float(fmax), str(y.device))
    out = ops.frames_fft(y, n_fft, hop_length, n // hop_length, window, banks=[(bank, int(power) == 2, False)])
    return _ft(out["bank0"], squeeze)


def mel_frequencies(n_mels=128, fmin=0.0, fmax=11025.0, htk=False, device="cpu"):
    lo = plan.hz_to_mel_f32(torch.tensor(float(fmin)))
    hi = plan.hz_to_mel_f32(torch.tensor(float(fmax)))
    return plan.mel_to_hz_f32(torch.linspace(float(lo), float(hi), n_mels)).to(device)
