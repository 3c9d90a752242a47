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
    n_frames = 1 + (n + 2 * (n_fft // 2) - n_fft) // hop_length
    if n_fft in (1024, 2048):
        out = ops.frames_fft(y, n_fft, hop_length, n_frames, window, want_complex=True)["complex"]
    else:  # any other length (e.g. the tempogram window of a clip shorter than 1024 frames): direct DFT
        out = ops.frames_dft(y, n_fft, hop_length, n_frames, window)
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


def _fm(spec):
    """reference-layout complex [F,T] / [B,F,T] -> frame-major real view [B,T,F,2] (copies only if needed)"""
    squeeze = spec.dim() == 2
    if squeeze:
        spec = spec[None]
    fm = torch.view_as_real(spec.transpose(1, 2).contiguous())
    return fm, squeeze


def istft(spec, n_fft=2048, hop_length=1024, center=True, window=torch.hann_window, length=None):
    """rosa/spectral.py:24-32 (torch.istft, center=True).  spec: complex [F, T] or [B, F, T]."""
    if not center:
        raise NotImplementedError("istft: only center=True is built")
    fm, squeeze = _fm(spec)
    out = ops.istft(fm, n_fft, hop_length, window, length)
    return out[0] if squeeze else out


def hpss(S, ks=31, power=2.0, margin=1.0):
    """rosa/spectral.py:145-161 on a complex STFT [F, T] -> (harmonic, percussive) complex STFTs."""
    if ks != 31 or power != 2.0:
        raise NotImplementedError("hpss: only ks=31, power=2 (the reference's only use) is built")
    if not torch.is_complex(S):
        raise NotImplementedError("hpss: complex STFT input expected")
    fm, squeeze = _fm(S)
    h, p = ops.hpss(fm, float(margin))
    h = _ft(torch.view_as_complex(h), squeeze)
    p = _ft(torch.view_as_complex(p), squeeze)
    return h, p


def chroma_cqt(y, sr, hop_length=1024, fmin=None, threshold=0.0, tuning=None, n_chroma=12, n_octaves=7, window=None,
               bins_per_octave=36, norm=True):
    """rosa/spectral.py:286-325 -> [n_chroma, T] (or [B, n_chroma, T])."""
    from .constantq import _tuning_index

    if fmin is not None or window is not None or n_octaves != 7 or bins_per_octave != 36:
        raise NotImplementedError("chroma_cqt: only fmin=C1, 7 octaves x 36 bins, window=None is built")
    if threshold is not None and threshold > 0.0:
        raise NotImplementedError("chroma_cqt: threshold > 0 is not used by the reference path")
    squeeze = y.dim() == 1
    B = 1 if squeeze else y.shape[0]
    C = ops.cqt(y, sr, hop_length, _tuning_index(tuning, B, y.device))
    chroma = ops.chroma_from_cqt(C, bins_per_octave, n_chroma, quantize=False)
    if norm:
        chroma = chroma / chroma.flatten(1).max(dim=1).values[:, None, None]
    return _ft(chroma, squeeze)


def chroma_cens(y, sr, hop_length=1024, fmin=None, tuning=None, n_chroma=12, n_octaves=7, bins_per_octave=36,
                window=None, win_len_smooth=41, smoothing_window=torch.hann_window):
    """rosa/spectral.py:239-280 -> [12, T] (or [B, 12, T]): CQT chroma -> L1 -> soft quantiser -> Hann smoothing -> L2."""
    from .constantq import _tuning_index

    if fmin is not None or window is not None or n_octaves != 7 or bins_per_octave != 36:
        raise NotImplementedError("chroma_cens: only fmin=C1, 7 octaves x 36 bins, window=None is built")
    if smoothing_window is not torch.hann_window:
        raise NotImplementedError("chroma_cens: only the Hann smoothing window is built")
    squeeze = y.dim() == 1
    B = 1 if squeeze else y.shape[0]
    C = ops.cqt(y, sr, hop_length, _tuning_index(tuning, B, y.device))
    q = ops.chroma_from_cqt(C, bins_per_octave, n_chroma, quantize=True)
    win = plan.cens_window(int(win_len_smooth), str(y.device)) if win_len_smooth else None
    return _ft(ops.cens_smooth(q, win, l2=True), squeeze)
