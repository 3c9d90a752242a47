"""Plan-time constants (twiddles, windows, banded filterbanks) cached per device.

These are the small immutable tables the kernels read; they are built once on the host (fp64 where the
reference's own construction is fp64-free we follow its fp32 op order so the weights agree to the last bits),
uploaded, and reused for every clip.  Nothing here touches waveform data.
"""
from __future__ import annotations

import functools
import math
from dataclasses import dataclass

import numpy as np
import torch


def _dev(device) -> torch.device:
    device = torch.device(device)
    if device.type != "cuda":
        from ._cabi import SsarError

        raise SsarError(f"ssar_b200 runs on CUDA devices only (got {device}); there is no CPU fallback")
    return device


@functools.lru_cache(maxsize=None)
def twiddle(n_fft: int, device) -> torch.Tensor:
    """[n_fft, 2] fp32: exp(-2*pi*i*k/n_fft), evaluated in fp64."""
    k = np.arange(n_fft, dtype=np.float64)
    ang = -2.0 * np.pi * k / n_fft
    tw = np.stack([np.cos(ang), np.sin(ang)], axis=1).astype(np.float32)
    return torch.from_numpy(tw).to(_dev(device))


@functools.lru_cache(maxsize=None)
def hann(n: int, device) -> torch.Tensor:
    """Periodic Hann, exactly `torch.hann_window(n)` as the reference builds it (rosa/spectral.py:17)."""
    return torch.hann_window(n, dtype=torch.float32).to(_dev(device))


def window_tensor(window, n: int, device):
    """`window` follows the reference API: a callable like torch.hann_window, a tensor, or None (rectangular)."""
    if window is None:
        return None
    if window is torch.hann_window:
        return hann(n, str(device))
    if callable(window):
        return window(n).to(dtype=torch.float32, device=_dev(device)).contiguous()
    return torch.as_tensor(window, dtype=torch.float32).to(_dev(device)).contiguous()


@dataclass(frozen=True)
class Banded:
    """Row-banded matrix: row r covers columns [start[r], start[r]+count[r]) with packed weights."""

    n_rows: int
    n_cols: int
    start: torch.Tensor   # int32 [n_rows]
    count: torch.Tensor   # int32 [n_rows]
    offset: torch.Tensor  # int32 [n_rows]
    weights: torch.Tensor  # fp32 [nnz] (or [nnz, 2] for complex)
    dense: torch.Tensor   # the dense matrix on the device (row-major), for GEMM-style consumers / inspection


def banded_from_dense(dense_cpu: torch.Tensor, device) -> Banded:
    device = _dev(device)
    is_c = dense_cpu.is_complex()
    mag = dense_cpu.abs()
    n_rows, n_cols = dense_cpu.shape
    starts, counts, offsets, chunks = [], [], [], []
    off = 0
    for r in range(n_rows):
        nz = torch.nonzero(mag[r] > 0).flatten()
        if len(nz) == 0:
            s, c = 0, 0
        else:
            s, c = int(nz[0]), int(nz[-1]) - int(nz[0]) + 1
        starts.append(s)
        counts.append(c)
        offsets.append(off)
        row = dense_cpu[r, s : s + c]
        chunks.append(torch.view_as_real(row.to(torch.complex64)) if is_c else row.to(torch.float32))
        off += c
    w = torch.cat(chunks) if off > 0 else torch.zeros((0, 2) if is_c else (0,), dtype=torch.float32)
    i32 = lambda v: torch.tensor(v, dtype=torch.int32, device=device)
    return Banded(
        n_rows, n_cols, i32(starts), i32(counts), i32(offsets), w.contiguous().to(device),
        (torch.view_as_real(dense_cpu.to(torch.complex64)) if is_c else dense_cpu.to(torch.float32)).contiguous().to(device),
    )


# ------------------------------------------------------------------------------------------------
# Slaney mel filterbank — same fp32 op order as rosa/spectral.py:73-110 + rosa/convert.py:15-66
# ------------------------------------------------------------------------------------------------
_F_SP = 200.0 / 3
_MIN_LOG_HZ = 1000.0
_MIN_LOG_MEL = _MIN_LOG_HZ / _F_SP
_LOGSTEP = math.log(6.4) / 27.0


def hz_to_mel_f32(f: torch.Tensor) -> torch.Tensor:
    f = f.to(torch.float32)
    mels = f / _F_SP
    log_t = f >= _MIN_LOG_HZ
    return torch.where(log_t, _MIN_LOG_MEL + torch.log(f / _MIN_LOG_HZ) / _LOGSTEP, mels)


def mel_to_hz_f32(m: torch.Tensor) -> torch.Tensor:
    m = m.to(torch.float32)
    freqs = _F_SP * m
    log_t = m >= _MIN_LOG_MEL
    return torch.where(log_t, _MIN_LOG_HZ * torch.exp(_LOGSTEP * (m - _MIN_LOG_MEL)), freqs)


def mel_dense_cpu(sr, n_fft: int, n_mels: int = 128, fmin: float = 0.0, fmax=None) -> torch.Tensor:
    if fmax is None:
        fmax = float(sr) / 2
    n_bins = 1 + n_fft // 2
    fftfreqs = torch.linspace(0, float(sr) / 2, n_bins)
    lo = hz_to_mel_f32(torch.tensor(float(fmin)))
    hi = hz_to_mel_f32(torch.tensor(float(fmax)))
    mel_f = mel_to_hz_f32(torch.linspace(float(lo), float(hi), n_mels + 2))
    fdiff = torch.diff(mel_f)
    ramps = mel_f.reshape(-1, 1) - fftfreqs
    lower = -ramps[:n_mels] / fdiff[:n_mels, None]
    upper = ramps[2 : n_mels + 2] / fdiff[1 : n_mels + 1, None]
    weights = torch.clamp_min(torch.minimum(lower, upper), 0.0)
    enorm = 2.0 / (mel_f[2 : n_mels + 2] - mel_f[:n_mels])
    return weights * enorm[:, None]


@functools.lru_cache(maxsize=None)
def mel_bank(sr: float, n_fft: int, n_mels: int, fmin: float, fmax, device) -> Banded:
    return banded_from_dense(mel_dense_cpu(sr, n_fft, n_mels, fmin, fmax), device)


@functools.lru_cache(maxsize=None)
def dct_matrix(n_out: int, n_in: int, device) -> torch.Tensor:
    """Orthonormal DCT-II rows: what rosa/spectral.py:35-56 `dct(x, norm="ortho")[:n_out]` evaluates via an FFT."""
    n = np.arange(n_in, dtype=np.float64)
    k = np.arange(n_out, dtype=np.float64)[:, None]
    d = np.cos(np.pi * k * (2 * n + 1) / (2 * n_in))
    d[0] *= 1.0 / np.sqrt(n_in)
    d[1:] *= np.sqrt(2.0 / n_in)
    return torch.from_numpy(d.astype(np.float32)).to(_dev(device))


# ------------------------------------------------------------------------------------------------
# torchaudio-style windowed-sinc resampling taps (algorithm of torchaudio.functional._get_sinc_resample_kernel,
# third-party / unpinned by the reference; evaluated in fp32 exactly like `resample(waveform_fp32, ...)` does)
# ------------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class ResamplePlan:
    orig: int
    new: int
    width: int
    max_taps: int
    taps: torch.Tensor   # fp32 [max_taps, new] (tap-major)
    first: torch.Tensor  # int32 [new]


def sinc_kernel_cpu(orig: int, new: int, method: str = "sinc_interp_hann", lowpass_filter_width: int = 6,
                    rolloff: float = 0.99, beta: float = 14.769656459379492):
    """Dense [new, 2*width+orig] fp32 kernel + width, same fp32 op order as torchaudio."""
    base_freq = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base_freq)
    idx = torch.arange(-width, width + orig, dtype=torch.float32)[None, :] / orig
    t = torch.arange(0, -new, -1, dtype=torch.float32)[:, None] / new + idx
    t *= base_freq
    t = t.clamp_(-lowpass_filter_width, lowpass_filter_width)
    if method in ("sinc_interp_hann", "sinc_interpolation"):
        window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    elif method in ("sinc_interp_kaiser", "kaiser_window"):
        b = torch.tensor(float(beta))
        window = torch.i0(b * torch.sqrt(1 - (t / lowpass_filter_width) ** 2)) / torch.i0(b)
    else:
        raise ValueError(method)
    t *= math.pi
    scale = base_freq / orig
    k = torch.where(t == 0, torch.tensor(1.0), t.sin() / t)
    k *= window * scale
    inside = (t.abs() < lowpass_filter_width * math.pi)  # un-clamped region; the clamped tail is ~1e-20
    return k, width, inside


@functools.lru_cache(maxsize=None)
def resample_plan(orig_freq: int, new_freq: int, method: str, device) -> ResamplePlan:
    if int(orig_freq) != orig_freq or int(new_freq) != new_freq:
        raise ValueError("Frequencies must be whole numbers (torchaudio raises on fractional rates)")
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    k, width, inside = sinc_kernel_cpu(orig, new, method)
    kw = k.shape[1]
    if kw <= 64:  # small kernels (the /2 kaiser decimator: 28 taps) keep every tap, clamped tail included
        inside = torch.ones_like(inside)
    any_in = inside.any(dim=1)
    first = torch.where(any_in, inside.float().argmax(dim=1), torch.zeros(new, dtype=torch.long))
    last = torch.where(any_in, kw - 1 - inside.flip(1).float().argmax(dim=1), torch.zeros(new, dtype=torch.long))
    max_taps = int((last - first + 1).max())
    cols = first[:, None] + torch.arange(max_taps)[None, :]
    valid = cols <= last[:, None]
    taps = torch.where(valid, k.gather(1, cols.clamp(max=kw - 1)), torch.zeros(()))
    dev = _dev(device)
    return ResamplePlan(orig, new, width, max_taps, taps.T.contiguous().to(dev), first.to(torch.int32).to(dev))


def biquad_coeffs(kind: str, sample_rate, cutoff_freq, Q: float = 0.707):
    """torchaudio lowpass_biquad / highpass_biquad coefficient design in fp32 tensors, normalised by a0."""
    cutoff = torch.as_tensor(cutoff_freq, dtype=torch.float32)
    Qt = torch.as_tensor(Q, dtype=torch.float32)
    w0 = 2 * math.pi * cutoff / sample_rate
    alpha = torch.sin(w0) / 2 / Qt
    if kind == "low":
        b0 = (1 - torch.cos(w0)) / 2
        b1 = 1 - torch.cos(w0)
    elif kind == "high":
        b0 = (1 + torch.cos(w0)) / 2
        b1 = -1 - torch.cos(w0)
    else:
        raise ValueError(kind)
    b2 = b0
    a0 = 1 + alpha
    a1 = -2 * torch.cos(w0)
    a2 = 1 - alpha
    return tuple(float(v) for v in (b0 / a0, b1 / a0, b2 / a0, a1 / a0, a2 / a0))
