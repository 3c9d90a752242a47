"""Thin Python wrappers over the C-ABI entry points (one function per `ssar_*` symbol).

Everything here takes and returns CUDA tensors; buffers come from torch's caching allocator and kernels run on
torch's current stream.  Layout convention: spectra are frame-major `[clip, frame, bin]`.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _cabi as cabi
from . import plan
from ._cabi import FramesArgs, SsarError, f32, f64, i32, i64, ptr

KEY_NEG_INF = 0x007FFFFF  # order-preserving key of -inf (see csrc/common.cuh f2key)
KEY_POS_INF = 0xFF800000


def _as_batch(x: torch.Tensor):
    if not x.is_cuda:
        raise SsarError("ssar_b200 needs CUDA tensors (no CPU fallback)")
    if x.dtype != torch.float32:
        x = x.float()
    squeeze = x.dim() == 1
    if squeeze:
        x = x[None]
    if x.dim() != 2:
        raise SsarError(f"expected waveform [N] or [clips, N], got {tuple(x.shape)}")
    if x.stride(1) != 1:
        x = x.contiguous()
    return x, squeeze


def fill_u32(t: torch.Tensor, value: int):
    cabi.call("ssar_fill_u32", ptr(t), i64(t.numel()), C.c_uint32(value))


def frames_fft(
    x: torch.Tensor,
    n_fft: int,
    hop: int,
    n_frames: int,
    window=torch.hann_window,
    *,
    want_complex: bool = False,
    spec_power: int | None = None,
    want_rms: bool = False,
    want_flatness: bool = False,
    banks=(),
    projection=None,
    projection_scale: float = 1.0,
):
    """Framed real FFT + fused epilogues over a batch of clips.

    banks: up to two `(plan.Banded, on_power: bool, track_max: bool)`.
    Returns a dict with the requested outputs (frame-major):
      complex [B,T,F,2], spec [B,T,F], rms [B,T], flatness [B,T], bank{i} [B,T,n_bands], bank{i}_max uint32 key [B],
      proj [B,T,n_rows,2]
    """
    x, _ = _as_batch(x)
    B, N = x.shape
    dev = x.device
    F = n_fft // 2 + 1
    a = FramesArgs()
    a.x = x.data_ptr()
    a.n_samples = N
    a.clip_stride = x.stride(0)
    a.n_clips = B
    a.n_fft = n_fft
    a.hop = hop
    a.n_frames = n_frames
    win = plan.window_tensor(window, n_fft, dev)
    a.window = win.data_ptr() if win is not None else None
    tw = plan.twiddle(n_fft, str(dev))
    a.twiddle = tw.data_ptr()
    out = {}
    keep = [x, win, tw]
    if want_complex:
        out["complex"] = torch.empty(B, n_frames, F, 2, dtype=torch.float32, device=dev)
        a.out_complex = out["complex"].data_ptr()
        a.ldc = F
    if spec_power is not None:
        out["spec"] = torch.empty(B, n_frames, F, dtype=torch.float32, device=dev)
        a.out_spec = out["spec"].data_ptr()
        a.lds = F
        a.spec_power = int(spec_power)
    if want_rms:
        out["rms"] = torch.empty(B, n_frames, dtype=torch.float32, device=dev)
        a.out_rms = out["rms"].data_ptr()
    if want_flatness:
        out["flatness"] = torch.empty(B, n_frames, dtype=torch.float32, device=dev)
        a.out_flatness = out["flatness"].data_ptr()
    if len(banks) > 2:
        raise SsarError("frames_fft: at most two filterbanks per launch")
    for i, (bank, on_power, track_max) in enumerate(banks):
        if bank.n_cols != F:
            raise SsarError(f"filterbank has {bank.n_cols} columns, spectrum has {F} bins")
        fb = a.fb[i]
        fb.n_bands = bank.n_rows
        fb.on_power = 1 if on_power else 0
        fb.start, fb.count, fb.offset = bank.start.data_ptr(), bank.count.data_ptr(), bank.offset.data_ptr()
        fb.weights = bank.weights.data_ptr()
        o = torch.empty(B, n_frames, bank.n_rows, dtype=torch.float32, device=dev)
        out[f"bank{i}"] = o
        fb.out = o.data_ptr()
        fb.ld = bank.n_rows
        if track_max:
            mk = torch.empty(B, dtype=torch.int32, device=dev)
            fill_u32(mk, KEY_NEG_INF)
            out[f"bank{i}_max"] = mk
            fb.max_key = mk.data_ptr()
        keep.append(bank)
    if projection is not None:
        if projection.n_cols != F:
            raise SsarError(f"projection has {projection.n_cols} columns, spectrum has {F} bins")
        pj = a.proj
        pj.n_rows = projection.n_rows
        pj.start, pj.count, pj.offset = (
            projection.start.data_ptr(), projection.count.data_ptr(), projection.offset.data_ptr())
        pj.weights = projection.weights.data_ptr()
        o = torch.empty(B, n_frames, projection.n_rows, 2, dtype=torch.float32, device=dev)
        out["proj"] = o
        pj.out = o.data_ptr()
        pj.ld = projection.n_rows
        pj.scale = float(projection_scale)
    cabi.call("ssar_frames_fft", C.byref(a))
    return out


def minmax_keys(x: torch.Tensor, n_clips: int = 1, want_min=True, want_max=True):
    """Per-clip min/max of a contiguous [n_clips, ...] tensor as order-preserving keys (int32 storage)."""
    x = x.contiguous()
    n_per = x.numel() // max(n_clips, 1)
    mn = mx = None
    if want_min:
        mn = torch.empty(n_clips, dtype=torch.int32, device=x.device)
        fill_u32(mn, KEY_POS_INF)
    if want_max:
        mx = torch.empty(n_clips, dtype=torch.int32, device=x.device)
        fill_u32(mx, KEY_NEG_INF)
    cabi.call("ssar_minmax_keys", ptr(x), i32(n_clips), i64(n_per), ptr(mn), ptr(mx))
    return mn, mx


def onset_flux(mel: torch.Tensor, max_key: torch.Tensor, pad_width: int, aggregate: str = "mean", top_db: float = 80.0):
    """mel [B,T,nb] -> onset envelope [B,T] (rosa/beat.py:10-23)."""
    B, T, nb = mel.shape
    out = torch.empty(B, T, dtype=torch.float32, device=mel.device)
    cabi.call("ssar_onset_flux", ptr(mel), ptr(max_key), i32(B), i32(T), i32(nb), i32(pad_width),
              i32({"mean": 0, "median": 1}[aggregate]), f32(top_db), ptr(out))
    return out


def mfcc_from_mel(mel: torch.Tensor, max_key: torch.Tensor, dct: torch.Tensor, top_db: float = 80.0):
    B, T, nb = mel.shape
    n_mfcc = dct.shape[0]
    out = torch.empty(B, T, n_mfcc, dtype=torch.float32, device=mel.device)
    cabi.call("ssar_mfcc", ptr(mel), ptr(max_key), ptr(dct), i32(B), i32(T), i32(nb), i32(n_mfcc), f32(top_db), ptr(out))
    return out


def power_to_db(x: torch.Tensor, n_clips: int = 1, top_db: float | None = 80.0, max_key=None):
    x = x.contiguous()
    out = torch.empty_like(x)
    if top_db is not None and max_key is None:
        _, max_key = minmax_keys(x, n_clips, want_min=False)
    cabi.call("ssar_power_to_db", ptr(x), i32(n_clips), i64(x.numel() // max(n_clips, 1)), ptr(max_key),
              f32(top_db if top_db is not None else 0.0), i32(top_db is not None), ptr(out))
    return out


def hpss(stft_fm: torch.Tensor, margin: float = 1.0, want_harm: bool = True, want_perc: bool = True):
    """stft_fm: frame-major complex STFT as real view [B, T1, F, 2].  Returns (harm, perc) in the same layout."""
    B, T1, F, _ = stft_fm.shape
    h = torch.empty_like(stft_fm) if want_harm else None
    p = torch.empty_like(stft_fm) if want_perc else None
    cabi.call("ssar_hpss", ptr(stft_fm), i32(B), i32(T1), i32(F), i64(F), f32(margin), f32(margin), ptr(h), ptr(p))
    return h, p


def istft(spec_fm: torch.Tensor, n_fft: int, hop: int, window=torch.hann_window, length: int | None = None):
    """spec_fm: frame-major complex [B, T1, F, 2] -> waveform [B, length] (torch.istft semantics, center=True)."""
    B, T1, F, _ = spec_fm.shape
    dev = spec_fm.device
    if length is None:
        length = hop * (T1 - 1)
    win = plan.window_tensor(window, n_fft, dev)
    frames = torch.empty(B, T1, n_fft, dtype=torch.float32, device=dev)
    cabi.call("ssar_irfft_frames", ptr(spec_fm), i64(F), i32(B), i32(T1), i32(n_fft), ptr(win),
              ptr(plan.twiddle(n_fft, str(dev))), ptr(frames))
    out = torch.empty(B, length, dtype=torch.float32, device=dev)
    cabi.call("ssar_overlap_add", ptr(frames), ptr(win), i32(B), i32(T1), i32(n_fft), i32(hop), i64(length), ptr(out))
    return out


def biquad(x: torch.Tensor, coeffs, clamp: bool = True):
    """x [B, N]; coeffs = (b0, b1, b2, a1, a2) already normalised by a0 (torchaudio lfilter semantics)."""
    x, _ = _as_batch(x)
    B, N = x.shape
    y = torch.empty(B, N, dtype=torch.float32, device=x.device)
    nbytes = cabi.lib().ssar_biquad_scratch_bytes(C.c_int(B), C.c_longlong(N))
    scratch = torch.empty(max(int(nbytes), 8), dtype=torch.uint8, device=x.device)
    b0, b1, b2, a1, a2 = coeffs
    cabi.call("ssar_biquad", ptr(x), i64(N), i64(x.stride(0)), i32(B), f32(b0), f32(b1), f32(b2), f32(a1), f32(a2),
              i32(1 if clamp else 0), ptr(y), i64(N), ptr(scratch))
    return y


def resample(x: torch.Tensor, rp, post_scale: float = 1.0):
    """x [B, N] -> [B, ceil(new*N/orig)] with a `plan.ResamplePlan`."""
    x, _ = _as_batch(x)
    B, N = x.shape
    n_out = -((-rp.new * N) // rp.orig)
    y = torch.empty(B, n_out, dtype=torch.float32, device=x.device)
    cabi.call("ssar_resample", ptr(x), i64(N), i64(x.stride(0)), i32(B), ptr(rp.taps), ptr(rp.first), i32(rp.orig),
              i32(rp.new), i32(rp.width), i32(rp.max_taps), f32(post_scale), ptr(y), i64(n_out), i64(n_out))
    return y


def frames_rms(x: torch.Tensor, frame_length: int, hop: int, n_frames: int):
    """features/audio.py:31-37 -> [B, n_frames]."""
    x, _ = _as_batch(x)
    B, N = x.shape
    out = torch.empty(B, n_frames, dtype=torch.float32, device=x.device)
    cabi.call("ssar_rms_frames", ptr(x), i64(N), i64(x.stride(0)), i32(B), i32(frame_length), i32(hop), i32(n_frames), ptr(out))
    return out
