"""Thin Python wrappers over the C-ABI entry points (one function per `ssar_*` symbol).

Everything here takes and returns CUDA tensors; buffers come from torch's caching allocator and kernels run on
torch's current stream.  Layout convention: spectra are frame-major `[clip, frame, bin]`.
"""
from __future__ import annotations

import ctypes as C
import math

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


def pcm16_to_float(pcm: torch.Tensor, out: torch.Tensor | None = None):
    """int16 PCM (CUDA) -> fp32 / 32768 (`ssar_pcm16_to_float`); `out`: optional preallocated fp32 tensor of the same shape."""
    if not pcm.is_cuda or pcm.dtype != torch.int16:
        raise SsarError("pcm16_to_float needs a CUDA int16 tensor")
    pcm = pcm.contiguous()
    if out is None:
        out = torch.empty(pcm.shape, dtype=torch.float32, device=pcm.device)
    elif not out.is_contiguous() or out.shape != pcm.shape or out.dtype != torch.float32:
        raise SsarError("pcm16_to_float: out must be a contiguous fp32 tensor of the input's shape")
    cabi.call("ssar_pcm16_to_float", ptr(pcm), i64(pcm.numel()), ptr(out), nbytes=cabi.nbytes_of(pcm, out))
    return out


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
    proj: dict | None = None,
    aux_frames: int | None = None,
    piptrack: dict | None = None,
    clip_samples: torch.Tensor | None = None,
    clip_frames: torch.Tensor | None = None,
    segment: tuple | None = None,
    sub_every: int = 1,
):
    """Framed real FFT + fused epilogues over a batch of clips.

    banks: up to two `(plan.Banded, on_power: bool, track_max: bool)`.
    Returns a dict with the requested outputs (frame-major):
      complex [B,T,F,2], spec [B,T,F], rms [B,T], flatness [B,T], bank{i} [B,T,n_bands], bank{i}_max uint32 key [B],
    proj: banded complex projection of the spectrum (the CQT response), a dict with
      n_rows, start, count, offset, weights (device tables), out (complex-as-real tensor or view whose element
      [b, t, r] sits at out.data_ptr + ((b*T + t)*ld + r) complex elements), ld, scale, and optionally
      slots (int32 [B]: per-clip table slot), row_scale (fp32 [n_slots, row_scale_ld] view), row_scale_ld.
    aux_frames: when given (<= n_frames), every output except `complex` covers only the first aux_frames frames of each
      clip ([B, aux_frames, ...]): stft (T + 1 frames) and the last-frame-dropping features of one signal in one launch.
    segment: (sample_offset, clip_total, frame_offset) — x holds samples [sample_offset, sample_offset + N) of clips that are
      clip_total long, local frame t = the clips' frame frame_offset + t (time segments of long tracks with n_fft / 2 halos).
    sub_every = s > 1: `complex` and `rms` hold every s-th frame only ([B, ceil(n_frames / s), F, 2] and [B, aux_frames / s]):
      the launch then also serves the consumer of the same signal's transform at hop s * hop.
    clip_samples (int64 [B]) / clip_frames (int32 [B]): RAGGED batch — clip b holds clip_samples[b] <= N samples and
      produces clip_frames[b] <= n_frames frames; rows past a clip's count are zeros in every output.
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
    if segment is not None:  # (sample_offset, clip_total, frame_offset): x is a time segment (with halos) of longer clips
        a.sample_offset, a.clip_total, a.frame_offset = int(segment[0]), int(segment[1]), int(segment[2])
    if clip_samples is not None or clip_frames is not None:
        if clip_samples is None or clip_frames is None or clip_samples.dtype != torch.int64 or clip_frames.dtype != torch.int32:
            raise SsarError("frames_fft: ragged batches need clip_samples (int64) and clip_frames (int32) together")
        a.clip_samples, a.clip_frames = clip_samples.data_ptr(), clip_frames.data_ptr()
    out = {}
    keep = [x, win, tw]
    n_cplx = n_frames
    if aux_frames is not None and int(aux_frames) != n_frames:
        a.aux_frames = int(aux_frames)
        n_frames = int(aux_frames)  # every allocation below except the complex spectrum
    sub = max(1, int(sub_every))
    if sub > 1:
        a.sub_every = sub
    if want_complex:
        out["complex"] = torch.empty(B, -(-n_cplx // sub), F, 2, dtype=torch.float32, device=dev)
        a.out_complex = out["complex"].data_ptr()
        a.ldc = F
    if spec_power is not None:
        out["spec"] = torch.empty(B, n_frames, F, dtype=torch.float32, device=dev)
        a.out_spec = out["spec"].data_ptr()
        a.lds = F
        a.spec_power = int(spec_power)
    if want_rms:
        out["rms"] = torch.empty(B, n_frames // sub, dtype=torch.float32, device=dev)
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
        if getattr(bank, "lanes", None) is not None and on_power:
            ln = bank.lanes
            fb.lane_start4, fb.lane_count4 = ln.start4.data_ptr(), ln.count4.data_ptr()
            fb.lane_offset4, fb.lane_weights = ln.offset4.data_ptr(), ln.weights.data_ptr()
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
    if proj is not None:
        pj = a.proj
        pj.n_rows = int(proj["n_rows"])
        pj.start, pj.count, pj.offset = proj["start"].data_ptr(), proj["count"].data_ptr(), proj["offset"].data_ptr()
        pj.weights = proj["weights"].data_ptr()
        pj.out = proj["out"].data_ptr()
        pj.ld = int(proj["ld"])
        pj.scale = float(proj.get("scale", 1.0))
        if proj.get("slots") is not None:
            pj.clip_slot = proj["slots"].data_ptr()
        if proj.get("row_scale") is not None:
            pj.row_scale = proj["row_scale"].data_ptr()
            pj.row_scale_ld = int(proj["row_scale_ld"])
        keep.append(proj)
    if piptrack is not None:  # dict(k_lo, k_hi, threshold, sr) -> out["pip_pitch"], out["pip_mag"], out["pip_count"]
        cap = int(cabi.lib().ssar_piptrack_capacity(C.c_int(n_frames), C.c_int(piptrack["k_lo"]), C.c_int(piptrack["k_hi"])))
        out["pip_pitch"] = torch.empty(B, cap, dtype=torch.float32, device=dev)
        out["pip_mag"] = torch.empty(B, cap, dtype=torch.float32, device=dev)
        out["pip_count"] = torch.empty(B, dtype=torch.int32, device=dev)
        a.pip.count, a.pip.pitch, a.pip.mag = out["pip_count"].data_ptr(), out["pip_pitch"].data_ptr(), out["pip_mag"].data_ptr()
        a.pip.capacity, a.pip.k_lo, a.pip.k_hi = cap, int(piptrack["k_lo"]), int(piptrack["k_hi"])
        a.pip.threshold, a.pip.sr = float(piptrack["threshold"]), float(piptrack["sr"])
    # (the piptrack lists are capacity-sized; what is written is a few peaks per frame: not counted as algorithmic bytes)
    outs = [v for k, v in out.items() if not k.startswith("pip_")] + ([proj["out"]] if proj is not None else [])
    nb = cabi.nbytes_of(x) + sum(o.numel() * o.element_size() for o in outs if proj is None or o is not proj["out"])
    if proj is not None:
        nb += B * n_frames * int(proj["n_rows"]) * 8
    cabi.call("ssar_frames_fft", C.byref(a), nbytes=nb)
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
    cabi.call("ssar_minmax_keys", ptr(x), i32(n_clips), i64(n_per), ptr(mn), ptr(mx), nbytes=cabi.nbytes_of(x))
    return mn, mx


def onset_flux(mel: torch.Tensor, max_key: torch.Tensor, pad_width: int, aggregate: str = "mean", top_db: float = 80.0):
    """mel [B,T,nb] -> onset envelope [B,T] (rosa/beat.py:10-23)."""
    B, T, nb = mel.shape
    out = torch.empty(B, T, dtype=torch.float32, device=mel.device)
    cabi.call("ssar_onset_flux", ptr(mel), ptr(max_key), i32(B), i32(T), i32(nb), i32(pad_width),
              i32({"mean": 0, "median": 1}[aggregate]), f32(top_db), ptr(out), nbytes=cabi.nbytes_of(mel, out))
    return out


def mfcc_from_mel(mel: torch.Tensor, max_key: torch.Tensor, dct: torch.Tensor, top_db: float = 80.0):
    B, T, nb = mel.shape
    n_mfcc = dct.shape[0]
    out = torch.empty(B, T, n_mfcc, dtype=torch.float32, device=mel.device)
    cabi.call("ssar_mfcc", ptr(mel), ptr(max_key), ptr(dct), i32(B), i32(T), i32(nb), i32(n_mfcc), f32(top_db), ptr(out),
              nbytes=cabi.nbytes_of(mel, out))
    return out


def power_to_db(x: torch.Tensor, n_clips: int = 1, top_db: float | None = 80.0, max_key=None):
    x = x.contiguous()
    out = torch.empty_like(x)
    if top_db is not None and max_key is None:
        _, max_key = minmax_keys(x, n_clips, want_min=False)
    cabi.call("ssar_power_to_db", ptr(x), i32(n_clips), i64(x.numel() // max(n_clips, 1)), ptr(max_key),
              f32(top_db if top_db is not None else 0.0), i32(top_db is not None), ptr(out), nbytes=cabi.nbytes_of(x, out))
    return out


def hpss(stft_fm: torch.Tensor, margin: float = 1.0, want_harm: bool = True, want_perc: bool = True, clip_frames=None):
    """stft_fm: frame-major complex STFT as real view [B, T1, F, 2].  Returns (harm, perc) in the same layout."""
    B, T1, F, _ = stft_fm.shape
    h = torch.empty_like(stft_fm) if want_harm else None
    p = torch.empty_like(stft_fm) if want_perc else None
    cabi.call("ssar_hpss", ptr(stft_fm), i32(B), i32(T1), i32(F), i64(F), f32(margin), f32(margin), ptr(h), ptr(p),
              ptr(clip_frames), nbytes=cabi.nbytes_of(stft_fm, h, p))
    return h, p


def istft(spec_fm: torch.Tensor, n_fft: int, hop: int, window=torch.hann_window, length: int | None = None,
          clip_frames=None, clip_len=None, out: torch.Tensor | None = None):
    """spec_fm: frame-major complex [B, T1, F, 2] -> waveform [B, length] (torch.istft semantics, center=True).
    clip_frames (int32 [B]) / clip_len (int64 [B]): ragged batch (per-clip frame counts and output lengths)."""
    B, T1, F, _ = spec_fm.shape
    dev = spec_fm.device
    if length is None:
        length = hop * (T1 - 1)
    if n_fft not in (1024, 2048):  # arbitrary length: direct inverse DFT over all bins
        return istft_band(spec_fm.contiguous(), n_fft, hop, window, 0, length)
    win = plan.window_tensor(window, n_fft, dev)
    if n_fft == 2048 and hop == 1024 and (length % 2 == 0 or B == 1):  # the HPSS case: fused transform + overlap-add
        # (rows must start 8-byte aligned: an even length, or a single clip)
        if out is None:
            out = torch.empty(B, length, dtype=torch.float32, device=dev)
        elif tuple(out.shape) != (B, length) or not out.is_contiguous() or out.dtype != torch.float32:
            raise SsarError("istft: out must be a contiguous fp32 [B, length] tensor")
        cabi.call("ssar_istft_half_overlap", ptr(spec_fm), i64(F), i32(B), i32(T1), i32(n_fft), ptr(win),
                  ptr(plan.twiddle(n_fft, str(dev))), i64(length), ptr(out), ptr(clip_frames), ptr(clip_len),
                  nbytes=cabi.nbytes_of(spec_fm, out))
        return out
    if out is not None:
        raise SsarError("istft: out= is only supported on the fused n_fft 2048 / hop 1024 path")
    frames = torch.empty(B, T1, n_fft, dtype=torch.float32, device=dev)
    cabi.call("ssar_irfft_frames", ptr(spec_fm), i64(F), i32(B), i32(T1), i32(n_fft), ptr(win),
              ptr(plan.twiddle(n_fft, str(dev))), ptr(frames), nbytes=cabi.nbytes_of(spec_fm, frames))
    out = torch.empty(B, length, dtype=torch.float32, device=dev)
    cabi.call("ssar_overlap_add", ptr(frames), ptr(win), i32(B), i32(T1), i32(n_fft), i32(hop), i64(length), ptr(out),
              ptr(clip_frames), ptr(clip_len), nbytes=cabi.nbytes_of(frames, out))
    return out


def biquad(x: torch.Tensor, coeffs, clamp: bool = True):
    """x [B, N]; coeffs = (b0, b1, b2, a1, a2) already normalised by a0 (torchaudio lfilter semantics)."""
    x, _ = _as_batch(x)
    B, N = x.shape
    y = torch.empty(B, N, dtype=torch.float32, device=x.device)
    nbytes = cabi.lib().ssar_biquad_scratch_bytes(C.c_int(B), C.c_longlong(N))
    scratch = torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=x.device)
    b0, b1, b2, a1, a2 = coeffs
    cabi.call("ssar_biquad", ptr(x), i64(N), i64(x.stride(0)), i32(B), f32(b0), f32(b1), f32(b2), f32(a1), f32(a2),
              i32(1 if clamp else 0), ptr(y), i64(N), ptr(scratch), nbytes=cabi.nbytes_of(x, y))
    return y


def biquad_rms(x: torch.Tensor, coeffs, clamp: bool = True, want_signal: bool = False, n_frames: int | None = None):
    """rms(biquad(x), 2048, 1024)[:, :n_frames] -> [B, n_frames] from the filter's own output pass (per-1024-sample sums
    of squares), without a second pass over the filtered signal; the signal itself is only written with want_signal
    (-> (rms, y)).  Same filter arithmetic as `biquad`."""
    x, _ = _as_batch(x)
    B, N = x.shape
    n_blocks = (N + 1023) // 1024
    T = N // 1024 if n_frames is None else int(n_frames)
    if n_blocks < 2 or T > n_blocks:
        raise SsarError("biquad_rms: clip too short for 2048-sample frames")
    y = torch.empty(B, N, dtype=torch.float32, device=x.device) if want_signal else None
    ss = torch.empty(B, n_blocks + 2, dtype=torch.float32, device=x.device)
    nbytes = cabi.lib().ssar_biquad_scratch_bytes(C.c_int(B), C.c_longlong(N))
    scratch = torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=x.device)
    b0, b1, b2, a1, a2 = coeffs
    cabi.call("ssar_biquad_ss", ptr(x), i64(N), i64(x.stride(0)), i32(B), f32(b0), f32(b1), f32(b2), f32(a1), f32(a2),
              i32(1 if clamp else 0), ptr(y), i64(N), ptr(ss), ptr(scratch), nbytes=cabi.nbytes_of(x, y))
    out = torch.empty(B, T, dtype=torch.float32, device=x.device)
    cabi.call("ssar_rms_from_block_ss", ptr(ss), i64(n_blocks), i32(B), i32(T), ptr(out), nbytes=cabi.nbytes_of(ss, out))
    return (out, y) if want_signal else out


def biquad_pair(x: torch.Tensor, coeffs0, coeffs1, clamp: bool = True, want_signal=(True, True), want_rms=(False, False),
                n_frames: int | None = None):
    """Two biquads over the SAME input in one set of passes (`ssar_biquad_pair`: the input is read once per phase, not
    once per filter).  -> [(y or None, rms [B, n_frames] or None) for the two filters]; same arithmetic per filter as
    `biquad` / `biquad_rms`."""
    x, _ = _as_batch(x)
    B, N = x.shape
    n_blocks = (N + 1023) // 1024
    T = N // 1024 if n_frames is None else int(n_frames)
    if any(want_rms) and (n_blocks < 2 or T > n_blocks):
        raise SsarError("biquad_pair: clip too short for 2048-sample frames")
    ys = [torch.empty(B, N, dtype=torch.float32, device=x.device) if w else None for w in want_signal]
    sss = [torch.empty(B, n_blocks + 2, dtype=torch.float32, device=x.device) if w else None for w in want_rms]
    nbytes = cabi.lib().ssar_biquad_scratch_bytes(C.c_int(B), C.c_longlong(N))
    scratch = torch.empty(max(int(nbytes), 8), dtype=torch.uint8, device=x.device)
    cf = (C.c_float * 10)(*[float(v) for v in (*coeffs0, *coeffs1)])
    cabi.call("ssar_biquad_pair", ptr(x), i64(N), i64(x.stride(0)), i32(B), cf, i32(1 if clamp else 0), ptr(ys[0]), ptr(ys[1]),
              i64(N), ptr(sss[0]), ptr(sss[1]), ptr(scratch), nbytes=cabi.nbytes_of(x, *ys))
    out = []
    for y, ss in zip(ys, sss):
        r = None
        if ss is not None:
            r = torch.empty(B, T, dtype=torch.float32, device=x.device)
            cabi.call("ssar_rms_from_block_ss", ptr(ss), i64(n_blocks), i32(B), i32(T), ptr(r), nbytes=cabi.nbytes_of(ss, r))
        out.append((y, r))
    return out


def biquad_chain(x: torch.Tensor, coeffs0, coeffs1, coeffs2, clamp: bool = True, want_signal=(True, True, True),
                 want_rms=(False, False, False), n_frames: int | None = None, outs=None):
    """biquad(x, coeffs0), biquad(x, coeffs1) and biquad(biquad(x, coeffs1), coeffs2) from one read of x
    (`ssar_biquad_chain`: low / high / mid pass of features/processing.py:141-150).  -> [(y or None, rms [B, n_frames] or
    None)] per filter; the framed RMS of all requested filters comes out of one `ssar_rms_from_block_ss` launch."""
    x, _ = _as_batch(x)
    B, N = x.shape
    n_blocks = (N + 1023) // 1024
    T = N // 1024 if n_frames is None else int(n_frames)
    if any(want_rms) and (n_blocks < 2 or T > n_blocks):
        raise SsarError("biquad_chain: clip too short for 2048-sample frames")
    if outs is not None:  # caller-provided contiguous [B, N] signals (e.g. rows of one stacked buffer)
        for o in outs:
            if o is not None and (tuple(o.shape) != (B, N) or not o.is_contiguous() or o.dtype != torch.float32):
                raise SsarError("biquad_chain: outs must be contiguous fp32 [B, N] tensors")
        ys = list(outs)
    else:
        ys = [torch.empty(B, N, dtype=torch.float32, device=x.device) if w else None for w in want_signal]
    n_rms = sum(bool(w) for w in want_rms)
    ss_all = torch.empty(n_rms, B, n_blocks + 2, dtype=torch.float32, device=x.device) if n_rms else None
    sss, k = [], 0
    for w in want_rms:
        sss.append(ss_all[k] if w else None)
        k += bool(w)
    nbytes = cabi.lib().ssar_biquad_scratch_bytes(C.c_int(B), C.c_longlong(N))
    scratch = torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=x.device)
    cf = (C.c_float * 15)(*[float(v) for v in (*coeffs0, *coeffs1, *coeffs2)])
    cabi.call("ssar_biquad_chain", ptr(x), i64(N), i64(x.stride(0)), i32(B), cf, i32(1 if clamp else 0), ptr(ys[0]), ptr(ys[1]),
              ptr(ys[2]), i64(N), ptr(sss[0]), ptr(sss[1]), ptr(sss[2]), ptr(scratch), nbytes=cabi.nbytes_of(x, *ys))
    rms = None
    if n_rms:
        rms = torch.empty(n_rms, B, T, dtype=torch.float32, device=x.device)
        cabi.call("ssar_rms_from_block_ss", ptr(ss_all), i64(n_blocks), i32(n_rms * B), i32(T), ptr(rms),
                  nbytes=cabi.nbytes_of(ss_all, rms))
    out, k = [], 0
    for y, w in zip(ys, want_rms):
        out.append((y, rms[k] if w else None))
        k += bool(w)
    return out


def resample(x: torch.Tensor, rp, post_scale: float = 1.0, clip_n_in=None):
    """x [B, N] -> [B, ceil(new*N/orig)] with a `plan.ResamplePlan`.  clip_n_in (int64 [B]): ragged batch — clip b holds
    clip_n_in[b] samples followed by zeros; its outputs past ceil(new * clip_n_in[b] / orig) are zeros."""
    x, _ = _as_batch(x)
    B, N = x.shape
    n_out = -((-rp.new * N) // rp.orig)
    y = torch.empty(B, n_out, dtype=torch.float32, device=x.device)
    cabi.call("ssar_resample", ptr(x), i64(N), i64(x.stride(0)), i32(B), ptr(rp.taps), ptr(rp.first), i32(rp.orig),
              i32(rp.new), i32(rp.width), i32(rp.max_taps), f32(post_scale), ptr(y), i64(n_out), i64(n_out), ptr(clip_n_in),
              nbytes=cabi.nbytes_of(x, y))
    return y


def frames_rms(x: torch.Tensor, frame_length: int, hop: int, n_frames: int):
    """features/audio.py:31-37 -> [B, n_frames]."""
    x, _ = _as_batch(x)
    B, N = x.shape
    out = torch.empty(B, n_frames, dtype=torch.float32, device=x.device)
    cabi.call("ssar_rms_frames", ptr(x), i64(N), i64(x.stride(0)), i32(B), i32(frame_length), i32(hop), i32(n_frames), ptr(out),
              nbytes=cabi.nbytes_of(x, out))
    return out


# ------------------------------------------------------------------------------------------------
# post-processing (csrc/post.cu)
# ------------------------------------------------------------------------------------------------
PAD_MODES = {"circular": 0, "reflect": 1, "replicate": 2}


def gauss1d(x: torch.Tensor, sigma: float, mode: str = "circular", batched: bool = False, causal="plain", clip_frames=None):
    """Gaussian smoothing along time.  x: [T, inner...] or, with batched=True, [B, T, inner...] -> same shape.
    causal: "plain" = features/processing.py's kernel; None / number = maua signal.py's `causal` argument."""
    if not x.is_cuda:
        raise SsarError("ssar_b200 needs CUDA tensors (no CPU fallback)")
    x = x.contiguous().float()
    n_clips, T = (x.shape[0], x.shape[1]) if batched else (1, x.shape[0])
    inner = x.numel() // max(n_clips * T, 1)
    radius = min(int(sigma * 4), 3 * T)
    if mode not in PAD_MODES:
        raise SsarError(f"gaussian_filter: unsupported padding mode {mode!r}")
    if mode == "reflect" and min(radius, T) >= T:
        # torch's reflect pad (features/processing.py:30,37) refuses a pad as long as the sequence; so does the mirror
        raise RuntimeError(f"Argument #4: Padding size should be less than the corresponding input dimension, but got: "
                           f"padding ({min(radius, T)}, {min(radius, T)}) at dimension 2 of input of {T} frames")
    w = (plan.gaussian_taps(float(sigma), radius, str(x.device)) if isinstance(causal, str) else
         plan.gaussian_taps_causal(float(sigma), radius, causal, str(x.device)))
    out = torch.empty_like(x)
    cabi.call("ssar_gauss1d", ptr(x), i32(n_clips), i32(T), i64(inner), ptr(w), i32(radius), i32(PAD_MODES[mode]), ptr(out),
              ptr(clip_frames), nbytes=cabi.nbytes_of(x, out))
    return out


def col_minmax(x: torch.Tensor, clip_frames=None):
    """x [B, T, C] -> (min [B, C], max [B, C])."""
    B, T, Cc = x.shape
    mn = torch.empty(B, Cc, dtype=torch.float32, device=x.device)
    mx = torch.empty(B, Cc, dtype=torch.float32, device=x.device)
    cabi.call("ssar_col_minmax", ptr(x), i32(B), i32(T), i32(Cc), ptr(mn), ptr(mx), ptr(clip_frames), nbytes=cabi.nbytes_of(x))
    return mn, mx


def col_quantile(x: torch.Tensor, q: float, subset: str = "all", interp: str = "linear", pre=None, want_count=False,
                 clip_frames=None):
    """x [B, T, C] contiguous -> per-(clip, column) quantile [B, C] (exact order statistics, radix select)."""
    B, T, Cc = x.shape
    out = torch.empty(B, Cc, dtype=torch.float32, device=x.device)
    cnt = torch.empty(B, Cc, dtype=torch.int32, device=x.device) if want_count else None
    ps, pd = (pre if pre is not None else (None, None))
    cabi.call("ssar_col_quantile", ptr(x), i32(B), i32(T), i32(Cc), i32({"all": 0, "peaks": 1}[subset]), f64(q),
              i32({"linear": 0, "midpoint": 3}[interp]), ptr(ps), ptr(pd), ptr(out), ptr(cnt), ptr(clip_frames),
              nbytes=cabi.nbytes_of(x))
    return (out, cnt) if want_count else out


def clamp_cols(x: torch.Tensor, lo=None, hi=None):
    B, T, Cc = x.shape
    out = torch.empty_like(x)
    cabi.call("ssar_clamp_cols", ptr(x), i32(B), i32(T), i32(Cc), ptr(lo), ptr(hi), ptr(out), nbytes=cabi.nbytes_of(x, out))
    return out


def normalize(x: torch.Tensor, n_clips: int = 1, eps: float = 1e-8, clip_rows=None):
    """(x - min) / ((max - min) + eps) with min/max over each clip's whole array.  clip_rows (int32 [n_clips]): ragged
    batch [n_clips, T, ...] — only the first clip_rows[c] rows of clip c are real (the rest comes back as zeros)."""
    x = x.contiguous()
    out = torch.empty_like(x)
    n_per = x.numel() // max(n_clips, 1)
    if 0 < n_per <= (1 << 18) and x.dtype == torch.float32:
        row_elems = n_per // max(x.shape[1], 1) if (clip_rows is not None and x.dim() >= 2) else 0
        cabi.call("ssar_normalize_fused", ptr(x), i32(n_clips), i64(n_per), f32(eps), ptr(out), ptr(clip_rows), i64(row_elems),
                  nbytes=cabi.nbytes_of(x, out))
        return out
    if clip_rows is not None:
        raise SsarError("normalize: ragged batches are limited to 262 144 elements per clip")
    mn, mx = minmax_keys(x, n_clips)
    cabi.call("ssar_normalize", ptr(x), i32(n_clips), i64(x.numel() // max(n_clips, 1)), ptr(mn), ptr(mx), f32(eps), ptr(out),
              nbytes=cabi.nbytes_of(x, out))
    return out


def emphasize(x: torch.Tensor, strength: float, percentile: float, clip_frames=None):
    B, T, Cc = x.shape
    mn, mx = col_minmax(x, clip_frames)
    scratch = torch.empty(2 * B * Cc, dtype=torch.float32, device=x.device)
    out = torch.empty_like(x)
    cabi.call("ssar_emphasize", ptr(x), i32(B), i32(T), i32(Cc), ptr(mn), ptr(mx), f32(strength), f64(percentile / 100),
              ptr(scratch), ptr(out), ptr(clip_frames), nbytes=cabi.nbytes_of(x, out))
    return out


def spectral_flux(x: torch.Tensor, n_clips: int = 1):
    x = x.contiguous()
    T = x.shape[1] if n_clips > 1 else x.shape[0]
    out = torch.empty_like(x)
    cabi.call("ssar_spectral_flux", ptr(x), i32(n_clips), i32(T), i64(x.numel() // max(n_clips * T, 1)), ptr(out),
              nbytes=cabi.nbytes_of(x, out))
    return out


def hwr_rowsum(x: torch.Tensor):
    """x [rows, inner] -> [rows]: sum of the half-wave rectified entries."""
    rows, inner = x.shape
    out = torch.empty(rows, dtype=torch.float32, device=x.device)
    cabi.call("ssar_hwr_rowsum", ptr(x), i64(rows), i32(inner), ptr(out), nbytes=cabi.nbytes_of(x, out))
    return out


_GROUP_TABLES = {}


def _group_tables(group_sizes, device):
    key = (tuple(int(g) for g in group_sizes), str(device))
    t = _GROUP_TABLES.get(key)
    if t is None:
        starts, cols = [], []
        for g, n in enumerate(key[0]):
            starts.append(len(cols))
            cols += [g] * n
        t = _GROUP_TABLES[key] = (torch.tensor(cols, dtype=torch.int32, device=device),
                                  torch.tensor(starts, dtype=torch.int32, device=device), starts)
    return t


def salience_features(raw: torch.Tensor, group_sizes, sigma: float = 2.0, short_sigma: float = 5.0, long_sigma: float = 80.0,
                      eps: float = 1e-8, clip_frames=None):
    """random/mir.py:44 for a batch: raw [B, T, C] holds the features of every clip side by side (column groups of
    `group_sizes`); every group g becomes normalize(salience_weighted(gaussian_filter(raw_g, sigma))) (mir.py:14-22).
    Returns a list of contiguous [B, T, C_g] tensors... as views of ONE feature-blocked buffer [B, T * C]: entry g is
    buf[:, T * start_g : T * (start_g + C_g)].view(B, T, C_g) (contiguous per clip)."""
    B, T, Cc = raw.shape
    if sum(group_sizes) != Cc:
        raise SsarError("salience_features: group sizes do not add up to the column count")
    col_group, group_start, starts = _group_tables(group_sizes, raw.device)
    G = len(group_sizes)
    env = gauss1d(raw, sigma, "circular", batched=True, clip_frames=clip_frames)
    short = gauss1d(env, short_sigma, "reflect", batched=True, clip_frames=clip_frames)
    long = gauss1d(env, long_sigma, "reflect", batched=True, clip_frames=clip_frames)
    keys = torch.empty(2, B, G, dtype=torch.int32, device=raw.device)
    fill_u32(keys[0], KEY_POS_INF)
    fill_u32(keys[1], KEY_NEG_INF)
    w = torch.empty_like(env)
    cabi.call("ssar_salience_weight", ptr(env), ptr(short), ptr(long), i32(B), i32(T), i32(Cc), ptr(col_group), i32(G), ptr(w),
              ptr(keys[0]), ptr(keys[1]), ptr(clip_frames), nbytes=cabi.nbytes_of(env, short, long, w))
    buf = torch.empty(B, T * Cc, dtype=torch.float32, device=raw.device)
    cabi.call("ssar_group_normalize", ptr(w), i32(B), i32(T), i32(Cc), ptr(col_group), ptr(group_start), i32(G), ptr(keys[0]),
              ptr(keys[1]), f32(eps), ptr(buf), nbytes=cabi.nbytes_of(w, buf))
    return [buf[:, T * st : T * (st + n)].view(B, T, n) for st, n in zip(starts, group_sizes)]


def window_offsets(n_rows: int, length: int):
    """Start rows of supervised/data.py:255-258's windows, in its order: for start in range(0, L, L // 4), the splits of
    tensor[start:] into chunks of L with the LAST chunk dropped (full or not)."""
    offs = []
    for start in range(0, length, length // 4):
        n_chunks = -(-(n_rows - start) // length) if n_rows > start else 0
        offs += [start + c * length for c in range(max(n_chunks - 1, 0))]
    return offs


def overlapping_slices(x: torch.Tensor, length: int):
    """supervised/data.py:255-258: x [T, ...] -> [n_windows, length, ...] (one gather launch, `ssar_window_slices`)."""
    if not x.is_cuda:
        raise SsarError("ssar_b200 needs CUDA tensors (no CPU fallback)")
    x = x.contiguous().float()
    T = x.shape[0]
    inner = x.numel() // max(T, 1)
    offs = window_offsets(T, int(length))
    if not offs:
        raise RuntimeError("stack expects a non-empty TensorList")  # torch.stack([]) in the reference
    o = torch.tensor(offs, dtype=torch.int64, device=x.device)
    out = torch.empty((len(offs), int(length)) + tuple(x.shape[1:]), dtype=torch.float32, device=x.device)
    cabi.call("ssar_window_slices", ptr(x), i64(T), i64(inner), ptr(o), i32(len(offs)), i32(length), ptr(out),
              nbytes=2 * cabi.nbytes_of(out))
    return out


def feature_stats(x: torch.Tensor, stats: torch.Tensor | None = None, clip_frames=None):
    """x [..., C] -> fp64 [3, C] = (count, sum, sum of squares) accumulated into `stats` (allocated if None).
    clip_frames (int32 [B]): ragged batch x [B, T, C] — rows past each clip's own count are skipped."""
    Cc = x.shape[-1]
    x = x.contiguous()
    if stats is None:
        stats = torch.zeros(3, Cc, dtype=torch.float64, device=x.device)
    cabi.call("ssar_feature_stats", ptr(x), i64(x.numel() // Cc), i32(Cc), ptr(stats),
              i32(x.shape[1] if clip_frames is not None else 0), ptr(clip_frames), nbytes=cabi.nbytes_of(x))
    return stats


def spectral_contrast(mag: torch.Tensor, bands, linear: bool = False, top_db: float = 80.0):
    """mag [B, T, F] magnitude spectrogram (frame-major) -> contrast [B, T, n_bands+1] (features/audio.py:72-123)."""
    B, T, F = mag.shape
    dev = mag.device
    peak = torch.empty(B, T, bands.n, dtype=torch.float32, device=dev)
    valley = torch.empty_like(peak)
    keys = torch.empty(2, B, dtype=torch.int32, device=dev)
    fill_u32(keys, KEY_NEG_INF)
    cabi.call("ssar_contrast_bands", ptr(mag), i64(F), i32(B), i32(T), i32(bands.n), ptr(bands.start), ptr(bands.count),
              ptr(bands.q), i32(bands.max_count), ptr(peak), ptr(valley), ptr(keys[0]), ptr(keys[1]),
              nbytes=cabi.nbytes_of(mag, peak, valley))
    out = torch.empty_like(peak)
    cabi.call("ssar_contrast_finish", ptr(peak), ptr(valley), ptr(keys[0]), ptr(keys[1]), i32(B), i64(T * bands.n),
              i32(1 if linear else 0), f32(top_db), ptr(out), nbytes=cabi.nbytes_of(peak, valley, out))
    return out


# ------------------------------------------------------------------------------------------------
# tuning / chroma / tonnetz (csrc/tuning.cu, csrc/chroma.cu)
# ------------------------------------------------------------------------------------------------
FUSED_PIPTRACK = True  # piptrack as an epilogue of the hop n_fft/4 transform (False: spectrogram to HBM + ssar_piptrack)


def estimate_tuning(y: torch.Tensor, sr, n_fft: int = 2048, bins_per_octave: int = 12, threshold: float = 0.1,
                    fmin: float = 150.0, fmax: float = 4000.0, lens=None, stft_out: dict | None = None,
                    defer: list | None = None):
    """rosa/pitch.py:9-24 for a batch [B, N] -> (tuning_idx int32 [B] in [0, 100], tuning_val fp32 [B]).
    lens: `ragged.ClipLengths` of a ragged batch (n_fft 2048 only).
    stft_out: a dict that receives by-products of the SAME launch (dense batches, fused epilogue only): the pitch tracker's
    transform runs at hop n_fft / 4, so every 2nd frame is a frame of stft(y, n_fft, hop n_fft / 2) (rosa/spectral.py:10-21) —
    "complex" [B, N // (n_fft / 2) + 1, F, 2] and "rms" [B, N // (n_fft / 2)] (features/audio.py:31-37), bit-equal to a
    frames_fft launch at hop n_fft / 2 with want_complex / want_rms.
    defer: a list -> the statistics kernel (`ssar_tuning_from_peaks`: one CTA per clip, latency-bound) is queued on a side
    stream and a zero-argument `join` is appended to the list; the caller runs work that does not need the tuning (the
    CQT's octave decimations) and calls join() before the first consumer of tuning_idx."""
    y, _ = _as_batch(y)
    B, N = y.shape
    hop = n_fft // 4  # hop_length=None reaches torch.stft's default through rosa/spectral.py:59-62
    T = N // hop
    k_lo, k_hi = plan.piptrack_bin_range(float(sr), n_fft, fmin, fmax)
    if FUSED_PIPTRACK and n_fft in (1024, 2048):
        # the peak picking runs in the transform's epilogue: the [frames, 1025] magnitudes never reach HBM
        rag = dict(clip_samples=lens.n, clip_frames=lens.P) if lens is not None else {}
        if stft_out is not None and lens is None and T % 2 == 0:
            # frames 0 .. T of the hop n_fft / 4 grid (one more than the tracker looks at): complex / rms rows = even frames
            o = frames_fft(y, n_fft, hop, T + 1, torch.hann_window, want_complex=True, want_rms=True, aux_frames=T, sub_every=2,
                           piptrack=dict(k_lo=k_lo, k_hi=k_hi, threshold=threshold, sr=sr))
            stft_out["complex"], stft_out["rms"] = o["complex"], o["rms"]
        else:
            o = frames_fft(y, n_fft, hop, T, torch.hann_window, piptrack=dict(k_lo=k_lo, k_hi=k_hi, threshold=threshold, sr=sr), **rag)
        pitch, mag, count = o["pip_pitch"], o["pip_mag"], o["pip_count"]
        cap = pitch.shape[1]
    else:
        if lens is not None:
            raise SsarError("estimate_tuning: ragged batches need the fused piptrack epilogue (n_fft 1024 / 2048)")
        S = frames_fft(y, n_fft, hop, T, torch.hann_window, spec_power=1)["spec"]
        cap = int(cabi.lib().ssar_piptrack_capacity(C.c_int(T), C.c_int(k_lo), C.c_int(k_hi)))
        pitch = torch.empty(B, cap, dtype=torch.float32, device=y.device)
        mag = torch.empty_like(pitch)
        count = torch.empty(B, dtype=torch.int32, device=y.device)
        cabi.call("ssar_piptrack", ptr(S), i64(S.shape[2]), i32(B), i32(T), i32(S.shape[2]), i32(k_lo), i32(k_hi),
                  f32(threshold), f32(float(sr)), i32(n_fft), i64(cap), ptr(pitch), ptr(mag), ptr(count), nbytes=cabi.nbytes_of(S))
    idx = torch.empty(B, dtype=torch.int32, device=y.device)
    val = torch.empty(B, dtype=torch.float32, device=y.device)
    edges = plan.tuning_edges(str(y.device))
    def launch():
        cabi.call("ssar_tuning_from_peaks", ptr(pitch), ptr(mag), i32(B), i64(cap), ptr(count), ptr(edges),
                  i32(edges.numel() - 1), f32(float(bins_per_octave)), ptr(idx), ptr(val))

    if defer is None:
        launch()
    else:
        main = torch.cuda.current_stream(y.device)
        side = _tuning_stream(y.device)
        side.wait_stream(main)
        with torch.cuda.stream(side):
            launch()
        keep = (pitch, mag, count, edges, idx, val)  # the peak lists stay allocated until the side kernel has been joined

        def join(_keep=keep):
            main.wait_stream(side)

        defer.append(join)
    return idx, val


_TUNING_STREAMS = {}


def _tuning_stream(device):
    import threading

    key = (torch.device(device).index, threading.get_ident())
    st = _TUNING_STREAMS.get(key)
    if st is None:
        st = _TUNING_STREAMS[key] = torch.cuda.Stream(device)
    return st


def chroma_from_cqt(C_fm: torch.Tensor, bins_per_octave: int = 36, n_chroma: int = 12, quantize: bool = False):
    """C_fm: frame-major complex-as-real CQT [B, T, n_bins, 2] -> chroma [B, T, n_chroma] (optionally L1-normalised and
    passed through the CENS soft quantiser)."""
    B, T, nb, _ = C_fm.shape
    dev = str(C_fm.device)
    cmap = plan.chroma_map(nb, bins_per_octave, n_chroma, dev)
    knots, sa, sb, sc, sd = plan.cens_spline(dev)
    out = torch.empty(B, T, n_chroma, dtype=torch.float32, device=C_fm.device)
    cabi.call("ssar_chroma_from_cqt", ptr(C_fm), i64(nb), i32(B), i32(T), i32(nb), ptr(cmap), i32(n_chroma),
              i32(1 if quantize else 0), ptr(knots), ptr(sa), ptr(sb), ptr(sc), ptr(sd), i32(knots.numel()),
              f32(0.25), f32(20.0), ptr(out), nbytes=cabi.nbytes_of(C_fm, out))
    return out


def cens_smooth(q: torch.Tensor, window: torch.Tensor | None, l2: bool = True, clip_frames=None):
    B, T, nc = q.shape
    if window is None:
        window = torch.ones(1, dtype=torch.float32, device=q.device)
    out = torch.empty_like(q)
    cabi.call("ssar_cens_smooth", ptr(q), i32(B), i32(T), i32(nc), ptr(window), i32(window.numel()), i32(1 if l2 else 0),
              ptr(out), ptr(clip_frames), nbytes=cabi.nbytes_of(q, out))
    return out


def tonnetz(chroma: torch.Tensor):
    """chroma [..., T, n_chroma] contiguous -> [..., T, 6]."""
    chroma = chroma.contiguous()
    nc = chroma.shape[-1]
    out = torch.empty(chroma.shape[:-1] + (6,), dtype=torch.float32, device=chroma.device)
    cabi.call("ssar_tonnetz", ptr(chroma), i64(chroma.numel() // nc), i32(nc), ptr(plan.tonnetz_phi(nc, str(chroma.device))),
              i32(6), ptr(out), nbytes=cabi.nbytes_of(chroma, out))
    return out


USE_TENSOR_CORE_CQT = True  # False: per-octave FFT + banded frequency-domain projection (frames_fft epilogue)


def cqt(y: torch.Tensor, sr, hop_length: int = 1024, tuning_idx=None, sync_free: bool = False, tensor_core=None, lens=None,
        join=None):
    """rosa/constantq.py:13-115 for the reference's only configuration (C1, 7 octaves x 36 bins, filter_scale 1,
    sparsity 0.01): returns the frame-major complex-as-real CQT [B, T, 252, 2] with 1/sqrt(length) applied.
    tuning_idx: int32 [B] from estimate_tuning (None -> estimate, like tuning=None in chroma_cqt)."""
    y, _ = _as_batch(y)
    B, N = y.shape
    n_oct, bpo = plan.CQT_N_OCTAVES, plan.CQT_BINS_PER_OCTAVE
    if hop_length % (1 << (n_oct - 1)) != 0:
        raise Exception("hop_length must be a positive integer multiple of 2^{0:d} for {1:d}-octave CQT/VQT".format(
            n_oct - 1, n_oct))
    if float(sr) / (1 << (n_oct - 1)) != int(float(sr) / (1 << (n_oct - 1))):
        raise ValueError("Frequencies must be of integer type to ensure quality resampling computation.")
    use_tc = USE_TENSOR_CORE_CQT if tensor_core is None else bool(tensor_core)
    use_tc = use_tc and hop_length % (8 << (n_oct - 1)) == 0 and N >= (plan.CQT_N_FFT << (n_oct - 1))
    join = list(join) if join else []  # pending `estimate_tuning(defer=...)` joins of a tuning_idx the caller passed in
    if tuning_idx is None:
        tuning_idx, _ = estimate_tuning(y, sr, bins_per_octave=bpo, lens=lens, defer=join if use_tc else None)
    if use_tc:
        return _cqt_tensor_core(y, sr, hop_length, tuning_idx, lens, join)
    for j in join:
        j()
    if lens is not None:
        raise SsarError("cqt: ragged batches run on the tensor-core path only (hop a multiple of 512, clips >= 65 536 samples)")
    tables = plan.cqt_tables(float(sr), str(y.device))
    if sync_free:  # all tuned bases resident; the per-clip choice is a device gather (no read-back: graph-capturable)
        slots = tables.slot_map()[tuning_idx.long()].contiguous()
    else:
        slots = tables.slots_for(tuning_idx.cpu().tolist())  # one small D2H sync per batch: the basis is data-dependent
    # octave signals: y_i = resample(y_{i-1}, /2, kaiser) * sqrt(2)
    sigs, hops = [y], [hop_length]
    rp = plan.resample_plan(2, 1, "kaiser_window", str(y.device))
    for i in range(1, n_oct):
        sigs.append(resample(sigs[-1], rp, post_scale=float(torch.tensor(2.0).sqrt())))
        hops.append(hops[-1] // 2)
    T = min(s.shape[1] // h for s, h in zip(sigs, hops))
    n_bins = n_oct * bpo
    out = torch.empty(B, T, n_bins, 2, dtype=torch.float32, device=y.device)
    for i, (s, h) in enumerate(zip(sigs, hops)):
        col0 = n_bins - bpo * (i + 1)
        # frames beyond T (octaves can be one frame longer) are simply not produced
        proj = dict(n_rows=bpo, start=tables.start, count=tables.count, offset=tables.offset, weights=tables.weights,
                    out=out[:, :, col0:], ld=n_bins, scale=float(2.0 ** i) ** 0.5, slots=slots,
                    row_scale=tables.row_scale[:, col0:], row_scale_ld=n_bins)
        frames_fft(s, plan.CQT_N_FFT, h, T, None, proj=proj)
    return out


def _cqt_tensor_core(y: torch.Tensor, sr, hop_length: int, tuning_idx: torch.Tensor, lens=None, join=()):
    """The 7 octave responses as ONE tcgen05 launch (`ssar_cqt_tc`): time-domain contraction of the overlapping frames
    (TMA tensor maps with row stride hop_i over the reflect-padded fp16 hi / lo signals) with the tuned basis.  Each
    clip's basis is picked on the device by its tuning index: no read-back, graph-capturable."""
    B, N = y.shape
    n_oct, bpo = plan.CQT_N_OCTAVES, plan.CQT_BINS_PER_OCTAVE
    tb = plan.cqt_tc_tables(float(sr), str(y.device))
    sigs = [y.contiguous()]
    rp = plan.resample_plan(2, 1, "kaiser_window", str(y.device))
    for i in range(1, n_oct):
        sigs.append(resample(sigs[-1], rp, post_scale=float(torch.tensor(2.0).sqrt()),
                             clip_n_in=lens.octave_n(i - 1) if lens is not None else None))
    T = min(s.shape[1] // (hop_length >> i) for i, s in enumerate(sigs))
    n_bins = n_oct * bpo
    out = torch.empty(B, T, n_bins, 2, dtype=torch.float32, device=y.device)
    ns = (C.c_longlong * n_oct)(*[s.shape[1] for s in sigs])
    ps = (C.c_void_p * n_oct)(*[s.data_ptr() for s in sigs])
    fn = cabi.lib().ssar_cqt_tc_scratch_bytes
    fn.restype = C.c_longlong
    scratch = torch.empty(int(fn(i32(B), ns, i32(n_oct))), dtype=torch.uint8, device=y.device)
    for j in join:  # the tuning statistics ran on a side stream next to the decimations above
        j()
    slots = tuning_idx.to(torch.int32).contiguous()
    cabi.call("ssar_cqt_tc", ps, ns, i32(n_oct), i32(B), i32(hop_length), i32(T), ptr(tb.basis_hi), ptr(tb.basis_lo),
              i32(tb.n_slots), ptr(slots), ptr(tb.row_scale), i64(n_bins), f32(tb.SCALE), ptr(scratch), ptr(out),
              ptr(lens.n if lens is not None else None), ptr(lens.T if lens is not None else None),
              nbytes=sum(cabi.nbytes_of(s) for s in sigs) + cabi.nbytes_of(out))
    return out


def frames_dft(x: torch.Tensor, n_fft: int, hop: int, n_frames: int, window=torch.hann_window, k_lo: int = 0,
               k_hi: int | None = None):
    """torch.stft(center=True, reflect) for ANY n_fft in [2, 4096], bins [k_lo, k_hi) only (direct DFT, `ssar_frames_dft`)
    -> complex-as-real [B, n_frames, k_hi - k_lo, 2]."""
    x, _ = _as_batch(x)
    B, N = x.shape
    k_hi = n_fft // 2 + 1 if k_hi is None else int(k_hi)
    win = plan.window_tensor(window, n_fft, x.device)
    out = torch.empty(B, n_frames, max(k_hi - k_lo, 0), 2, dtype=torch.float32, device=x.device)
    cabi.call("ssar_frames_dft", ptr(x), i64(N), i64(x.stride(0)), i32(B), i32(n_fft), i32(hop), i32(n_frames), ptr(win),
              i32(k_lo), i32(k_hi), ptr(out), nbytes=cabi.nbytes_of(x, out))
    return out


def istft_band(spec_band: torch.Tensor, n_fft: int, hop: int, window, k_lo: int, length: int):
    """torch.istft(center=True, length) for ANY n_fft from a spectrum that is zero outside bins [k_lo, k_lo + nb):
    spec_band complex-as-real [B, n_frames, nb, 2] -> [B, length] (`ssar_istft_band`)."""
    B, T1, nb, _ = spec_band.shape
    win = plan.window_tensor(window, n_fft, spec_band.device)
    out = torch.empty(B, length, dtype=torch.float32, device=spec_band.device)
    cabi.call("ssar_istft_band", ptr(spec_band), i32(B), i32(T1), i32(n_fft), i32(hop), ptr(win), i32(k_lo), i32(k_lo + nb),
              i64(length), ptr(out), nbytes=cabi.nbytes_of(spec_band, out))
    return out


def tempo_bin_range(sr, hop_length: int, win: int, tempo_min, tempo_max):
    """Bins of the Fourier tempogram that rosa/beat.py:52-57 keeps: [k_lo, k_hi) of linspace(0, rate / 2, 1 + win // 2)."""
    F = win // 2 + 1
    freqs = torch.linspace(0, float(sr * 60 / float(hop_length)) / 2, F)
    keep = torch.ones(F, dtype=torch.bool)
    if tempo_min is not None:
        keep &= ~(freqs < tempo_min)
    if tempo_max is not None:
        keep &= ~(freqs > tempo_max)
    idx = keep.nonzero().flatten()
    return (int(idx[0]), int(idx[-1]) + 1) if len(idx) else (0, 0)


def plp_from_onset(onset_env: torch.Tensor, sr, hop_length: int = 1024, win_length: int = 1024, tempo_min=60, tempo_max=180,
                   lens=None):
    """rosa/beat.py:46-75 after the onset envelope: onset_env [B, T] -> pulse [B, T] (normalised per clip).
    The tempogram window is min(T, win_length) (beat.py:47): 1024 / 2048 run on the radix FFT kernels; any other length
    (a clip shorter than win_length frames) on the band-limited direct DFT, which only ever forms the 60-180 BPM bins."""
    onset_env, _ = _as_batch(onset_env)
    B, T = onset_env.shape
    win = min(T, win_length)
    if lens is not None and (win_length not in (1024, 2048) or min(lens.host_T) < win_length):
        raise SsarError("plp: every clip of a ragged batch needs at least win_length frames (the window would differ per clip)")
    k_lo, k_hi = tempo_bin_range(sr, hop_length, win, tempo_min, tempo_max)
    n_frames = 1 + (T + 2 * (win // 2) - win)  # torch.stft, hop 1, centre-padded by win // 2 on both sides
    if win in (1024, 2048):
        F = win // 2 + 1
        rag = dict(clip_samples=lens.T64, clip_frames=lens.T1) if lens is not None else {}
        tg = frames_fft(onset_env, win, 1, n_frames, torch.hann_window, want_complex=True, **rag)["complex"]  # [B, T+1, F, 2]
        cabi.call("ssar_plp_peaks", ptr(tg), i64(F), i32(B), i32(n_frames), i32(F), i32(k_lo), i32(k_hi),
                  nbytes=2 * cabi.nbytes_of(tg))
        pulse = istft(tg, win, 1, torch.hann_window, T, clip_frames=lens.T1 if lens is not None else None,
                      clip_len=lens.T64 if lens is not None else None)
    else:
        if win < 2:
            raise SsarError("plp: the onset envelope needs at least 2 frames")
        nb = k_hi - k_lo
        tg = frames_dft(onset_env, win, 1, n_frames, torch.hann_window, k_lo, k_hi)  # [B, n_frames, nb, 2]
        if nb > 0:
            cabi.call("ssar_plp_peaks", ptr(tg), i64(nb), i32(B), i32(n_frames), i32(nb), i32(0), i32(nb),
                      nbytes=2 * cabi.nbytes_of(tg))
        pulse = (istft_band(tg, win, 1, torch.hann_window, k_lo, T) if nb > 0 else
                 torch.zeros(B, T, dtype=torch.float32, device=onset_env.device))
    zero = torch.zeros(B, 1, dtype=torch.float32, device=pulse.device)
    pulse = clamp_cols(pulse.view(B, T, 1), zero, None).view(B, T)  # clamp(pulse, 0, pulse.max())
    return normalize(pulse, B, clip_rows=lens.T if lens is not None else None)


# ------------------------------------------------------------------------------------------------
# trajectory construction (csrc/traj_mix.cu, csrc/traj_noise.cu)
# ------------------------------------------------------------------------------------------------
USE_TENSOR_CORE_MIX = True  # False: always the CUDA-core fp32 GEMM (ssar_mix_gemm)


class SplitBasis:
    """A basis [K, ...] (a latent palette) split ONCE into the tensor-core mix's fp16 (hi, lo) panels
    (`ssar_mix_gemm_tc_split_basis`): pass it as `mix_gemm(..., basis_split=...)` for every product with that basis —
    a batch of clips renders against one palette.  Valid while `basis` is not modified."""

    def __init__(self, basis: torch.Tensor):
        if not basis.is_cuda:
            raise SsarError("ssar_b200 needs CUDA tensors (no CPU fallback)")
        self.basis = basis.contiguous().float()
        self.K = self.basis.shape[0]
        self.N = self.basis.numel() // max(self.K, 1)
        self.block = None
        if 0 < self.K <= 192 and self.N % 128 == 0:
            fn = cabi.lib().ssar_mix_gemm_tc_basis_bytes
            fn.restype = C.c_longlong
            self.block = torch.empty(int(fn(i32(self.K), i32(self.N))), dtype=torch.uint8, device=basis.device)
            cabi.call("ssar_mix_gemm_tc_split_basis", ptr(self.basis), i64(self.N), i32(self.K), i32(self.N), ptr(self.block),
                      nbytes=cabi.nbytes_of(self.basis) + self.K * self.N * 4)


def mix_gemm(weights: torch.Tensor, basis: torch.Tensor, out: torch.Tensor | None = None, group_cols: int | None = None,
             col_begin: int = 0, col_end: int | None = None, row_alpha=None, row_beta=None, basis_split: SplitBasis | None = None):
    """out[t, n] = row_alpha[t] * out[t, n] + row_beta[t] * sum_k weights[t, g(n), k] * basis[k, n].

    weights: [T, K] (one group) or [T, G, K]; basis: [K, ...] flattened to [K, N]; out: [T, ...] with N elements per row
    (allocated when None).  g(n) = n // group_cols (default N // G)."""
    if not (weights.is_cuda and basis.is_cuda):
        raise SsarError("ssar_b200 needs CUDA tensors (no CPU fallback)")
    weights = weights.contiguous().float()
    basis = basis.contiguous().float()
    if weights.dim() == 2:
        weights = weights[:, None, :]
    T, G, K = weights.shape
    if basis.shape[0] != K:
        raise SsarError(f"mix_gemm: weights have {K} columns but the basis has {basis.shape[0]} rows")
    N = basis.numel() // max(K, 1)
    if out is None:
        out = torch.empty((T,) + tuple(basis.shape[1:]), dtype=torch.float32, device=basis.device)
    if not out.is_contiguous() or out.numel() != T * N:
        raise SsarError("mix_gemm: out must be a contiguous [T, N] tensor")
    if group_cols is None:
        group_cols = N // G
    if col_end is None:
        col_end = N
    # tensor-core path (tcgen05, fp16 x 3-term split, fp32-accurate): whole-row renders of up to 192 basis rows
    if (USE_TENSOR_CORE_MIX and row_alpha is None and row_beta is None and col_begin == 0 and col_end == N and K <= 192
            and T >= 128 and group_cols % 128 == 0 and N % group_cols == 0 and N // group_cols == G):
        nb = int(cabi.lib().ssar_mix_gemm_tc_scratch_bytes(C.c_int(T), C.c_int(G), C.c_int(K), C.c_int(N)))
        scratch = torch.empty(nb, dtype=torch.uint8, device=out.device)
        pre = None
        if basis_split is not None and basis_split.block is not None:
            if basis_split.K != K or basis_split.N != N:
                raise SsarError("mix_gemm: basis_split was made for another basis shape")
            pre = basis_split.block
        cabi.call("ssar_mix_gemm_tc", ptr(weights), i64(G * K), i64(K), i32(group_cols), ptr(basis), i64(N), ptr(out), i64(N),
                  i32(T), i32(K), i32(N), ptr(pre), ptr(scratch), nbytes=cabi.nbytes_of(weights, basis) + T * N * 4)
        return out
    cabi.call("ssar_mix_gemm", ptr(weights), i64(G * K), i64(K), i32(group_cols), ptr(basis), i64(N), ptr(out), i64(N),
              i32(T), i32(K), i32(col_begin), i32(col_end), ptr(row_alpha), ptr(row_beta),
              nbytes=cabi.nbytes_of(weights, basis) + T * (col_end - col_begin) * 4 * (2 if row_alpha is not None else 1))
    return out


def weights_merge(weights: torch.Tensor, seq: torch.Tensor, g_begin: int, g_end: int, alpha, beta):
    """In place: weights[:, g, :] = alpha * weights[:, g, :] + beta * seq for g in [g_begin, g_end); alpha / beta are
    floats or [T] tensors."""
    T, G, K = weights.shape
    a_t, b_t = (alpha if torch.is_tensor(alpha) else None), (beta if torch.is_tensor(beta) else None)
    cabi.call("ssar_weights_merge", ptr(weights), i32(T), i32(G), i32(K), i32(g_begin), i32(g_end), ptr(seq.contiguous()),
              ptr(a_t), ptr(b_t), f32(0.0 if a_t is not None else alpha), f32(0.0 if b_t is not None else beta),
              nbytes=T * (g_end - g_begin) * K * 8 + cabi.nbytes_of(seq))
    return weights


def subpatch_merge(weights: torch.Tensor, g_begin: int, g_end: int, selection: torch.Tensor, src: torch.Tensor,
                   src_scale: float = 1.0, lerp: bool = False, sigma: float = 1.0, mod: torch.Tensor | None = None,
                   mod_scale: float = 1.0, alpha: float = 0.5, beta: float = 0.5):
    """One latent sub-patch into weights [T, G, P] in place (see `ssar_subpatch_merge`).  selection: int64 [K] on the
    GPU; src: [T, K] (or [T, >=1] with lerp=True: K = 2, S = (f, 1 - f)); sigma: circular Gaussian along time (0 = none);
    mod: [T, >=1] modulation feature (alpha = 1 - mod_scale * mod, beta = mod_scale * mod) or None (scalars)."""
    T, G, P = weights.shape
    K = int(selection.numel())
    if sigma:
        radius = min(int(sigma * 4), 3 * T)
        taps = plan.gaussian_taps(float(sigma), radius, str(weights.device))
    else:
        radius, taps = 0, None
    cabi.call("ssar_subpatch_merge", ptr(weights), i32(T), i32(G), i32(P), i32(g_begin), i32(g_end), ptr(selection), i32(K),
              ptr(src), i64(src.stride(0)), f32(src_scale), i32(1 if lerp else 0), ptr(taps), i32(radius), ptr(mod),
              i64(mod.stride(0) if mod is not None else 0), f32(mod_scale), f32(alpha), f32(beta),
              nbytes=T * (g_end - g_begin) * P * (4 if (mod is None and alpha == 0.0) else 8) + T * K * 4)
    return weights


def onehot_gauss(labels: torch.Tensor, n_classes: int, sigma: float, causal="plain"):
    """labels int64 [T] on the GPU -> circularly smoothed one-hot weights [T, n_classes] (latent.py:39-41)."""
    T = labels.shape[0]
    radius = min(int(sigma * 4), 3 * T)
    taps = (plan.gaussian_taps(float(sigma), radius, str(labels.device)) if isinstance(causal, str) else
            plan.gaussian_taps_causal(float(sigma), radius, causal, str(labels.device)))
    out = torch.empty(T, n_classes, dtype=torch.float32, device=labels.device)
    cabi.call("ssar_onehot_gauss", ptr(labels), i32(T), i32(n_classes), ptr(taps), i32(radius), ptr(out),
              nbytes=cabi.nbytes_of(labels, out))
    return out


def rows_affine(x: torch.Tensor, mu=None, sig=None):
    """out[t, ...] = mu[t] + sig[t] * x[t, ...]."""
    x = x.contiguous()
    rows = x.shape[0]
    out = torch.empty_like(x)
    cabi.call("ssar_rows_affine", ptr(x), i64(rows), i64(x.numel() // max(rows, 1)), ptr(mu), ptr(sig), ptr(out),
              nbytes=cabi.nbytes_of(x, out))
    return out


def noise_program(ops_list, device):
    """A list of `_cabi.NoiseOp` as (host array, device copy, n_ops, n_loops): made once per program, passed to
    `noise_eval` for every frame batch."""
    n = len(ops_list)
    arr = (cabi.NoiseOp * n)(*ops_list)
    dev_ops = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(device)
    return arr, dev_ops, n, sum(1 for o in ops_list if o.kind == 0)


def noise_eval(ops_list, i: int, b: int, height: int, width: int, device):
    """Runs a compiled noise program (list of `_cabi.NoiseOp`, or the tuple `noise_program` made of one) for frames
    [i, i + b) -> [b, height, width]."""
    arr, dev_ops, n, n_loops = ops_list if isinstance(ops_list, tuple) else noise_program(ops_list, device)
    scratch = torch.empty(max(n_loops * b, 1), dtype=torch.float32, device=device)
    out = torch.empty(b, height, width, dtype=torch.float32, device=device)
    cabi.call("ssar_noise_eval", ptr(dev_ops), C.cast(arr, C.c_void_p), i32(n), i32(i), i32(b), i32(height), i32(width),
              ptr(scratch), ptr(out), nbytes=cabi.nbytes_of(out))
    return out


# ------------------------------------------------------------------------------------------------
# batched generator draws (csrc/philox_draws.cu)
# ------------------------------------------------------------------------------------------------
_PHILOX_OK = {}


def philox_normal_increment(numel: int) -> int:
    fn = cabi.lib().ssar_philox_normal_increment
    fn.restype = C.c_longlong
    return int(fn(C.c_longlong(int(numel))))


def _philox_normal_raw(shapes, seed: int, first_offset: int, device, lazy_views: bool = False):
    numels = [int(math.prod(s)) for s in shapes]
    total = sum(numels)
    flat = torch.empty(max(total, 1), dtype=torch.float32, device=device)
    n = len(shapes)
    base = flat.data_ptr()
    ptrs, o = (C.c_void_p * max(n, 1))(), 0
    for i, ne in enumerate(numels):
        ptrs[i] = base + 4 * o
        o += ne
    ne_arr = (C.c_longlong * max(n, 1))(*numels)
    end = C.c_longlong(0)
    cabi.call("ssar_philox_normal_segments", ptrs, ne_arr, i32(n), i64(seed), i64(first_offset), C.byref(end),
              nbytes=4 * total)
    if lazy_views:  # the caller slices `flat` when (if) the tensors are looked at
        return (flat, numels, total), int(end.value)
    return _segment_views(flat, numels, total, shapes), int(end.value)


def _segment_views(flat, numels, total, shapes):
    parts = flat[:total].split(numels) if total else [flat[:0]] * len(shapes)  # one call for all slices
    return [p.view(s) for p, s in zip(parts, shapes)]


def philox_batched_ok(device) -> bool:
    """One-time check per device that the batched draws equal torch's per-call CUDA draws bit for bit (values and
    generator offsets, small / medium / grid-saturating sizes).  When it fails (another torch build with a different
    kernel geometry) the callers keep issuing the per-call torch draws — still on the GPU, just one launch per draw."""
    device = torch.device(device)
    key = device.index if device.index is not None else torch.cuda.current_device()
    ok = _PHILOX_OK.get(key)
    if ok is None:
        with torch.cuda.device(key):
            g = torch.Generator(device)
            g.manual_seed(0x5EED1234)
            g.set_offset(40)
            shapes = [(3, 4, 4), (2, 7, 64, 64), (3, 256, 256), (2, 20, 256, 256)]
            ref = [torch.randn(s, generator=g, device=device) for s in shapes]
            got, end = _philox_normal_raw(shapes, 0x5EED1234, 40, device)
            ok = end == g.get_offset() and all(torch.equal(a, b) for a, b in zip(ref, got))
        _PHILOX_OK[key] = ok
        if not ok:
            import warnings

            warnings.warn("ssar_b200: batched Philox draws differ from this torch build's per-call draws; "
                          "using one torch call per draw")
    return ok


def philox_normal_segments(shapes, rng: torch.Generator):
    """[torch.randn(s, generator=rng, device=rng.device) for s in shapes] in ONE launch (a few when there are more than
    96 shapes); advances `rng` exactly as the per-call sequence would.  CUDA generators only."""
    dev = rng.device
    views, end = _philox_normal_raw(shapes, rng.initial_seed(), rng.get_offset(), dev)
    rng.set_offset(end)
    return views


def philox_normal_segments_deferred(shapes, rng: torch.Generator):
    """Same launch and generator advance as `philox_normal_segments`; returns a zero-argument function that slices the
    drawn buffer into the per-shape tensors when called (the ~100 view objects of a Patch cost more host time than the
    launch itself, and lazy noise modules may never be evaluated)."""
    (flat, numels, total), end = _philox_normal_raw(shapes, rng.initial_seed(), rng.get_offset(), rng.device, lazy_views=True)
    rng.set_offset(end)
    shapes = list(shapes)
    return lambda: _segment_views(flat, numels, total, shapes)


def patch_sample(seeds: torch.Tensor, program: torch.Tensor, n_head: int, n_lat: int, n_noise: int, min_sub: int,
                 max_sub: int, capacity: int):
    """seeds int64 [B] (device), program int32 [2 * n_ops] (device) -> (floats [B, capacity], counts int32 [B, 3],
    end_offset int64 [B]) — see `ssar_patch_sample`."""
    B = seeds.numel()
    out = torch.empty(B, capacity, dtype=torch.float32, device=seeds.device)
    counts = torch.empty(B, 3, dtype=torch.int32, device=seeds.device)
    end = torch.empty(B, dtype=torch.int64, device=seeds.device)
    cabi.call("ssar_patch_sample", ptr(seeds), i32(B), ptr(program), i32(n_head), i32(n_lat), i32(n_noise), i32(min_sub),
              i32(max_sub), i32(capacity), ptr(out), ptr(counts), ptr(end))
    return out, counts, end


_RANDPERM = {}


def randperm_bits(n: int) -> int:
    """Key width ATen's randperm_out_cuda uses for n elements (TensorFactories.cu, 'Algorithm of randperm')."""
    nd = float(n)
    return min(64, int(math.ceil(math.log2(nd - (6 * nd * nd + 1) / (math.log(0.9) * 12)))))


def _philox_randperm_raw(seed: int, first_offset: int, stride: int, n: int, count: int, device):
    out = torch.empty(count, n, dtype=torch.int64, device=device)
    cabi.call("ssar_philox_randperm", i64(seed), i64(first_offset), i64(stride), i32(n), i32(randperm_bits(n)), i32(count),
              ptr(out), nbytes=8 * count * n)
    return out


def philox_randperm_stride(device, n: int):
    """Generator-offset increment of one torch.randperm(n) on `device` when `ssar_philox_randperm` reproduces torch's
    permutations bit for bit (checked once per (device, n) on 256 consecutive calls: ~15 of them contain runs of equal
    keys and exercise the duplicate shuffle), else None — callers then keep calling torch.randperm."""
    device = torch.device(device)
    key = (device.index if device.index is not None else torch.cuda.current_device(), int(n))
    st = _RANDPERM.get(key, 0)
    if st == 0:
        st = None
        try:
            if 0 < n <= 1024 and randperm_bits(n) <= 32:
                with torch.cuda.device(key[0]):
                    g = torch.Generator(device)
                    g.manual_seed(0xC0FFEE)
                    first = torch.randperm(n, generator=g, device=device)
                    inc = g.get_offset()
                    ref = torch.stack([first] + [torch.randperm(n, generator=g, device=device) for _ in range(255)])
                    got = _philox_randperm_raw(0xC0FFEE, 0, inc, n, 256, device)
                    if inc > 0 and inc % 4 == 0 and g.get_offset() == 256 * inc and torch.equal(ref, got):
                        st = int(inc)
        except Exception:  # pragma: no cover
            st = None
        _RANDPERM[key] = st
        if st is None:
            import warnings

            warnings.warn("ssar_b200: batched randperm differs from this torch build's torch.randperm; using torch.randperm")
    return st


def philox_randperm(rng: torch.Generator, n: int, count: int, stride: int):
    """torch.stack([torch.randperm(n, generator=rng, device=rng.device) for _ in range(count)]) in one launch;
    advances `rng` as the calls would.  stride: `philox_randperm_stride(rng.device, n)` (not None)."""
    first = rng.get_offset()
    out = _philox_randperm_raw(rng.initial_seed(), first, stride, n, count, rng.device)
    rng.set_offset(first + count * stride)
    return out


STACK_MAX = 24  # sub-patches per ssar_latent_stack launch


def latent_stack(weights: torch.Tensor, descs, taps: torch.Tensor | None, load_initial: bool):
    """Applies a list of `_cabi.SubpatchDesc` to weights [T, G, P] in place, in order, one launch per 24 descriptors."""
    T, G, P = weights.shape
    for c0 in range(0, len(descs), STACK_MAX):
        part = descs[c0 : c0 + STACK_MAX]
        arr = (cabi.SubpatchDesc * len(part))(*part)
        cabi.call("ssar_latent_stack", ptr(weights), i32(T), i32(G), i32(P), arr, i32(len(part)), ptr(taps),
                  i32(1 if (load_initial or c0 > 0) else 0), nbytes=T * G * P * (8 if (load_initial or c0 > 0) else 4))
    return weights


def slerp_loops(y: torch.Tensor, size: int, n_loops: int):
    """maua audioreactive/latent.py:69-81: y [K, W, L] -> [size, W, L] (see `ssar_slerp_loops`)."""
    y = y.contiguous().float()
    K, Wn, L = y.shape
    n_t = round(size / (K * n_loops + 1))  # Python round, like the reference
    if n_t < 1:
        raise RuntimeError("slerp_loops: size is smaller than half the number of knots (linspace of 0 steps)")
    fn = cabi.lib().ssar_slerp_loops_scratch_floats
    fn.restype = C.c_longlong
    scratch = torch.empty(int(fn(i32(K), i32(n_loops), i32(Wn), i32(L))), dtype=torch.float32, device=y.device)
    out = torch.empty(size, Wn, L, dtype=torch.float32, device=y.device)
    cabi.call("ssar_slerp_loops", ptr(y), i32(K), i32(Wn), i32(L), i32(n_loops), i32(n_t), i32(size), ptr(scratch), ptr(out),
              nbytes=cabi.nbytes_of(y, out))
    return out
