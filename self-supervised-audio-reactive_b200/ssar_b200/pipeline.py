"""Batched entry points for the BASELINE.json configurations (one call = all clips of a batch).

These compose the same kernels the per-function mirrors use, but with every clip of the batch in one launch
and with everything that can share an FFT fused into that FFT's epilogue.
"""
from __future__ import annotations

import functools
import threading

import torch

from . import ops, plan


def spectral_frontend(x: torch.Tensor, sr, n_fft: int = 2048, hop: int = 512, want_spec: bool = True):
    """BASELINE config 2: STFT power spectrogram + mel + onset strength + RMS for a batch of clips.

    x: [B, N] fp32 on the GPU.  Returns frame-major tensors:
      spec  [B, T, 1+n_fft/2]  == spectrogram(y, n_fft, hop, power=2)      (rosa/spectral.py:59-62)
      mel   [B, T, 128]        == melspectrogram(y, sr, n_fft, hop)        (rosa/spectral.py:65-70)
      onset [B, T]             == onset_strength(y, sr, hop, n_fft)        (rosa/beat.py:10-23)
      rms   [B, T]             == rms(y, sr, n_fft, hop)[:, 0]             (features/audio.py:31-37)
    with T = N // hop.  Two launches: the fused frame kernel, then the tiny dB/flux kernel (it needs the
    clip-global mel maximum, which the first kernel accumulates in its epilogue).
    """
    dev = str(x.device)
    T = x.shape[-1] // hop
    bank = plan.mel_bank(float(sr), n_fft, 128, 0.0, None, dev)
    bank_onset = plan.mel_bank(float(sr), n_fft, 128, 0.0, 11025.0, dev)
    out = ops.frames_fft(
        x, n_fft, hop, T, torch.hann_window,
        spec_power=2 if want_spec else None, want_rms=True,
        banks=[(bank, True, False), (bank_onset, True, True)],
    )
    onset = ops.onset_flux(out["bank1"], out["bank1_max"], 1 + n_fft // (2 * hop), "mean")
    return {"spec": out.get("spec"), "mel": out["bank0"], "onset": onset, "rms": out["rms"]}


def spectral_frontend_segmented(x: torch.Tensor, sr, n_fft: int = 2048, hop: int = 512, seg_frames: int = 4096,
                                want_spec: bool = True, device=None):
    """`spectral_frontend` for LONG tracks cut into time segments with FFT-window halos (north star: "long tracks are split
    by time segment with FFT-window halos"): each segment of seg_frames frames is transformed from the samples
    [f0 * hop - n_fft / 2, (f1 - 1) * hop + n_fft / 2] (clamped to the track; reflection only at the track's true ends).
    x [B, N] may live in (pinned) HOST memory: segment s + 1 crosses PCIe on a copy stream while segment s is
    transformed, and only one segment (+ halo) of the waveform is resident at a time.  The one track-global quantity of
    this front end — power_to_db's maximum over the whole mel spectrogram (rosa/beat.py:14-15) — is the maximum of the
    segments' maxima, so the onset envelope is computed after the last segment.  Same outputs as `spectral_frontend`,
    bit for bit (tests/test_gpu_frontend.py)."""
    B, N = x.shape
    dev = torch.device(device) if device is not None else (x.device if x.is_cuda else torch.device("cuda"))
    T = N // hop
    M = n_fft // 2
    bank = plan.mel_bank(float(sr), n_fft, 128, 0.0, None, str(dev))
    bank_onset = plan.mel_bank(float(sr), n_fft, 128, 0.0, 11025.0, str(dev))
    spec = torch.empty(B, T, n_fft // 2 + 1, dtype=torch.float32, device=dev) if want_spec else None
    mel = torch.empty(B, T, 128, dtype=torch.float32, device=dev)
    mel_on = torch.empty(B, T, 128, dtype=torch.float32, device=dev)
    rms = torch.empty(B, T, dtype=torch.float32, device=dev)
    max_key = None
    main = torch.cuda.current_stream(dev)
    copy_stream = _side_stream(dev, "h2d") if not x.is_cuda else None
    bounds = [(f0, min(f0 + seg_frames, T)) for f0 in range(0, T, seg_frames)]

    def fetch(f0, f1):
        lo, hi = max(0, f0 * hop - M), min(N, (f1 - 1) * hop + M)
        if f0 == 0:
            hi = max(hi, min(N, M + 1))          # reflection of the leading pad reads samples 1 .. n_fft / 2
        if (f1 - 1) * hop + M > N:
            lo = min(lo, max(0, N - 1 - M))      # ... and of the trailing pad the last n_fft / 2
        if x.is_cuda:
            return x[:, lo:hi], lo, None
        with torch.cuda.stream(copy_stream):
            seg = torch.empty(B, hi - lo, dtype=torch.float32, device=dev)
            seg.copy_(x[:, lo:hi], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return seg, lo, ev

    nxt = fetch(*bounds[0])
    for si, (f0, f1) in enumerate(bounds):
        seg, lo, ev = nxt
        nxt = fetch(*bounds[si + 1]) if si + 1 < len(bounds) else None
        if ev is not None:
            main.wait_event(ev)
            seg.record_stream(main)
        o = ops.frames_fft(seg, n_fft, hop, f1 - f0, torch.hann_window, spec_power=2 if want_spec else None, want_rms=True,
                           banks=[(bank, True, False), (bank_onset, True, True)], segment=(lo, N, f0))
        if want_spec:
            spec[:, f0:f1] = o["spec"]
        mel[:, f0:f1], mel_on[:, f0:f1], rms[:, f0:f1] = o["bank0"], o["bank1"], o["rms"]
        # order-preserving keys: the running maximum is an integer maximum (keys are stored as int32 bit patterns)
        k = o["bank1_max"].view(torch.int32).to(torch.int64) & 0xFFFFFFFF
        max_key = k if max_key is None else torch.maximum(max_key, k)
    key32 = torch.where(max_key >= 2 ** 31, max_key - 2 ** 32, max_key).to(torch.int32)
    onset = ops.onset_flux(mel_on, key32, 1 + n_fft // (2 * hop), "mean")
    return {"spec": spec, "mel": mel, "onset": onset, "rms": rms}


def hpss_pair(x: torch.Tensor, margin: float = 8.0, n_fft: int = 2048, hop: int = 1024, D=None, want_harm=True,
              want_perc=True, lens=None, perc_out=None):
    """harmonic(x), percussive(x) (features/audio.py:13-24) from ONE forward STFT and ONE pair of median passes.
    x [B, N] -> (harm [B, N], perc [B, N]) (None for a part that is not wanted).  Exact: the two reference calls differ
    only in the mask applied."""
    n = x.shape[-1]
    rag = dict(clip_samples=lens.n, clip_frames=lens.T1) if lens is not None else {}
    if D is None:
        D = ops.frames_fft(x, n_fft, hop, 1 + n // hop, torch.hann_window, want_complex=True, **rag)["complex"]
    h, p = ops.hpss(D, margin, want_harm, want_perc, clip_frames=lens.T1 if lens is not None else None)
    irag = dict(clip_frames=lens.T1, clip_len=lens.n) if lens is not None else {}
    return (ops.istft(h, n_fft, hop, torch.hann_window, n, **irag) if want_harm else None,
            ops.istft(p, n_fft, hop, torch.hann_window, n, out=perc_out, **irag) if want_perc else None)


def band_split(x: torch.Tensor, sr):
    """low_pass(x), mid_pass(x), high_pass(x) (features/processing.py:141-150) from ONE read of x: mid_pass is literally
    low_pass(high_pass(x, 4000), 200), and high_pass(x, 4000) is also high_pass's default."""
    lo_c, hi_c = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
    (low, _), (high, _), (mid, _) = ops.biquad_chain(x, lo_c, hi_c, lo_c)
    return low, mid, high


def band_rms(x: torch.Tensor, sr, n_frames: int):
    """rms(low_pass(x)), rms(mid_pass(x)), rms(high_pass(x)) (frame 2048 / hop 1024) -> three [B, n_frames] tensors, from
    the filters' own output sweeps: no filtered signal is ever written."""
    lo_c, hi_c = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
    (_, r_lo), (_, r_hi), (_, r_mid) = ops.biquad_chain(x, lo_c, hi_c, lo_c, want_signal=(False, False, False),
                                                        want_rms=(True, True, True), n_frames=n_frames)
    return r_lo, r_mid, r_hi


def onset_pair(x: torch.Tensor, sr, hop: int = 1024, n_fft: int = 2048, want_median: bool = False, lens=None):
    """onset_strength(x) with mean aggregation (+ the median-aggregated envelope plp needs) from one mel pass."""
    bank = plan.mel_bank(float(sr), n_fft, 128, 0.0, 11025.0, str(x.device))
    rag = dict(clip_samples=lens.n, clip_frames=lens.T) if lens is not None else {}
    out = ops.frames_fft(x, n_fft, hop, x.shape[-1] // hop, torch.hann_window, banks=[(bank, True, True)], **rag)
    pad = 1 + n_fft // (2 * hop)
    mean = ops.onset_flux(out["bank0"], out["bank0_max"], pad, "mean")
    med = ops.onset_flux(out["bank0"], out["bank0_max"], pad, "median") if want_median else None
    return mean, med


def _check_peak_counts(cnt):
    if int(cnt.min()) == 0:  # torch.quantile raises on a column without interior peaks (processing.py:117)
        raise RuntimeError("quantile() input tensor must be non-empty")


# random/mir.py:8 (AFEATFNS, in its order) with each feature's column count: the inputs of `Patch`
MIR_GROUPS = (("chromagram", 12), ("tonnetz", 6), ("mfcc", 20), ("spectral_contrast", 7), ("spectral_flatness", 1),
              ("rms", 1), ("drop_strength", 1), ("onsets", 1))


def _cens_and_tonnetz(y: torch.Tensor, sr, hop: int, sync_free: bool, lens=None, stft_out=None):
    """chroma_cens(y, sr).T and the tonnetz of it (rosa/spectral.py:239-280, features/audio.py:48-57) -> [B, T, 12], [B, T, 6].
    The tuning is estimated from `y` itself (chroma_cqt's tuning=None).  stft_out: see ops.estimate_tuning (the pitch tracker's
    launch can hand out stft(y, 2048, 1024) and rms(y) as by-products)."""
    idx, join = None, []
    if stft_out is not None:
        idx, _ = ops.estimate_tuning(y, sr, bins_per_octave=plan.CQT_BINS_PER_OCTAVE, lens=lens, stft_out=stft_out, defer=join)
    C = ops.cqt(y, sr, hop, tuning_idx=idx, sync_free=sync_free, lens=lens, join=join)
    cens = ops.cens_smooth(ops.chroma_from_cqt(C, quantize=True), plan.cens_window(41, str(y.device)), l2=True,
                           clip_frames=lens.T if lens is not None else None)
    return cens, ops.tonnetz(cens)


def feature_chain(x: torch.Tensor, sr, want_a2f: bool = True, want_mir: bool = False, sync_free: bool = False,
                  pulse_of: str = "perc", lens=None):
    """Every per-frame feature of a batch of clips x [B, N] that is ALREADY at sr = 1024 * fps (n_fft 2048, hop 1024):

      a2f: the 59 un-clamped, un-smoothed columns of supervised/data.py:64-99 -> [B, T, 59]
      mir: the 8 raw feature functions random/mir.py:26 calls on the audio itself, side by side -> [B, T, 49] in
           MIR_GROUPS order (before mir.py:44's smoothing / salience weighting / normalisation)

    Same arithmetic per clip as the reference's call sequences; work the reference repeats on identical inputs is done
    once: every signal is transformed ONCE per (n_fft, hop) — the launch over `x` emits the complex STFT the HPSS needs
    (T + 1 frames) together with magnitude (contrast), mel (mfcc), flatness and RMS of the first T frames; harmonic /
    percussive share the STFT and the medians; the three band filters share the 4 kHz high-pass.  The two paths differ in
    which signal the chroma features see: mir.py's chromagram(audio) = chroma_cens(harmonic(audio)), while data.py:68-69
    passes audio_harm in, i.e. chroma_cens(harmonic(harmonic(audio))) — both are built when both are wanted.
    pulse_of: "perc" = plp(audio_perc) (data.py:89), "audio" = plp(audio) (features/generate.py:123).
    lens: `ragged.ClipLengths` when the rows of x are clips of DIFFERENT lengths (valid samples followed by zeros): every
    kernel then works on each clip's own length and the rows past a clip's frame count come back as zeros / don't-care."""
    B, N = x.shape
    n_fft, hop = 2048, 1024
    T = N // hop
    dev = str(x.device)

    # --- one pass over the audio: complex STFT (T + 1 frames) + mel(mfcc) / magnitude(contrast) / flatness / rms (T frames)
    mel_bank = plan.mel_bank(float(sr), n_fft, 128, 0.0, None, dev)
    rag1 = dict(clip_samples=lens.n, clip_frames=lens.T1) if lens is not None else {}
    cf = lens.T if lens is not None else None
    fa = ops.frames_fft(x, n_fft, hop, T + 1, torch.hann_window, want_complex=True, spec_power=1, want_flatness=True,
                        want_rms=True, banks=[(mel_bank, True, True)], aux_frames=T, **rag1)
    mfcc = ops.mfcc_from_mel(fa["bank0"], fa["bank0_max"], plan.dct_matrix(20, 128, dev))
    contrast = ops.spectral_contrast(fa["spec"], plan.contrast_bands(float(sr), n_fft, 200.0, 6, 0.02, dev))
    flatness = fa["flatness"]

    # --- harmonic / percussive separation
    # a2f, dense batch: the percussive part and its three band signals sit in ONE [4, B, N] buffer, so their onset
    # envelopes (supervised/data.py:70-88) come out of one transform launch and one flux launch over 4 B rows
    pstack = torch.empty(4, B, N, dtype=torch.float32, device=x.device) if (want_a2f and lens is None and N % 2 == 0) else None
    harm, perc = hpss_pair(x, D=fa["complex"], lens=lens, perc_out=pstack[0] if pstack is not None else None)
    del fa["complex"], fa["spec"]
    if pstack is not None:
        lo_c, hi_c = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
        ops.biquad_chain(perc, lo_c, hi_c, lo_c, outs=(pstack[1], pstack[3], pstack[2]))  # rows: perc, low, mid, high
        bank = plan.mel_bank(float(sr), n_fft, 128, 0.0, 11025.0, dev)
        pm = ops.frames_fft(pstack.view(4 * B, N), n_fft, hop, T, torch.hann_window, banks=[(bank, True, True)])
        pad = 1 + n_fft // (2 * hop)
        on4 = ops.onset_flux(pm["bank0"], pm["bank0_max"], pad, "mean").view(4, B, T)
        on, on_lo, on_mid, on_hi = on4[0], on4[1], on4[2], on4[3]
        on_med = ops.onset_flux(pm["bank0"][:B], pm["bank0_max"][:B], pad, "median") if pulse_of == "perc" else None
        del pm
    else:
        on, on_med = onset_pair(perc, sr, hop, n_fft, want_median=want_a2f and pulse_of == "perc", lens=lens)
    if want_a2f and pulse_of != "perc":
        on_med = onset_pair(x, sr, hop, n_fft, want_median=True, lens=lens)[1]

    out = {}
    cens1 = ton1 = None
    harm_stft = {} if (want_mir and want_a2f and lens is None and n_fft == 2 * hop) else None
    if want_mir:
        cens1, ton1 = _cens_and_tonnetz(harm, sr, hop, sync_free, lens, stft_out=harm_stft)  # chromagram(audio), tonnetz(audio)
    if want_a2f or want_mir:
        rms_x_bands = band_rms(x, sr, T) if want_a2f else ()
        rms_a = torch.stack([fa["rms"], *rms_x_bands], dim=-1)  # [B, T, 4] (or [B, T, 1])
        drop = ops.emphasize(ops.gauss1d(rms_a, 10, "circular", batched=True, clip_frames=cf), 10, 50, clip_frames=cf)
    if want_a2f:
        # harmonic(audio_harm): features/audio.py:44-45 applied to data.py:68-69's argument; the same launch gives rms(harm)
        if harm_stft:  # the tuning estimate of `harm` above ran its STFT at hop 512: its even frames are this transform
            fh = harm_stft
        else:
            fh = ops.frames_fft(harm, n_fft, hop, T + 1, torch.hann_window, want_complex=True, want_rms=True, aux_frames=T, **rag1)
        harm2 = hpss_pair(harm, D=fh["complex"], want_perc=False, lens=lens)[0]
        del fh["complex"]
        cens, ton = _cens_and_tonnetz(harm2, sr, hop, sync_free, lens)
        # onset envelopes of the percussive part's three bands; pulse
        if pstack is None:
            p_lo, p_mid, p_hi = band_split(perc, sr)
            on_lo, on_mid, on_hi = (onset_pair(b, sr, hop, n_fft, lens=lens)[0] for b in (p_lo, p_mid, p_hi))
        pulse = ops.plp_from_onset(on_med, sr, hop, lens=lens)
        # band signals of harm / x are only ever looked at through their RMS: taken from the filters' output passes
        rms_h = [fh["rms"], *band_rms(harm, sr, T)]
        Tm = min(T, cens.shape[1])
        cols = [mfcc[:, :Tm], cens[:, :Tm], ton[:, :Tm], contrast[:, :Tm]]
        cols += [s[:, :Tm, None] for s in (flatness, on, on_lo, on_mid, on_hi, pulse, *rms_h)]
        cols.append(drop[:, :Tm])
        out["a2f"] = torch.cat(cols, dim=2).contiguous()
    if want_mir:
        Tm = min(T, cens1.shape[1])
        onsets = ops.normalize(on, B, clip_rows=cf)  # features/audio.py:27-28
        cols = [cens1[:, :Tm], ton1[:, :Tm], mfcc[:, :Tm], contrast[:, :Tm]]
        cols += [s[:, :Tm, None] for s in (flatness, fa["rms"], drop[:, :, 0], onsets)]
        out["mir"] = torch.cat(cols, dim=2).contiguous()
    return out


def finish_features(feats: torch.Tensor, fps, clamp=True, smooth=True, emphasis=False, deferred_checks=None, lens=None):
    """supervised/data.py:100-106 on [B, T, 59]: peak-percentile clamp, lower-percentile clamp, smoothing, emphasis."""
    cf = lens.T if lens is not None else None
    if clamp:
        thr, cnt = ops.col_quantile(feats, (100 - 2.5) / 100, "peaks", "linear", want_count=True, clip_frames=cf)
        feats = ops.clamp_cols(feats, None, thr)
        feats = ops.clamp_cols(feats, ops.col_quantile(feats, 4 * 2.5 / 100, clip_frames=cf), None)
        # the reference raises when a column has no interior peak; reading the count back synchronises the stream, so a
        # batched caller may collect the check and run it after it has queued the rest of its work
        if deferred_checks is None:
            _check_peak_counts(cnt)
        else:
            deferred_checks.append(functools.partial(_check_peak_counts, cnt))
    if smooth:
        feats = ops.gauss1d(feats, 0.1 * fps, "circular", batched=True, clip_frames=cf)
    if emphasis:
        feats = ops.emphasize(feats, 2, 75, clip_frames=cf)
    return feats


def patch_inputs_from_raw(mir_raw: torch.Tensor, lens=None):
    """random/mir.py:44 for a batch: [B, T, 49] raw features (MIR_GROUPS order) -> dict name -> [B, T, C] of
    normalize(salience_weighted(gaussian_filter(f, sigma=2))), each (clip, feature) block normalised on its own."""
    parts = ops.salience_features(mir_raw, [n for _, n in MIR_GROUPS], sigma=2.0, short_sigma=5.0, long_sigma=80.0,
                                  clip_frames=lens.T if lens is not None else None)
    return {name: p for (name, _), p in zip(MIR_GROUPS, parts)}


def audio2features_batch(audio: torch.Tensor, sr, fps, clamp=True, smooth=True, emphasis=False, deferred_checks=None,
                         sync_free=False):
    """supervised/data.py:60-107 for a batch of mono clips [B, N] at `sr` -> [B, T, 59]."""
    from .resampling import resample

    if audio.dim() != 2:
        raise ValueError("audio2features_batch expects [clips, samples]")
    x = resample(audio, sr, fps * 1024)
    raw = feature_chain(x, fps * 1024, want_a2f=True, want_mir=False, sync_free=sync_free)["a2f"]
    return finish_features(raw, fps, clamp, smooth, emphasis, deferred_checks)


def music_features_batch(x: torch.Tensor, sr, sync_free=False):
    """The feature part of random/mir.py:25-46 (`retrieve_music_information`) for a batch of clips x [B, N] at sr:
    dict name -> [B, T, C], the `features` argument of `Patch`.  Tempo, beats and the Laplacian segmentations
    (mir.py:28-42: librosa / scipy on the CPU) are inputs of `Patch` here, not computed."""
    return patch_inputs_from_raw(feature_chain(x, sr, want_a2f=False, want_mir=True, sync_free=sync_free)["mir"])


def features_and_patch_inputs(audio: torch.Tensor, sr, fps, deferred_checks=None, sync_free=False, lens=None):
    """One pass over a batch of clips [B, N] at `sr` for both consumers of the north-star unit: the dataset features
    `audio2features(audio, sr, fps)` [B, T, 59] (supervised/data.py:60-107) and the `Patch` inputs
    `retrieve_music_information(resample(audio, sr, 1024 * fps), 1024 * fps)[0]` (random/sample.py:34,61; mir.py:26,44)."""
    from .resampling import resample

    if audio.dtype == torch.int16:  # 16-bit PCM: x / 32768 like torchaudio.load's normalisation (exact)
        audio = ops.pcm16_to_float(audio)
    x = resample(audio, sr, fps * 1024, clip_n_in=lens.n_in if lens is not None else None)
    ch = feature_chain(x, fps * 1024, want_a2f=True, want_mir=True, sync_free=sync_free, lens=lens)
    return (finish_features(ch["a2f"], fps, deferred_checks=deferred_checks, lens=lens),
            patch_inputs_from_raw(ch["mir"], lens=lens))


def _count_replayed(n: int):
    """A graph replay launches the kernels of the n C-ABI calls it captured: they count as launches (`_cabi.launch_count`
    is what bench.py reports as `gpu_launches`)."""
    from . import _cabi

    with _cabi._count_lock:
        _cabi.launch_count += n


class FeatureGraph:
    """`features_and_patch_inputs` for a fixed (clips, samples, sr, fps) as ONE replayed CUDA graph.

    The feature path of a chunk of clips is ~300 launches whose arguments depend only on the shapes (the one
    data-dependent choice, each clip's tuned CQT basis, is a device gather with `sync_free=True`), so it is captured once
    and replayed: the host cost of a chunk drops from ~12 ms of Python + launch calls to one graph launch.  Inputs are
    copied (or staged from pinned host memory) into the static buffer `x`; outputs are static too and must be consumed or
    cloned before the next replay (`run` returns clones of the small ones)."""

    _cache = {}     # key -> FeatureGraph, insertion-ordered: least recently used first
    _seen = {}      # key -> how often the shape was asked for
    MAX_GRAPHS = 3  # each graph pins its private pool (~0.4 GB per clip): keep only the most recently used shapes
    _lock = threading.RLock()

    @staticmethod
    def _key(n_clips, n_samples, sr, fps, device):
        # a graph owns static input / output buffers: it belongs to the host thread that asked for it (two threads that
        # process the same shape concurrently get two graphs, never each other's buffers)
        return (int(n_clips), int(n_samples), float(sr), int(fps), device.index, threading.get_ident())

    @classmethod
    def get(cls, n_clips, n_samples, sr, fps, device):
        device = torch.device(device)
        key = cls._key(n_clips, n_samples, sr, fps, device)
        with cls._lock:
            g = cls._cache.pop(key, None)
            if g is None:
                g = cls(n_clips, n_samples, sr, fps, device)
                mine = [k for k in cls._cache if k[-1] == key[-1]]  # only this thread's graphs are ever evicted here:
                while len(mine) >= cls.MAX_GRAPHS:                  # another thread may be replaying its own
                    cls._cache.pop(mine.pop(0))  # drops the graph and its pool
            cls._cache[key] = g
        return g

    @classmethod
    def get_if_recurring(cls, n_clips, n_samples, sr, fps, device):
        """The graph for this shape from its SECOND request on (None the first time): a dataset of clips of many different
        lengths runs eagerly instead of capturing — and pinning memory for — a graph per length."""
        device = torch.device(device)
        key = cls._key(n_clips, n_samples, sr, fps, device)
        with cls._lock:
            n = cls._seen[key] = cls._seen.get(key, 0) + 1
            if len(cls._seen) > 4096:
                cls._seen.clear()
            cached = key in cls._cache
        return cls.get(n_clips, n_samples, sr, fps, device) if (n >= 2 or cached) else None

    def __init__(self, n_clips, n_samples, sr, fps, device):
        self.device, self.sr, self.fps = device, sr, fps
        with torch.cuda.device(device):
            self.x = torch.zeros(n_clips, n_samples, dtype=torch.float32, device=device)
            self.x[:, ::997] = 0.25  # not silent: the warm-up pass must not hit the reference's empty-peak-set error
            self.x[:, 1::1499] = -0.5
            cs = torch.cuda.Stream(device)
            cs.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(cs):  # eager warm-up: plan tables, function attributes, allocator pools
                features_and_patch_inputs(self.x, sr, fps, deferred_checks=[], sync_free=True)
            cs.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            checks = []
            from . import _cabi

            n0 = _cabi.launch_count
            with torch.cuda.graph(self.graph, stream=cs, capture_error_mode="thread_local"):
                self.feats, self.patch_feats = features_and_patch_inputs(self.x, sr, fps, deferred_checks=checks,
                                                                         sync_free=True)
            self.n_launches = _cabi.launch_count - n0  # C-ABI entry points captured: every replay launches them again
            self._checks = checks
            self.peak_count = checks[0].args[0] if checks else None  # static [B, 59] interior-peak counts

    def run(self, x=None, deferred_checks=None):
        """Replays on the current stream.  x: [clips, samples] CUDA tensor copied into the static input (None: the caller
        already filled `self.x`).  -> (features [B, T, 59], patch feature dict) as fresh tensors."""
        if x is not None:
            if x.dtype == torch.int16:
                ops.pcm16_to_float(x, out=self.x)  # 16-bit PCM staging: converted straight into the static input
            else:
                self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        _count_replayed(self.n_launches)
        feats = self.feats.clone()
        pf = {k: v.clone() for k, v in self.patch_feats.items()}
        if self.peak_count is not None:
            cnt = self.peak_count.clone()
            if deferred_checks is not None:
                deferred_checks.append(functools.partial(_check_peak_counts, cnt))
            else:
                _check_peak_counts(cnt)
        return feats, pf


class RaggedFeatureGraph:
    """`features_and_patch_inputs` for up to `n_clips` clips of DIFFERENT lengths (each <= `cap_samples`) as one replayed
    CUDA graph: the clips sit zero-padded in the static input, their lengths are device data (`ragged.ClipLengths`, fixed
    addresses), so one capture serves every combination of lengths — the per-track eager pass (~400 launches) a dataset
    of 1000 distinct durations would otherwise need is replaced by one graph launch per chunk."""

    _cache = {}
    MAX_GRAPHS = 10  # one per capacity bucket (~0.05 GB per clip-minute each)

    @classmethod
    def get(cls, n_clips, cap_samples, sr, fps, device):
        device = torch.device(device)
        key = (int(n_clips), int(cap_samples), float(sr), int(fps), device.index, threading.get_ident())
        with FeatureGraph._lock:
            g = cls._cache.pop(key, None)
            if g is None:
                mine = [k for k in cls._cache if k[-1] == key[-1]]
                while len(mine) >= cls.MAX_GRAPHS:
                    cls._cache.pop(mine.pop(0))
                g = cls(n_clips, cap_samples, sr, fps, device)
            cls._cache[key] = g
        return g

    def __init__(self, n_clips, cap_samples, sr, fps, device):
        from .ragged import ClipLengths

        self.device, self.sr, self.fps, self.B, self.cap = device, sr, fps, int(n_clips), int(cap_samples)
        with torch.cuda.device(device):
            self.x = torch.zeros(n_clips, cap_samples, dtype=torch.float32, device=device)
            self.x[:, ::997] = 0.25  # not silent: the warm-up pass must not hit the reference's empty-peak-set error
            self.x[:, 1::1499] = -0.5
            self.lens = ClipLengths(n_clips, device).update([cap_samples] * n_clips, sr, fps * 1024)
            cs = torch.cuda.Stream(device)
            cs.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(cs):
                features_and_patch_inputs(self.x, sr, fps, deferred_checks=[], sync_free=True, lens=self.lens)
            cs.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            checks = []
            from . import _cabi

            n0 = _cabi.launch_count
            with torch.cuda.graph(self.graph, stream=cs, capture_error_mode="thread_local"):
                self.feats, self.patch_feats = features_and_patch_inputs(self.x, sr, fps, deferred_checks=checks,
                                                                         sync_free=True, lens=self.lens)
            self.n_launches = _cabi.launch_count - n0
            self.peak_count = checks[0].args[0] if checks else None

    def run(self, clips, deferred_checks=None):
        """clips: up to n_clips 1-D waveforms (CUDA or pinned host), each <= cap_samples long.  Replays on the current
        stream -> (features [B, T_cap, 59], patch feature dict, frames per clip); rows past a clip's own frame count and
        clips beyond len(clips) are don't-care."""
        n = len(clips)
        lengths = [int(c.shape[-1]) for c in clips]
        pad = lengths + [lengths[-1]] * (self.B - n)  # unused slots repeat the last clip's length (and hold its audio)
        self.lens.update(pad, self.sr, self.fps * 1024)
        for b in range(self.B):
            c, L = clips[min(b, n - 1)], pad[b]
            self.x[b, :L].copy_(c, non_blocking=True)
            if L < self.cap:
                self.x[b, L:].zero_()
        self.graph.replay()
        _count_replayed(self.n_launches)
        feats = self.feats[:n].clone()
        pf = {k: v[:n].clone() for k, v in self.patch_feats.items()}
        if self.peak_count is not None:
            cnt = self.peak_count[:n].clone()
            if deferred_checks is not None:
                deferred_checks.append(functools.partial(_check_peak_counts, cnt))
            else:
                _check_peak_counts(cnt)
        return feats, pf, list(self.lens.host_T[:n])


def features_and_latents_ragged(clips, sr, fps, palette: torch.Tensor, segmentations, tempo=120.0, seed=42,
                                rng_device="cuda", chunk=8, bucket_seconds=30.0, on_clip=None, use_graphs=True,
                                stats=None):
    """The north-star unit for tracks of DIFFERENT durations (SURVEY 8d cfg 5): clips = list of 1-D waveforms at `sr`
    (CUDA, or pinned host: staged per chunk).  Tracks are processed longest first in chunks of `chunk`, each chunk
    zero-padded to its longest member rounded up to `bucket_seconds` — ONE ragged launch set (a replayed CUDA graph per
    capacity bucket) serves the chunk, results identical to per-track calls (tests/test_gpu_ragged.py).

    segmentations: dict (name, k) -> label tensor of at least the longest track's frame count (sliced per track), or a
    callable (track index, n_frames) -> dict.  on_clip(i, features [T_i, 59], latents [T_i, 18, 512]) is called per
    track while its latents are alive (1000 four-minute tracks of latents are 212 GB: they cannot all stay resident);
    without it the features and latents of every track are collected and returned as two lists (small sets only).
    stats: optional fp64 [3, 59] accumulator of the dataset statistic (count, sum, sum of squares over valid frames)."""
    from .random.patch import Patch, precapture_sampler_graphs, sampler_kernel_ready

    dev = palette.device
    if dev.type != "cuda":
        raise ops.SsarError("features_and_latents_ragged needs the palette (and the compute) on a CUDA device")
    n_tracks = len(clips)
    if n_tracks == 0:
        return [], []
    order = sorted(range(n_tracks), key=lambda i: (-int(clips[i].shape[-1]), i))
    chunk = max(1, int(chunk))
    bucket = max(1, int(round(bucket_seconds * sr)))
    stream = torch.cuda.current_stream(dev)
    fstream = _side_stream(dev, "features")
    seg0 = segmentations if isinstance(segmentations, dict) else segmentations(order[0], int(clips[order[0]].shape[-1]) * fps // sr)
    if str(rng_device).startswith("cuda"):
        with torch.cuda.device(dev):
            if not sampler_kernel_ready(dev, seg0):
                precapture_sampler_graphs(dev, seg0)
    with torch.cuda.device(dev), torch.cuda.stream(stream):
        palette_split = ops.SplitBasis(palette)
    feats_out, lat_out = [None] * n_tracks, [None] * n_tracks
    checks = []
    W, L = palette.shape[1], palette.shape[2]

    class _Prefix(dict):
        """segmentations[(name, k)][:T], sliced when (if) a sub-patch asks for that key."""

        def __init__(self, full, T):
            super().__init__()
            self._full, self._T = full, T

        def __missing__(self, key):
            v = self[key] = self._full[key][: self._T]
            return v

        def keys(self):
            return self._full.keys()

        def __iter__(self):
            return iter(self._full)

        def __len__(self):
            return len(self._full)

    # the config draws of every track (patch.py:41-66 needs no feature values) on a helper thread and a high-priority
    # side stream, 64 seeds per sampler launch + read-back, in processing order: the chunk loop below waits only for
    # the batch of 64 its tracks are in, and otherwise never on the GPU, so the host queues ahead of the kernels
    patch_of, sampled, sample_error = {}, {}, []
    batches = [order[c0 : c0 + 64] for c0 in range(0, n_tracks, 64)]
    for part in batches:
        ev = threading.Event()
        for i in part:
            sampled[i] = ev

    def sample_job():
        try:
            with torch.cuda.device(dev), torch.cuda.stream(_side_stream(dev)):
                for part in batches:
                    for i, pt in zip(part, Patch.sample_batch([seed + i for i in part], seg0, tempo, fps=fps, device=rng_device)):
                        patch_of[i] = pt
                    sampled[part[0]].set()
        except BaseException as e:  # surfaced by the calling thread
            sample_error.append(e)
            for ev in sampled.values():
                ev.set()

    sampler = _pool(3).submit(sample_job)

    def render(item):
        idxs, f, pf, Ts, done, frames_dev = item
        with torch.cuda.device(dev), torch.cuda.stream(stream):
            stream.wait_event(done)
            if stats is not None:
                ops.feature_stats(f, stats, clip_frames=frames_dev)  # valid frames of the whole chunk in one launch
            lat = torch.empty(len(idxs), max(Ts), W, L, dtype=torch.float32, device=dev)
            for j, i in enumerate(idxs):
                T = Ts[j]
                sampled[i].wait()
                if sample_error:
                    raise sample_error[0]
                patch = patch_of.pop(i)
                patch.segmentations = _Prefix(segmentations, T) if isinstance(segmentations, dict) else segmentations(i, T)
                patch.bind({k: v[j, :T] for k, v in pf.items()})
                patch.forward(palette, downscale_factor=4, latents_out=lat[j, :T], palette_split=palette_split)
                if on_clip is not None:
                    on_clip(i, f[j, :T], lat[j, :T])
                else:
                    feats_out[i], lat_out[i] = f[j, :T], lat[j, :T]

    pending = None
    fstream.wait_stream(stream)
    for c0 in range(0, n_tracks, chunk):
        idxs = order[c0 : c0 + chunk]
        cap = -(-int(clips[idxs[0]].shape[-1]) // bucket) * bucket  # longest first: the chunk's first track
        with torch.cuda.device(dev), torch.cuda.stream(fstream):
            if use_graphs:
                g = RaggedFeatureGraph.get(chunk, cap, sr, fps, dev)
                f, pf, Ts = g.run([clips[i] for i in idxs], deferred_checks=checks)
                frames_dev = g.lens.T[: len(idxs)].clone()
            else:
                from .ragged import ClipLengths

                lens = ClipLengths(len(idxs), dev).update([int(clips[i].shape[-1]) for i in idxs], sr, fps * 1024)
                x = torch.zeros(len(idxs), cap, dtype=torch.float32, device=dev)
                for b, i in enumerate(idxs):
                    x[b, : clips[i].shape[-1]].copy_(clips[i], non_blocking=True)
                f, pf = features_and_patch_inputs(x, sr, fps, deferred_checks=checks, lens=lens)
                Ts, frames_dev = list(lens.host_T), lens.T
            done = torch.cuda.Event()
            done.record(fstream)
            for t in (f, frames_dev, *pf.values()):
                t.record_stream(stream)
        if pending is not None:
            render(pending)  # chunk k's patches are queued while chunk k + 1's feature kernels run
        pending = (idxs, f, pf, Ts, done, frames_dev)
    render(pending)
    sampler.result()
    fstream.wait_stream(stream)
    for check in checks:
        check()
    return feats_out, lat_out


USE_FEATURE_GRAPHS = True  # batched pipeline: replay the feature path of each chunk as a CUDA graph

_POOLS = {}
_SIDE_STREAMS = {}


def _side_stream(device, purpose="sampler"):
    key = (torch.device(device).index, threading.get_ident(), purpose)
    st = _SIDE_STREAMS.get(key)
    if st is None:
        # sampler: high priority, so its tiny generator kernels jump the queue of feature kernels
        st = _SIDE_STREAMS[key] = torch.cuda.Stream(device, priority=-1 if purpose == "sampler" else 0)
    return st


_STAGING = {}


def _staging(device, shape, dtype):
    """Two device buffers of `shape` for the H2D ping-pong of host-resident audio, kept per (device, thread, shape);
    `free[i]` is the event after which buffer i may be overwritten."""
    key = (torch.device(device).index, threading.get_ident(), tuple(shape), dtype)
    st = _STAGING.get(key)
    if st is None:
        with FeatureGraph._lock:
            for k in [k for k in _STAGING if k[1] == key[1] and len(k) == 4][:-1]:  # this thread's older shapes only; their copies were
                _STAGING.pop(k)                                     # consumed by the calls that issued them
            st = _STAGING[key] = {"buf": [torch.empty(shape, dtype=dtype, device=device) for _ in range(2)],
                                  "free": [None, None]}
    return st


class StagedBatch:
    """A host batch on its way to (or already in) one of two persistent device buffers: what `stage_batch` returns and
    `features_and_latents` accepts in place of the batch."""

    def __init__(self, tensor, ready, state, slot):
        self.tensor, self.ready, self._state, self._slot = tensor, ready, state, slot

    @property
    def shape(self):
        return self.tensor.shape

    def consumed(self, stream):
        """Called by the consumer once every kernel that reads the batch is queued on `stream`."""
        ev = torch.cuda.Event()
        ev.record(stream)
        self._state["free"][self._slot] = ev


def stage_batch(audio: torch.Tensor, device) -> StagedBatch:
    """Start the H2D copy of a (pinned) host batch [B, N] (fp32 waveforms or int16 PCM) on the copy stream and return at
    once.  A dataset job calls it for batch k + 1 before `features_and_latents(batch k)`: the PCIe transfer of a batch then
    overlaps the kernels of the previous one instead of standing in front of its own (two persistent device buffers per
    (device, thread, shape); a buffer is only overwritten after the call that consumed it has finished reading it).
    Input staging only — there is no CPU compute path."""
    device = torch.device(device)
    if audio.is_cuda:
        raise ValueError("stage_batch expects a host tensor")
    key = (device.index, threading.get_ident(), tuple(audio.shape), audio.dtype, "batch")
    st = _STAGING.get(key)
    if st is None:
        with FeatureGraph._lock:
            for k in [k for k in _STAGING if k[1] == key[1] and k[-1] == "batch"]:
                _STAGING.pop(k)
            st = _STAGING[key] = {"buf": [torch.empty(audio.shape, dtype=audio.dtype, device=device) for _ in range(2)],
                                  "free": [None, None], "next": 0}
    slot = st["next"]
    st["next"] = 1 - slot
    copy_stream = _side_stream(device, "h2d")
    with torch.cuda.device(device), torch.cuda.stream(copy_stream):
        if st["free"][slot] is not None:
            copy_stream.wait_event(st["free"][slot])
        st["buf"][slot].copy_(audio, non_blocking=True)
        ready = torch.cuda.Event()
        ready.record(copy_stream)
    return StagedBatch(st["buf"][slot], ready, st, slot)


def _pool(workers: int):
    from concurrent.futures import ThreadPoolExecutor

    p = _POOLS.get(workers)
    if p is None:
        p = _POOLS[workers] = ThreadPoolExecutor(max_workers=workers, thread_name_prefix="ssar-patch")
    return p


def default_patch_workers(rng_device="cpu") -> int:
    """Host threads for the per-clip Patch construction.  With a CPU generator the large noise-basis draws dominate and
    release the GIL, so threads help: the process's share of the host cores (one process per GPU under torchrun:
    LOCAL_WORLD_SIZE processes share the box), at most 8.  With a CUDA generator every draw is a ~5 us launch made under
    the GIL and threads only add contention (measured: 286 ms vs 415 ms per 32-clip step): 1."""
    if str(rng_device).startswith("cuda"):
        return 1
    import os

    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    share = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1"))))
    return max(1, min(8, cores // share))


def features_and_latents(audio, sr, fps, palette: torch.Tensor, segmentations, tempo=120.0, seed=42,
                         rng_device="cuda", workers=None, chunk=8, serial=False):
    """The north-star unit for a batch of clips: waveforms [B, N] -> (features [B, T, 59], latents [B, T, 18, 512]).
    audio: a device or (pinned) host tensor, or the `StagedBatch` handle of `stage_batch` (a host batch whose copy was started
    ahead of time); fp32 waveforms, or int16 PCM (converted on the device as x / 32768, exactly torchaudio.load's normalisation of
    the 16-bit files the reference reads — a host-resident batch then crosses PCIe at half the bytes).
    Tempo and segmentations are inputs (random/mir.py:28-42 is CPU librosa / scipy work outside the path).
    rng_device: where the Patch generator lives ("cpu" reproduces a CPU-seeded reference run bit for bit in its draws;
    the reference's own default is the compute device, random/__init__.py:1-3).

    Host pipeline (the step is bound by host launch work, not by the kernels): three threads on the caller's stream +
    one side stream, all in stream order so no extra events are needed —
      * feature thread: queues `audio2features_batch` for `chunk` clips at a time (it blocks once per chunk on the clip
        tunings; torch releases the GIL while it waits);
      * sampler thread(s): `Patch(...)` for every clip on a high-priority side stream (patch.py:41-66 needs no feature
        values; its read-backs then never queue behind the feature kernels);
      * calling thread: as soon as a chunk's features are queued, binds them to that chunk's patches and renders
        (one fused launch per latent sub-patch, one mixing GEMM per clip, the noise-basis draws).
    workers: sampler threads (CPU generators only: their large basis draws release the GIL; CUDA generators use one).
    serial: queue everything from the calling thread on the current stream (no helper threads, no side streams) — the
    mode bench.py uses for its per-kernel CUDA-event pass, where concurrent streams would stretch the event intervals."""
    import queue
    import threading

    from .random.patch import Patch, precapture_sampler_graphs, sampler_kernel_ready

    # `audio` may live in (pinned) HOST memory: the feature thread then stages it chunk by chunk on a copy stream, one
    # chunk ahead of the kernels (input staging only — there is no CPU compute path)
    staged = audio if isinstance(audio, StagedBatch) else None
    if staged is not None:  # `stage_batch`: the copy was started earlier; this call only waits for it in stream order
        audio = staged.tensor
    dev = palette.device if not audio.is_cuda else audio.device
    if dev.type != "cuda":
        raise ops.SsarError("features_and_latents needs the palette (and the compute) on a CUDA device")
    stream = torch.cuda.current_stream(dev)
    if staged is not None:
        stream.wait_event(staged.ready)
    if str(rng_device).startswith("cuda"):
        with torch.cuda.device(dev):
            if not sampler_kernel_ready(dev, segmentations):  # fallback sampler: one CUDA graph per sub-patch count
                precapture_sampler_graphs(dev, segmentations)  # no-op once captured
    workers = default_patch_workers(rng_device) if workers is None else int(workers)
    with torch.cuda.device(dev), torch.cuda.stream(stream):
        palette_split = ops.SplitBasis(palette)  # one fp16 (hi, lo) split of the palette for every clip's render
    B = audio.shape[0]
    chunk = max(1, min(int(chunk), B))
    bounds = [(c0, min(c0 + chunk, B)) for c0 in range(0, B, chunk)]
    if not audio.is_cuda and not serial and chunk >= 4 and chunk % 2 == 0 and B % chunk == 0 and B >= 3 * chunk:
        # host-resident audio: tapered schedule (half chunks first and last, two graph shapes) — nothing can run until
        # the first chunk has crossed PCIe, so it is a short one, and only half a chunk of patches is left to render after
        # the last feature kernels.  Device-resident audio keeps whole chunks: the extra chunk costs host time per step
        # that 8 ranks sharing one host do not have (N = 8: 52.1 ms tapered against 47.5 ms)
        h = chunk // 2
        cuts = [0, h] + list(range(h + chunk, B - h, chunk)) + [B - h, B]
        cuts = sorted(set(cuts))
        bounds = list(zip(cuts[:-1], cuts[1:]))
    checks = []
    fgraphs = None
    if USE_FEATURE_GRAPHS and not serial:
        # captured here, before the helper threads start launching (one graph per distinct chunk size)
        with torch.cuda.device(dev), torch.cuda.stream(stream):
            fgraphs = {}
            for n in sorted({c1 - c0 for c0, c1 in bounds}):
                g = FeatureGraph.get_if_recurring(n, audio.shape[1], sr, fps, dev)
                if g is not None:
                    fgraphs[n] = g
    if serial:
        feats_chunks, latents, noises = [], None, [None] * B
        with torch.cuda.device(dev), torch.cuda.stream(stream):
            for c0, c1 in bounds:
                f, pf = features_and_patch_inputs(audio[c0:c1].to(dev, non_blocking=True), sr, fps, deferred_checks=checks)
                feats_chunks.append(f)
                if latents is None:
                    latents = torch.empty(B, f.shape[1], palette.shape[1], palette.shape[2], dtype=torch.float32, device=dev)
                chunk_patches = Patch.sample_batch([seed + b for b in range(c0, c1)], segmentations, tempo, fps=fps,
                                                   device=rng_device)
                for b in range(c0, c1):
                    patch = chunk_patches[b - c0]
                    patch.bind({k: v[b - c0] for k, v in pf.items()})
                    noises[b] = patch.forward(palette, downscale_factor=4, latents_out=latents[b], palette_split=palette_split)[1]
            feats = feats_chunks[0] if len(feats_chunks) == 1 else torch.cat(feats_chunks, dim=0)
        if staged is not None:
            staged.consumed(stream)
        for check in checks:
            check()
        return feats, latents, noises
    pool = _pool(max(workers, 1) + 2)
    ready_feats = queue.Queue()

    # features run on their own stream: the patch rendering of chunk k (a chain of small kernels on the caller's stream)
    # then overlaps the feature kernels of chunks k + 1.. on the GPU instead of queueing behind them
    fstream = _side_stream(dev, "features")
    fstream.wait_stream(stream)

    def features_job():
        try:
            with torch.cuda.device(dev), torch.cuda.stream(fstream):
                staged = None
                if not audio.is_cuda:
                    copy_stream = _side_stream(dev, "h2d")
                    # two persistent device buffers (ping-pong): a fresh allocation per chunk on the copy stream makes
                    # the caching allocator cudaMalloc (a device-wide sync) whenever its blocks are still held by the
                    # feature stream, which stalled the copy / compute overlap
                    ring = _staging(dev, (max(c1 - c0 for c0, c1 in bounds),) + tuple(audio.shape[1:]), audio.dtype)

                    def stage(ci, c0, c1):
                        slot = ci % 2
                        with torch.cuda.stream(copy_stream):
                            if ring["free"][slot] is not None:
                                copy_stream.wait_event(ring["free"][slot])  # its previous chunk has been consumed
                            xd = ring["buf"][slot][: c1 - c0]
                            xd.copy_(audio[c0:c1], non_blocking=True)
                            ev = torch.cuda.Event()
                            ev.record(copy_stream)
                        return xd, ev

                    staged = stage(0, *bounds[0])
                for ci, (c0, c1) in enumerate(bounds):
                    if staged is not None:
                        xd, ev = staged
                        staged = stage(ci + 1, *bounds[ci + 1]) if ci + 1 < len(bounds) else None  # one chunk ahead
                        fstream.wait_event(ev)
                    else:
                        xd = audio[c0:c1]
                    if fgraphs is not None and (c1 - c0) in fgraphs:
                        f, pf = fgraphs[c1 - c0].run(xd, deferred_checks=checks)
                    else:
                        f, pf = features_and_patch_inputs(xd, sr, fps, deferred_checks=checks)
                    done = torch.cuda.Event()
                    done.record(fstream)
                    if not audio.is_cuda:
                        ring["free"][ci % 2] = done
                    for t in (f, *pf.values()):
                        t.record_stream(stream)  # allocated on the feature stream, consumed on the caller's
                    ready_feats.put((c0, c1, f, pf, done))
        except BaseException as e:  # surfaced by the calling thread
            ready_feats.put(e)

    patches = [None] * B
    sampled = [threading.Event() for _ in range(B)]
    sample_error = []

    def sample_job(indices):
        try:
            with torch.cuda.device(dev), torch.cuda.stream(_side_stream(dev)):
                if str(rng_device).startswith("cuda"):
                    # one sampler kernel + one read-back per chunk of clips (Patch.sample_batch)
                    for c0 in range(0, len(indices), chunk):
                        part = indices[c0 : c0 + chunk]
                        for b, pt in zip(part, Patch.sample_batch([seed + b for b in part], segmentations, tempo, fps=fps,
                                                                  device=rng_device)):
                            patches[b] = pt
                            sampled[b].set()
                    return
                for b in indices:
                    patches[b] = Patch(None, segmentations, tempo, fps=fps, seed=seed + b, device=rng_device)
                    sampled[b].set()
        except BaseException as e:
            sample_error.append(e)
            for ev in sampled:
                ev.set()

    futs = [pool.submit(features_job)]
    n_samplers = 1 if str(rng_device).startswith("cuda") else max(1, workers)
    for w in range(n_samplers):
        futs.append(pool.submit(sample_job, list(range(w, B, n_samplers))))

    feats_chunks, latents, noises = [], None, [None] * B
    with torch.cuda.device(dev), torch.cuda.stream(stream):
        for _ in bounds:
            item = ready_feats.get()
            if isinstance(item, BaseException):
                raise item
            c0, c1, f, pf, done = item
            stream.wait_event(done)
            feats_chunks.append(f)
            if latents is None:
                latents = torch.empty(B, f.shape[1], palette.shape[1], palette.shape[2], dtype=torch.float32, device=dev)
            for b in range(c0, c1):
                sampled[b].wait()
                if sample_error:
                    raise sample_error[0]
                patches[b].bind({k: v[b - c0] for k, v in pf.items()})
                noises[b] = patches[b].forward(palette, downscale_factor=4, latents_out=latents[b],
                                               palette_split=palette_split)[1]
        feats = feats_chunks[0] if len(feats_chunks) == 1 else torch.cat(feats_chunks, dim=0)
    for fu in futs:
        fu.result()
    if staged is not None:
        staged.consumed(stream)  # (the caller's stream has waited for every chunk's feature kernels)
    for check in checks:  # error behaviour of the reference (empty peak set), read back once everything is queued
        check()
    return feats, latents, noises
