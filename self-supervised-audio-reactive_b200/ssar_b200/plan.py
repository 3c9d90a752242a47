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
    lanes: object = None  # LaneBanks (real banks only): the lane-interleaved form the frames_fft epilogue streams


LANE_PAD = 64  # zeroed floats the frames_fft kernel keeps behind the n_cols spectrum bins of its power tile


@dataclass
class LaneBanks:
    """Rows in passes of 32 (row 32 p + lane): lane reads 4 k4 + c of its row's 16-byte-aligned window
    [start4, start4 + 4 count4[p]) at weights[((offset4[p] + k4) * 32 + lane) * 4 + c]; zero weights pad every row to
    its pass's longest, so a warp streams coalesced 512-byte rows of weights whatever the row widths."""

    start4: torch.Tensor   # int32 [n_pass * 32]
    count4: torch.Tensor   # int32 [n_pass]
    offset4: torch.Tensor  # int32 [n_pass]
    weights: torch.Tensor  # fp32 [sum(count4) * 32 * 4]


def lane_banks(starts, counts, rows_cpu, n_cols, device):
    """rows_cpu[r]: the packed fp32 weights of row r.  None when a padded window would leave the zeroed tile tail."""
    n_rows = len(starts)
    n_pass = (n_rows + 31) // 32
    start4 = [0] * (n_pass * 32)
    count4, offset4, blocks = [], [], []
    off = 0
    for p_ in range(n_pass):
        rows = range(32 * p_, min(32 * p_ + 32, n_rows))
        n4 = max([((starts[r] & 3) + counts[r] + 3) // 4 for r in rows] + [0])
        blk = torch.zeros(n4, 32, 4, dtype=torch.float32)
        for r in rows:
            s4 = starts[r] & ~3
            if s4 + 4 * n4 > n_cols + LANE_PAD:
                return None
            start4[r] = s4
            lead = starts[r] - s4
            flat = torch.zeros(4 * n4, dtype=torch.float32)
            flat[lead:lead + counts[r]] = rows_cpu[r]
            blk[:, r - 32 * p_, :] = flat.view(n4, 4)
        count4.append(n4)
        offset4.append(off)
        off += n4
        blocks.append(blk.reshape(-1))
    i32 = lambda v: torch.tensor(v, dtype=torch.int32, device=device)
    w = torch.cat(blocks) if blocks else torch.zeros(0)
    return LaneBanks(i32(start4), i32(count4), i32(offset4), w.contiguous().to(device))


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
    lanes = None if is_c else lane_banks(starts, counts, chunks, n_cols, device)
    return Banded(
        n_rows, n_cols, i32(starts), i32(counts), i32(offsets), w.contiguous().to(device),
        (torch.view_as_real(dense_cpu.to(torch.complex64)) if is_c else dense_cpu.to(torch.float32)).contiguous().to(device),
        lanes,
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


@functools.lru_cache(maxsize=None)
def gaussian_taps(sigma: float, radius: int, device) -> torch.Tensor:
    """features/processing.py:21-24: exp(-0.5/sigma^2 * k^2) / sum over k in [-radius, radius], fp32 op order kept."""
    k = torch.arange(-radius, radius + 1, dtype=torch.float32)
    k = torch.exp(-0.5 / sigma ** 2 * k ** 2)
    k = k / k.sum()
    return k.to(_dev(device))


@functools.lru_cache(maxsize=64)
def gaussian_taps_causal(sigma: float, radius: int, causal, device) -> torch.Tensor:
    """maua audioreactive/signal.py:124-130: the same Gaussian with the taps right of the centre scaled by `causal`
    (by 0 when `causal` is not a float) BEFORE the normalisation; `causal=None` = the plain kernel."""
    k = torch.arange(-radius, radius + 1, dtype=torch.float32)
    k = torch.exp(-0.5 / sigma ** 2 * k ** 2)
    if causal is not None:
        k[radius + 1:] *= 0 if not isinstance(causal, float) else causal
    k = k / k.sum()
    return k.to(_dev(device))


@dataclass(frozen=True)
class ContrastBands:
    start: torch.Tensor  # int32 [n_bands + 1]
    count: torch.Tensor
    q: torch.Tensor
    n: int
    max_count: int


@functools.lru_cache(maxsize=None)
def contrast_bands(sr: float, n_fft: int, fmin: float, n_bands: int, quantile: float, device) -> ContrastBands:
    """Bin ranges and extreme counts of features/audio.py:86-113, evaluated once with the same fp32 comparisons."""
    freq = torch.linspace(0, float(sr) / 2, int(1 + n_fft // 2))
    octa = torch.zeros(n_bands + 2)
    octa[1:] = fmin * (2.0 ** torch.arange(0, n_bands + 1))
    starts, counts, qs = [], [], []
    for k, (f_low, f_high) in enumerate(zip(octa[:-1], octa[1:])):
        band = torch.logical_and(freq >= f_low, freq <= f_high)
        idx = band.nonzero().flatten()
        if k > 0:
            band[idx[0] - 1] = True
        if k == n_bands:
            band[idx[-1] + 1:] = True
        bins = band.nonzero().flatten()
        assert int(bins[-1] - bins[0]) + 1 == len(bins), "contrast band is not a contiguous bin range"
        n_used = len(bins) - (1 if k < n_bands else 0)
        q = int(torch.maximum(torch.round(torch.tensor(quantile) * band.sum()), torch.ones(())))
        if q > n_used or q > 32:
            raise ValueError(f"spectral_contrast: band {k} needs {q} extremes of {n_used} bins (limit 32)")
        starts.append(int(bins[0]))
        counts.append(n_used)
        qs.append(q)
    dev = _dev(device)
    i32 = lambda v: torch.tensor(v, dtype=torch.int32, device=dev)
    return ContrastBands(i32(starts), i32(counts), i32(qs), n_bands + 1, max(counts))


# ------------------------------------------------------------------------------------------------
# Constant-Q basis (rosa/constantq.py:118-164, 215-269) for every tuning the estimator can return.
# estimate_tuning quantises to linspace(-0.5, 0.5, 101)[argmax of a 100-bin histogram] (rosa/pitch.py:115-120), so
# only 100 distinct bases exist; one basis serves all octaves (sr and fmin halve together, exact in fp32).
# ------------------------------------------------------------------------------------------------
CQT_BINS_PER_OCTAVE = 36
CQT_N_OCTAVES = 7
CQT_N_FFT = 1024
C1_HZ = 440.0 * 2.0 ** ((24 - 69) / 12.0)  # librosa.note_to_hz("C1"), rosa/convert.py:129-130


def cqt_tuning_values() -> torch.Tensor:
    return torch.linspace(-0.5, 0.5, 101)


def cqt_fmin_tuned(tuning: torch.Tensor) -> torch.Tensor:
    """rosa/constantq.py:48-58: C1 * 2^(tuning / bins_per_octave), fp32."""
    return torch.tensor(C1_HZ).float() * 2.0 ** (tuning / CQT_BINS_PER_OCTAVE)


def cqt_octave_basis_cpu(sr: float, tuning: torch.Tensor, sparsity: float = 0.01):
    """Top-octave FFT basis [36, 513] complex64 (dense, sparsified entries zeroed) + all 252 filter lengths."""
    bpo, n_bins = CQT_BINS_PER_OCTAVE, CQT_BINS_PER_OCTAVE * CQT_N_OCTAVES
    alpha = 2.0 ** (1.0 / bpo) - 1.0
    Q = 1.0 / alpha
    fmin = cqt_fmin_tuned(tuning)
    ratios = 2.0 ** (torch.arange(n_bins, dtype=torch.float) / bpo)
    all_freqs = 1.0 * fmin * ratios                      # cqt_frequencies (tuning correction 2^0 = 1)
    all_lengths = Q * sr / (fmin * ratios + 0.0 / alpha)  # constant_q_lengths over all bins (the final 1/sqrt(len))
    fmin_t = torch.min(all_freqs[-bpo:])
    freqs = fmin_t * ratios[:bpo]
    lengths = Q * sr / (freqs + 0.0 / alpha)
    n_fft = int(2.0 ** torch.ceil(torch.log2(lengths.max())))
    rows = []
    for ilen, freq in zip(lengths, freqs):
        half = torch.div(ilen, 2, rounding_mode="floor")
        phase = torch.arange(-half, half, dtype=torch.float) * 1j * 2 * torch.pi * freq / sr
        sig = torch.exp(phase) * torch.hann_window(len(phase))
        sig = sig / sig.norm(p=1, dim=0)
        lpad = (n_fft - len(sig)) // 2
        rows.append(torch.nn.functional.pad(sig, (lpad, n_fft - len(sig) - lpad)))
    basis = torch.stack(rows) * (lengths[:, None] / float(n_fft))
    fb = torch.fft.fft(basis, n=n_fft, dim=1)[:, : n_fft // 2 + 1]
    # sparsify_rows: drop the smallest entries holding < `sparsity` of each row's L1 mass
    mags = fb.abs()
    srt = torch.sort(mags, dim=1).values
    cum = torch.cumsum(srt / mags.sum(1, keepdim=True), dim=1)
    thr_idx = torch.argmin((cum < sparsity).to(torch.uint8), dim=1)
    thr = srt.gather(1, thr_idx[:, None])
    fb = torch.where(mags >= thr, fb, torch.zeros((), dtype=fb.dtype))
    return fb, n_fft, all_lengths


def cqt_banded_cpu(sr: float, tuning_index: int):
    """Banded form of one tuning's top-octave basis: (start[36], count[36], weights [nnz, 2] fp32, lengths[252])."""
    fb, n_fft, lengths = cqt_octave_basis_cpu(float(sr), cqt_tuning_values()[tuning_index])
    assert n_fft == CQT_N_FFT, f"CQT filters need n_fft={n_fft}; only {CQT_N_FFT} is built (sr={sr})"
    nz = fb.abs() > 0
    starts, counts, chunks = [], [], []
    for r in range(fb.shape[0]):
        idx = nz[r].nonzero().flatten()
        s, c = int(idx[0]), int(idx[-1]) - int(idx[0]) + 1
        starts.append(s)
        counts.append(c)
        chunks.append(torch.view_as_real(fb[r, s : s + c].contiguous()))
    return starts, counts, torch.cat(chunks), lengths


class CqtTables:
    """Device tables of the tuned CQT bases, grown on demand: slot s holds the basis of tuning index `slot_tuning[s]`."""

    def __init__(self, sr: float, device):
        self.sr, self.device = float(sr), _dev(device)
        self.slot_of = {}
        self._starts, self._counts, self._offsets, self._weights, self._rscale = [], [], [], [], []
        self._nnz = 0
        self.start = self.count = self.offset = self.weights = self.row_scale = None

    def slots_for(self, tuning_idx_host):
        """tuning indices (host ints, one per clip) -> int32 device tensor of slots; uploads any new basis."""
        grew = False
        for ti in sorted(set(int(v) for v in tuning_idx_host)):
            if ti in self.slot_of:
                continue
            st, cn, w, lengths = cqt_banded_cpu(self.sr, ti)
            self.slot_of[ti] = len(self.slot_of)
            self._starts += st
            self._counts += cn
            off = self._nnz
            for c in cn:
                self._offsets.append(off)
                off += c
            self._nnz = off
            self._weights.append(w)
            self._rscale.append(1.0 / torch.sqrt(lengths))  # V /= sqrt(lengths) (rosa/constantq.py:113)
            grew = True
        if grew:
            i32 = lambda v: torch.tensor(v, dtype=torch.int32, device=self.device)
            self.start, self.count, self.offset = i32(self._starts), i32(self._counts), i32(self._offsets)
            self.weights = torch.cat(self._weights).contiguous().to(self.device)
            self.row_scale = torch.stack(self._rscale).contiguous().to(self.device)  # [n_slots, 252]
        return torch.tensor([self.slot_of[int(v)] for v in tuning_idx_host], dtype=torch.int32, device=self.device)

    def slot_map(self):
        """int32 device table tuning index -> slot with EVERY tuning's basis resident (101 x 36 rows, ~0.5 MB): lets a
        batch pipeline pick each clip's basis on the device (`slot_map[tuning_idx]`) with no read-back — the form a
        CUDA-graph capture of the feature path needs."""
        n = len(cqt_tuning_values())
        if getattr(self, "_slot_map", None) is None or len(self.slot_of) < n:
            self.slots_for(range(n))
            self._slot_map = torch.tensor([self.slot_of[i] for i in range(n)], dtype=torch.int32, device=self.device)
        return self._slot_map


@functools.lru_cache(maxsize=None)
def cqt_tables(sr: float, device) -> CqtTables:
    return CqtTables(sr, device)


class CqtTcTables:
    """Operands of `ssar_cqt_tc`: the TIME-domain form of every tuning's basis, b[r, n] = sum_k fft_basis[r, k]
    exp(-2 pi i k n / 1024) (the inverse of what rosa/constantq.py:118-144 does to the filters, applied to the SPARSIFIED
    FFT basis so the contraction equals fft_basis @ rfft(frame) exactly), computed in fp64 from the fp32 basis, scaled by
    2^6 into fp16's normal range and split into (hi, lo) fp16 panels [101 * 80, 1024]: rows 2r / 2r + 1 of slot s hold
    Re / Im of b_s[r, :], rows 72..79 are zero.  Slot = tuning index (all 101 resident, 33 MB)."""

    ROWS = 80
    SCALE = 64.0

    def __init__(self, sr: float, device):
        n = len(cqt_tuning_values())
        panel = torch.zeros(n, self.ROWS, CQT_N_FFT, dtype=torch.float64)
        rscale = []
        for ti in range(n):
            fb, n_fft, lengths = cqt_octave_basis_cpu(float(sr), cqt_tuning_values()[ti])
            assert n_fft == CQT_N_FFT, f"CQT filters need n_fft={n_fft}; only {CQT_N_FFT} is built (sr={sr})"
            full = torch.zeros(fb.shape[0], n_fft, dtype=torch.complex128)
            full[:, : fb.shape[1]] = fb.to(torch.complex128)
            b = torch.fft.fft(full, dim=1)  # [36, 1024]
            panel[ti, 0 : 2 * fb.shape[0] : 2] = b.real
            panel[ti, 1 : 2 * fb.shape[0] : 2] = b.imag
            rscale.append(1.0 / torch.sqrt(lengths))  # V /= sqrt(lengths) (rosa/constantq.py:113)
        panel = (panel * self.SCALE).reshape(n * self.ROWS, CQT_N_FFT)
        hi = panel.to(torch.float16)
        lo = (panel - hi.double()).to(torch.float16)
        dev = _dev(device)
        self.n_slots = n
        self.basis_hi, self.basis_lo = hi.contiguous().to(dev), lo.contiguous().to(dev)
        self.row_scale = torch.stack(rscale).float().contiguous().to(dev)  # [101, 252]


@functools.lru_cache(maxsize=None)
def cqt_tc_tables(sr: float, device) -> CqtTcTables:
    return CqtTcTables(sr, device)


@functools.lru_cache(maxsize=None)
def tuning_edges(device) -> torch.Tensor:
    return cqt_tuning_values().to(_dev(device))


def piptrack_bin_range(sr: float, n_fft: int, fmin: float = 150.0, fmax: float = 4000.0):
    """rosa/pitch.py:44-66: bins with fmin <= fft_freqs < fmax (fp32 linspace comparison, contiguous)."""
    fmax = min(fmax, float(sr) / 2)
    freqs = torch.linspace(0, float(sr) / 2, int(1 + n_fft // 2))
    m = ((max(fmin, 0) <= freqs) & (freqs < fmax)).nonzero().flatten()
    return int(m[0]), int(m[-1]) + 1


def chroma_map_cpu(n_input: int, bins_per_octave: int = 36, n_chroma: int = 12) -> torch.Tensor:
    """rosa/convert.py:69-117 cq_to_chroma with fmin = C1, base_c=True, window=None as a class index per CQT bin.
    C1 is MIDI 24 -> roll 0.  Bin j of every octave belongs to chroma ((j + n_merge // 2) // n_merge) % n_chroma."""
    n_merge = bins_per_octave // n_chroma
    assert n_merge * n_chroma == bins_per_octave
    midi_0 = (12 * (math.log2(C1_HZ) - math.log2(440.0)) + 69) % 12
    roll = int(round(midi_0 * (n_chroma / 12.0))) % n_chroma
    j = torch.arange(n_input) % bins_per_octave
    return ((((j + n_merge // 2) // n_merge) + roll) % n_chroma).to(torch.int32)


@functools.lru_cache(maxsize=None)
def chroma_map(n_input: int, bins_per_octave: int, n_chroma: int, device) -> torch.Tensor:
    return chroma_map_cpu(n_input, bins_per_octave, n_chroma).to(_dev(device))


def natural_cubic_spline_np(t: np.ndarray, y: np.ndarray):
    """Natural cubic spline through (t, y) in the (a, b, two_c, three_d) parameterisation of torchcubicspline
    (published algorithm; the package itself is un-vendored and unpinned): value = a + b f + (two_c / 2) f^2 +
    (three_d / 3) f^3 upstream — the reference's own spline_eval uses two_c and three_d directly (rosa/spectral.py:200)."""
    t, y = t.astype(np.float64), y.astype(np.float64)
    n = len(t)
    dt = np.diff(t)
    dx = np.diff(y)
    A = np.zeros((n, n))
    rhs = np.zeros(n)
    for i in range(n):
        if i > 0:
            A[i, i] += 2.0 / dt[i - 1]
            A[i, i - 1] = 1.0 / dt[i - 1]
            rhs[i] += 3.0 * dx[i - 1] / dt[i - 1] ** 2
        if i < n - 1:
            A[i, i] += 2.0 / dt[i]
            A[i, i + 1] = 1.0 / dt[i]
            rhs[i] += 3.0 * dx[i] / dt[i] ** 2
    k = np.linalg.solve(A, rhs)
    a = y[:-1]
    b = k[:-1]
    two_c = (6.0 * dx / dt - 4.0 * k[:-1] - 2.0 * k[1:]) / dt
    three_d = (-6.0 * dx / dt + 3.0 * (k[:-1] + k[1:])) / dt ** 2
    return a, b, two_c, three_d


def cens_quantiser_knots():
    """The 240 knots of rosa/spectral.py:164-188 (QUANT_STEPS 0.4/0.2/0.1/0.05 -> plateaus 0.5 .. 4.5)."""
    steps = [0.4, 0.2, 0.1, 0.05]
    p1, p2, p3, p4 = np.diff(list(reversed(steps + [0])))
    xs = [
        torch.linspace(-0.1, 0.025, 101)[:-1], torch.linspace(0.025, p1, 11)[:-1], torch.linspace(p1, p1 + p2, 11)[:-1],
        torch.linspace(p1 + p2, p1 + p2 + p3, 11)[:-1], torch.linspace(p1 + p2 + p3, 0.5, 11)[:-1],
        torch.linspace(0.5, 1.1, 100),
    ]
    ys = torch.cat((
        0.5 * torch.ones(len(xs[0])), xs[1] / p1, (xs[2] - p1) / p2 + 1, (xs[3] - p1 - p2) / p3 + 2,
        (xs[4] - p1 - p2 - p3) / p4 + 3, 4.5 * torch.ones(len(xs[5])),
    ))
    return torch.cat(xs).float(), ys.float()


@functools.lru_cache(maxsize=None)
def cens_spline(device):
    xs, ys = cens_quantiser_knots()
    a, b, c, d = natural_cubic_spline_np(xs.numpy(), ys.numpy())
    dev = _dev(device)
    f = lambda v: torch.from_numpy(np.asarray(v, dtype=np.float32)).to(dev)
    return xs.to(dev), f(a), f(b), f(c), f(d)


@functools.lru_cache(maxsize=None)
def cens_window(win_len_smooth: int, device) -> torch.Tensor:
    """rosa/spectral.py:274-275: periodic hann(win_len_smooth + 2) / sum."""
    w = torch.hann_window(win_len_smooth + 2)
    return (w / torch.sum(w)).to(_dev(device))


@functools.lru_cache(maxsize=None)
def tonnetz_phi(n_chroma: int, device) -> torch.Tensor:
    """features/audio.py:50-55 (endpoint-inclusive linspace(0, 12, n_chroma) kept as written)."""
    dim_map = torch.linspace(0, 12, n_chroma)
    scale = torch.tensor([7.0 / 6, 7.0 / 6, 3.0 / 2, 3.0 / 2, 2.0 / 3, 2.0 / 3])
    V = scale.reshape(-1, 1) * dim_map
    V[::2] -= 0.5
    R = torch.tensor([1, 1, 1, 1, 0.5, 0.5])
    return (R[:, None] * torch.cos(torch.pi * V)).contiguous().to(_dev(device))


@functools.lru_cache(maxsize=64)
def _cardinal_spline_coeffs(n_points: int):
    """(a, b, two_c, three_d) [n_points, n_points - 1] of the natural cubic splines through the unit vectors on
    linspace(0, 1, n_points) (fp64)."""
    t_in = torch.linspace(0, 1, n_points).numpy()
    eye = np.eye(n_points)
    per_knot = [natural_cubic_spline_np(t_in, eye[k]) for k in range(n_points)]
    return tuple(np.stack([c[j] for c in per_knot]) for j in range(4))


def spline_loop_weights_cpu(n_knots: int, size: int, n_loops=1) -> torch.Tensor:
    """Cardinal weights of ssar/random/latent.py:7-13: out[t] = sum_k W[t, k] y[k] for the natural cubic spline through
    (y_0 .. y_{K-1}, y_0) on linspace(0, 1, K + 1), evaluated at linspace(0, n_loops, size) % 1 (upstream
    torchcubicspline evaluate: index = clamp(bucketize(t) - 1), value = a + b f + two_c f^2 / 2 + three_d f^3 / 3).
    The closing knot's weight is folded onto knot 0.  Returns fp32 [size, n_knots]."""
    K1 = n_knots + 1
    t_in = torch.linspace(0, 1, K1)
    t_out = torch.linspace(0, n_loops, size) % 1
    idx = (torch.bucketize(t_out, t_in) - 1).clamp(0, K1 - 2)
    frac = (t_out - t_in[idx]).double().numpy()[:, None]
    idx = idx.numpy()
    if K1 == 2:
        t64 = t_in.double().numpy()
        slope = 1.0 / (t64[1] - t64[0])
        W = np.concatenate([1.0 - slope * frac, slope * frac], axis=1)
    else:
        a, b, c2, d3 = _cardinal_spline_coeffs(K1)  # each [K1, K1 - 1]
        W = a.T[idx] + (b.T[idx] + (0.5 * c2.T[idx] + d3.T[idx] * frac / 3.0) * frac) * frac
    W = W.copy()
    W[:, 0] += W[:, -1]
    return torch.from_numpy(W[:, :-1].astype(np.float32))


def knot_loop_weights_cpu(n_latents: int, size: int, n_loops: int) -> torch.Tensor:
    """Cardinal weights of maua audioreactive/latent.py:84-92 (spline_loops): natural cubic spline through
    cat([y] * n_loops + [y[0]]) on linspace(0, 1, K n_loops + 1), evaluated at linspace(0, 1, size); the weight of knot j
    is folded onto latent j % K.  Returns fp32 [size, n_latents]."""
    K1 = n_latents * n_loops + 1
    t_in = torch.linspace(0, 1, K1)
    t_out = torch.linspace(0, 1, size)
    idx = (torch.bucketize(t_out, t_in) - 1).clamp(0, K1 - 2)
    frac = (t_out - t_in[idx]).double().numpy()[:, None]
    idx = idx.numpy()
    a, b, c2, d3 = _cardinal_spline_coeffs(K1)
    Wf = a.T[idx] + (b.T[idx] + (0.5 * c2.T[idx] + d3.T[idx] * frac / 3.0) * frac) * frac  # [size, K1]
    W = np.zeros((size, n_latents))
    for j in range(K1):
        W[:, j % n_latents] += Wf[:, j]
    return torch.from_numpy(W.astype(np.float32))


def _upload_async(t: torch.Tensor, device) -> torch.Tensor:
    """Host tensor -> device through PINNED staging (torch's caching host allocator keeps the block alive until the copy
    has run).  A copy from pageable memory synchronises the stream before it starts: per-track tables (a dataset of
    distinct track lengths misses every length-keyed cache) then stall the host behind all queued kernels."""
    pinned = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    pinned.copy_(t)
    return pinned.to(_dev(device), non_blocking=True)


@functools.lru_cache(maxsize=64)
def _cardinal_tables_device(K1: int, device):
    """(t_in fp32 [K1], coefficient tables fp64 [K1 - 1, 4, K1]) of `_cardinal_spline_coeffs` on the device."""
    a, b, c2, d3 = _cardinal_spline_coeffs(K1)
    # (np.stack of transposed views comes back F-ordered: make the [segment, 4, knot] layout the kernel reads explicit)
    tab = torch.from_numpy(np.ascontiguousarray(np.stack([a.T, b.T, c2.T, d3.T], axis=1)))
    return torch.linspace(0, 1, K1).to(_dev(device)), tab.to(_dev(device)).contiguous()


@functools.lru_cache(maxsize=256)
def spline_loop_weights(n_knots: int, size: int, n_loops, device) -> torch.Tensor:
    """`spline_loop_weights_cpu` on the device, cached (the weights depend only on the knot count, length and loop count,
    so every sub-patch / clip of a batch of equal-length clips shares them).  The evaluation points are made on the host
    ONE kernel (`ssar_spline_loop_weights`): the evaluation points `linspace(0, n_loops, size) % 1` (latent.py:10), the
    table lookup and the cubic in fp64 with the host function's expression — a dataset of tracks of distinct lengths
    misses the cache for every track, and the numpy version cost 0.6 ms (a torch-op version 0.24 ms) of host time per
    call.  The device linspace can differ from torch's CPU linspace by an ulp of t (ATen's vectorised CPU kernel rounds
    twice per element); the spline is continuous, so the weights move by ~1e-7.  Read-only result."""
    K1 = n_knots + 1
    if K1 == 2:
        return _upload_async(spline_loop_weights_cpu(n_knots, size, n_loops), device)
    t_in, tab = _cardinal_tables_device(K1, device)
    from . import _cabi as cabi

    out = torch.empty(size, n_knots, dtype=torch.float32, device=t_in.device)
    cabi.call("ssar_spline_loop_weights", cabi.ptr(tab), cabi.ptr(t_in), cabi.i32(K1), cabi.i32(size), cabi.f32(float(n_loops)),
              cabi.ptr(out), nbytes=cabi.nbytes_of(out))
    return out


@functools.lru_cache(maxsize=256)
def loop_phase(n_loops, length: int, device) -> torch.Tensor:
    """ssar/random/noise.py:49: linspace(0, n_loops * 2 pi, length), evaluated on the CPU like the reference (the
    module is built on the default device and moved afterwards), cached on the device.  Read-only."""
    return _upload_async(torch.linspace(0, n_loops * 2 * torch.pi, length), device)
