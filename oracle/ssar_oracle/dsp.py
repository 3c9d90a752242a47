"""Oracle: framing, STFT-family transforms and per-frame spectral features (torch CPU, fp32).

TEST INFRASTRUCTURE (see package docstring).  Written as explicit framing + rFFT rather than torch.stft so the
restatement is independent of the op the reference happens to call.
"""
from __future__ import annotations

import math

import numpy as np
import torch


# ----------------------------------------------------------------------------------------------
# framing / transforms
# ----------------------------------------------------------------------------------------------
def reflect_pad(y: torch.Tensor, p: int) -> torch.Tensor:
    """torch.nn.functional.pad(mode="reflect") on the last axis: edge sample not repeated."""
    if p == 0:
        return y
    left = y[..., 1 : p + 1].flip(-1)
    right = y[..., -p - 1 : -1].flip(-1)
    return torch.cat([left, y, right], dim=-1)


def frames(y: torch.Tensor, n_fft: int, hop: int) -> torch.Tensor:
    """[..., N] -> [..., 1 + N//hop, n_fft] centred, reflect-padded frames (torch.stft center=True)."""
    return reflect_pad(y, n_fft // 2).unfold(-1, n_fft, hop)


def hann_periodic(n: int) -> torch.Tensor:
    """torch.hann_window(n) (periodic): 0.5 - 0.5 cos(2 pi k / n)."""
    return torch.hann_window(n, dtype=torch.float32)


def stft(y, n_fft=2048, hop_length=1024, window="hann"):
    """ssar/features/rosa/spectral.py:10-21 -> complex64 [F, 1 + N//hop] (or [B, F, T+1])."""
    fr = frames(y.float(), n_fft, hop_length)
    if window is not None:
        fr = fr * (hann_periodic(n_fft) if isinstance(window, str) else window)
    return torch.fft.rfft(fr, dim=-1).transpose(-1, -2)


def istft(spec, n_fft=2048, hop_length=1024, window="hann", length=None):
    """ssar/features/rosa/spectral.py:24-32 (torch.istft semantics): irfft per frame, window, overlap-add,
    divide by the summed squared window, drop n_fft/2 from the start, trim / zero-pad to `length`."""
    w = hann_periodic(n_fft) if isinstance(window, str) else (torch.ones(n_fft) if window is None else window)
    fr = torch.fft.irfft(spec.transpose(-1, -2), n=n_fft, dim=-1) * w  # [T, n_fft]
    T = fr.shape[-2]
    total = n_fft + hop_length * (T - 1)
    out = torch.zeros(total, dtype=torch.float32)
    env = torch.zeros(total, dtype=torch.float32)
    idx = (torch.arange(T)[:, None] * hop_length + torch.arange(n_fft)[None, :]).reshape(-1)
    out.index_add_(0, idx, fr.reshape(-1))
    env.index_add_(0, idx, (w * w).repeat(T))
    start = n_fft // 2
    end = total - n_fft // 2 if length is None else start + length
    out, env = out[start:end], env[start:end]
    out = out / env
    if length is not None and out.shape[0] < length:
        out = torch.nn.functional.pad(out, (0, length - out.shape[0]))
    return out


def spectrogram(y, n_fft=2048, hop_length=1024, power=1, window="hann"):
    """rosa/spectral.py:59-62: last STFT frame dropped -> [F, N//hop]."""
    if hop_length is None:
        hop_length = n_fft // 4
    return stft(y, n_fft, hop_length, window)[..., :-1].abs() ** power


# ----------------------------------------------------------------------------------------------
# mel
# ----------------------------------------------------------------------------------------------
F_SP = 200.0 / 3
MIN_LOG_HZ = 1000.0
MIN_LOG_MEL = MIN_LOG_HZ / F_SP
LOGSTEP = np.log(6.4) / 27.0


def hz_to_mel(f):
    """rosa/convert.py:15-41 (Slaney)."""
    f = torch.as_tensor(f, dtype=torch.float32)
    lin = f / F_SP
    return torch.where(f >= MIN_LOG_HZ, MIN_LOG_MEL + torch.log(f / MIN_LOG_HZ) / LOGSTEP, lin)


def mel_to_hz(m):
    """rosa/convert.py:44-66."""
    m = torch.as_tensor(m, dtype=torch.float32)
    return torch.where(m >= MIN_LOG_MEL, MIN_LOG_HZ * torch.exp(LOGSTEP * (m - MIN_LOG_MEL)), F_SP * m)


def mel_basis(sr, n_fft, n_mels=128, fmin=0.0, fmax=None):
    """rosa/spectral.py:73-110: triangular, area-normalised filters [n_mels, 1 + n_fft/2]."""
    fmax = float(sr) / 2 if fmax is None else fmax
    fftfreqs = torch.linspace(0, float(sr) / 2, 1 + n_fft // 2)
    edges = mel_to_hz(torch.linspace(float(hz_to_mel(fmin)), float(hz_to_mel(fmax)), n_mels + 2))
    width = torch.diff(edges)
    ramps = edges[:, None] - fftfreqs[None, :]
    W = torch.zeros(n_mels, 1 + n_fft // 2)
    for i in range(n_mels):
        W[i] = torch.clamp(torch.minimum(-ramps[i] / width[i], ramps[i + 2] / width[i + 1]), min=0)
    return W * (2.0 / (edges[2:] - edges[:-2]))[:, None]


def melspectrogram(y, sr, n_fft=2048, hop_length=1024, power=2.0, fmax=None):
    """rosa/spectral.py:65-70."""
    return mel_basis(sr, n_fft, fmax=fmax) @ spectrogram(y, n_fft, hop_length, power)


def power_to_db(x, top_db=80.0):
    """rosa/convert.py:7-12 with ref_value = 1, amin = 1e-10."""
    amin = torch.tensor(1e-10)
    db = 10.0 * torch.log10(torch.maximum(amin, x))
    db = db - 10.0 * torch.log10(torch.maximum(amin, torch.ones(())))
    if top_db is not None:
        db = torch.maximum(db, db.max() - top_db)
    return db


def dct_ortho(x):
    """rosa/spectral.py:35-56 with norm="ortho", over the last axis (restated as the explicit cosine sum)."""
    N = x.shape[-1]
    n = torch.arange(N, dtype=torch.float64)
    k = torch.arange(N, dtype=torch.float64)[:, None]
    D = torch.cos(math.pi * k * (2 * n + 1) / (2 * N))
    D[0] /= math.sqrt(N)
    D[1:] *= math.sqrt(2.0 / N)
    return (x.double() @ D.T).float()


# ----------------------------------------------------------------------------------------------
# per-frame features
# ----------------------------------------------------------------------------------------------
def rms(y, sr=None, frame_length=2048, hop_length=1024):
    """features/audio.py:31-37 -> [T, 1]."""
    fr = frames(y.float(), frame_length, hop_length)[:-1]
    return torch.sqrt(torch.mean(fr.abs() ** 2, dim=1)).unsqueeze(-1)


def onset_strength(y, sr, hop_length=1024, n_fft=2048, aggregate="mean"):
    """rosa/beat.py:10-23: mel (fmax 11025) -> dB -> positive first difference -> mean/median over bands ->
    shifted right by 1 + n_fft // (2 hop) frames, truncated to T."""
    S = power_to_db(melspectrogram(y, sr, n_fft=n_fft, hop_length=hop_length, fmax=11025.0).abs())
    flux = torch.clamp(S[:, 1:] - S[:, :-1], min=0)
    env = flux.mean(dim=0) if aggregate == "mean" else flux.median(dim=0).values
    shift = 1 + n_fft // (2 * hop_length)
    return torch.cat([torch.zeros(shift), env])[: S.shape[1]]


def plp(y, sr, hop_length=1024, win_length=1024, tempo_min=60, tempo_max=180):
    """rosa/beat.py:41-75: median-aggregated onset envelope -> Fourier tempogram (hop 1) -> keep the per-frame peak bin
    of the 60-180 BPM band at unit magnitude -> inverse STFT -> positive half -> min/max normalise."""
    env = onset_strength(y, sr, hop_length, aggregate="median")
    T = len(env)
    win = min(T, win_length)
    G = stft(env, win, 1)
    bpm = torch.linspace(0, float(sr * 60 / float(hop_length)) / 2, 1 + win // 2)
    G = torch.where(((bpm < tempo_min) | (bpm > tempo_max))[:, None], torch.zeros((), dtype=G.dtype), G)
    strength = torch.log1p(1e6 * G.abs())
    G = torch.where(strength < strength.max(0, keepdim=True).values, torch.zeros((), dtype=G.dtype), G)
    G = G / (torch.finfo(torch.float32).tiny ** 0.5 + G.abs().max(0, keepdim=True).values)
    p = istft(G, win, 1, length=T).clamp(min=0)
    p = p - p.min()
    return p / (p.max() + 1e-8)


def mfcc(y, sr, n_mfcc=20):
    """features/audio.py:60-65 -> [T, n_mfcc]."""
    S = power_to_db(melspectrogram(y, sr))
    return dct_ortho(S.T)[:, :n_mfcc]


def spectral_flatness(y, sr=None, n_fft=2048, hop_length=1024, amin=1e-10, power=2.0):
    """features/audio.py:126-133 -> [T, 1]."""
    S = torch.clamp(spectrogram(y, n_fft, hop_length, power=1.0) ** power, min=amin)
    return (torch.exp(torch.mean(torch.log(S), dim=0)) / torch.mean(S, dim=0)).unsqueeze(-1)


def spectral_contrast(y, sr, n_fft=2048, hop_length=1024, fmin=200.0, n_bands=6, quantile=0.02, linear=False):
    """features/audio.py:72-123 -> [T, n_bands + 1]: octave bands with edges 0, fmin, 2 fmin, ...; each band also
    takes the bin just below its lower edge (k > 0), the top band everything above; all but the top band drop their
    last bin; q = max(1, round(quantile * bins_in_band)) extremes are averaged."""
    S = spectrogram(y, n_fft, hop_length, power=1)
    freq = torch.linspace(0, float(sr) / 2, 1 + n_fft // 2)
    edges = torch.cat([torch.zeros(1), fmin * 2.0 ** torch.arange(0, n_bands + 1)])
    peak, valley = [], []
    for k in range(n_bands + 1):
        inside = torch.nonzero((freq >= edges[k]) & (freq <= edges[k + 1])).flatten()
        lo, hi = int(inside[0]), int(inside[-1])
        if k > 0:
            lo -= 1
        if k == n_bands:
            hi = len(freq) - 1
        n_in_band = hi - lo + 1
        if k < n_bands:
            hi -= 1
        q = max(1, int(torch.round(torch.tensor(quantile * n_in_band))))
        srt = torch.sort(S[lo : hi + 1], dim=0).values
        valley.append(srt[:q].mean(0))
        peak.append(srt[-q:].mean(0))
    peak, valley = torch.stack(peak), torch.stack(valley)
    if linear:
        return (peak - valley).T
    return (power_to_db(peak) - power_to_db(valley)).T


# ----------------------------------------------------------------------------------------------
# harmonic / percussive separation
# ----------------------------------------------------------------------------------------------
def median_filter_1d(S, k, dim):
    """features/processing.py:74-84 restricted to a 1-D window: reflect pad k//2, sliding window, torch.median
    (= lower median; exact middle for odd k), here via a full sort."""
    h = k // 2
    S = S.transpose(dim, -1)
    win = reflect_pad(S, h).unfold(-1, k, 1)
    out = torch.sort(win, dim=-1).values[..., (k - 1) // 2]
    return out.transpose(dim, -1)


def softmask(X, X_ref, power=2.0, split_zeros=False):
    """rosa/spectral.py:120-142."""
    Z = torch.maximum(X, X_ref)
    bad = Z < torch.finfo(torch.float32).tiny
    Z = torch.where(bad, torch.ones(()), Z)
    m, r = (X / Z) ** power, (X_ref / Z) ** power
    mask = m / (m + r)
    return torch.where(bad, torch.full((), 0.5 if split_zeros else 0.0), mask)


def hpss(D, ks=31, power=2.0, margin=1.0):
    """rosa/spectral.py:145-161 on a complex STFT [F, T]."""
    S = D.abs()
    phase = torch.exp(1j * torch.angle(D))
    harm = median_filter_1d(S, ks, dim=1)   # along time
    perc = median_filter_1d(S, ks, dim=0)   # along frequency
    split = margin == 1
    mh = softmask(harm, perc * margin, power, split)
    mp = softmask(perc, harm * margin, power, split)
    return (S * mh) * phase, (S * mp) * phase


def harmonic(y, margin=8.0):
    """features/audio.py:13-17."""
    return istft(hpss(stft(y), margin=margin)[0], length=len(y))


def percussive(y, margin=8.0):
    """features/audio.py:20-24."""
    return istft(hpss(stft(y), margin=margin)[1], length=len(y))


# ----------------------------------------------------------------------------------------------
# IIR band splits and resampling (third-party torchaudio algorithms restated; torchaudio is unpinned by env.yaml)
# ----------------------------------------------------------------------------------------------
def biquad_design(kind, sr, fc, Q=0.707):
    """torchaudio lowpass_biquad / highpass_biquad: RBJ cookbook, coefficient arithmetic in fp32."""
    w0 = torch.tensor(2 * math.pi * fc / sr, dtype=torch.float32)
    w0 = 2 * math.pi * torch.tensor(fc, dtype=torch.float32) / sr
    alpha = torch.sin(w0) / 2 / torch.tensor(Q, dtype=torch.float32)
    c = torch.cos(w0)
    if kind == "low":
        b = [(1 - c) / 2, 1 - c, (1 - c) / 2]
    else:
        b = [(1 + c) / 2, -1 - c, (1 + c) / 2]
    a = [1 + alpha, -2 * c, 1 - alpha]
    return np.array([float(v / a[0]) for v in b], np.float32), np.array([float(v / a[0]) for v in a], np.float32)


def lfilter_clamped(x, b, a):
    """torchaudio lfilter(clamp=True): direct-form recurrence from a zero state, output clamped to [-1, 1]."""
    from scipy.signal import lfilter

    y = lfilter(b.astype(np.float64), a.astype(np.float64), x.numpy().astype(np.float64))
    return torch.from_numpy(np.clip(y, -1.0, 1.0).astype(np.float32))


def low_pass(x, sr, fmax=200):
    """features/processing.py:141-142."""
    return lfilter_clamped(x, *biquad_design("low", sr, fmax))


def high_pass(x, sr, fmin=4000):
    """features/processing.py:149-150."""
    return lfilter_clamped(x, *biquad_design("high", sr, fmin))


def mid_pass(x, sr, fmin=200, fmax=4000):
    """features/processing.py:145-146 (high-pass at fmax, then low-pass at fmin — as written in the reference)."""
    return low_pass(high_pass(x, sr, fmax), sr, fmin)


def sinc_resample_kernel(orig, new, method="hann", zeros=6, rolloff=0.99, beta=14.769656459379492):
    """torchaudio _get_sinc_resample_kernel (fp32, as reached from resample(fp32 waveform))."""
    base = min(orig, new) * rolloff
    width = math.ceil(zeros * orig / base)
    idx = torch.arange(-width, width + orig, dtype=torch.float32)[None] / orig
    t = (torch.arange(0, -new, -1, dtype=torch.float32)[:, None] / new + idx) * base
    t = t.clamp(-zeros, zeros)
    if method == "hann":
        win = torch.cos(t * math.pi / zeros / 2) ** 2
    else:
        bt = torch.tensor(float(beta))
        win = torch.i0(bt * torch.sqrt(1 - (t / zeros) ** 2)) / torch.i0(bt)
    t = t * math.pi
    k = torch.where(t == 0, torch.ones(()), torch.sin(t) / t) * win * (base / orig)
    return k, width


def resample(x, orig_freq, new_freq, method="hann"):
    """torchaudio.functional.resample (supervised/data.py:63, rosa/constantq.py:83)."""
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    k, width = sinc_resample_kernel(orig, new, method)
    n = x.shape[-1]
    xp = torch.nn.functional.pad(x.float()[None, None], (width, width + orig))
    y = torch.nn.functional.conv1d(xp, k[:, None], stride=orig)  # [1, new, blocks]
    y = y[0].T.reshape(-1)
    return y[: math.ceil(new * n / orig)]
