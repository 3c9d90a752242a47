"""Oracle: tuning estimate, constant-Q transform, chroma / CENS, tonnetz (torch CPU, fp32).  TEST INFRASTRUCTURE ONLY.

Restates ssar/features/rosa/pitch.py:9-120, rosa/constantq.py:13-269, rosa/convert.py:69-130,
rosa/spectral.py:164-325 and features/audio.py:44-57 for the one configuration the reference reaches
(C1, 7 octaves x 36 bins per octave, filter_scale 1, sparsity 0.01, gamma 0).

torchcubicspline boundary (UNPINNED, see package docstring): `natural_spline` below follows the published upstream
algorithm; SPLINE_AS_WRITTEN selects how the reference's own spline_eval (rosa/spectral.py:192-202) reads the
coefficients — True (default) = exactly as written (two_c and three_d used undivided), False = upstream's evaluate.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import dsp

SPLINE_AS_WRITTEN = True
BPO, N_OCT, N_BINS = 36, 7, 252
C1 = 440.0 * 2.0 ** ((24 - 69) / 12.0)


# ----------------------------------------------------------------------------------------------
# tuning
# ----------------------------------------------------------------------------------------------
def piptrack(y, sr, n_fft=2048, fmin=150.0, fmax=4000.0, threshold=0.1):
    """rosa/pitch.py:27-83 -> (pitches, mags) [1 + n_fft/2, N // (n_fft/4)]."""
    S = dsp.spectrogram(y, n_fft, n_fft // 4, power=1)
    F = S.shape[0]
    freqs = torch.linspace(0, float(sr) / 2, F)
    up, dn, mid = S[2:], S[:-2], S[1:-1]
    avg = torch.zeros_like(S)
    shift = torch.zeros_like(S)
    avg[1:-1] = 0.5 * (up - dn)
    den = 2 * mid - up - dn
    shift[1:-1] = avg[1:-1] / (den + (den.abs() < torch.finfo(torch.float32).tiny))
    gated = S * (S > threshold * S.max(0).values)
    z = torch.zeros(1, S.shape[1])
    is_peak = (gated > torch.cat([z, gated[:-1]])) & (gated >= torch.cat([gated[1:], z]))
    ok = is_peak & ((freqs >= max(fmin, 0)) & (freqs < min(fmax, float(sr) / 2)))[:, None]
    k = torch.arange(F, dtype=torch.float32)[:, None]
    pitches = torch.where(ok, (k + shift) * float(sr) / n_fft, torch.zeros(()))
    mags = torch.where(ok, S + 0.5 * avg * shift, torch.zeros(()))
    return pitches, mags


def pitch_tuning(freqs, resolution=0.01, bins_per_octave=12):
    """rosa/pitch.py:98-120."""
    freqs = freqs[freqs > 0]
    if freqs.numel() == 0:
        return torch.tensor(0.0)
    res = (bins_per_octave * torch.log2(freqs / (440.0 / 16))) % 1.0
    res = torch.where(res >= 0.5, res - 1.0, res)
    n = int(np.ceil(1.0 / resolution))
    counts = torch.histc(res, bins=n, min=-0.5, max=0.5)
    return torch.linspace(-0.5, 0.5, n + 1)[int(torch.argmax(counts))]


def estimate_tuning(y, sr, n_fft=2048, bins_per_octave=12):
    """rosa/pitch.py:9-24."""
    p, m = piptrack(y, sr, n_fft)
    tracked = p > 0
    thr = torch.median(m[tracked]) if tracked.any() else 0.0
    return pitch_tuning(p[(m >= thr) & tracked], bins_per_octave=bins_per_octave)


# ----------------------------------------------------------------------------------------------
# constant-Q
# ----------------------------------------------------------------------------------------------
def filter_lengths(sr, fmin, n_bins):
    """rosa/constantq.py:215-220 with filter_scale 1, gamma 0."""
    alpha = 2.0 ** (1.0 / BPO) - 1.0
    freq = fmin * (2.0 ** (torch.arange(n_bins, dtype=torch.float) / BPO))
    return (1.0 / alpha) * sr / (freq + 0.0 / alpha)


def octave_fft_basis(sr, fmin, sparsity=0.01):
    """rosa/constantq.py:118-164 + 223-263: 36 Hann-windowed complex exponentials, L1-normalised, centred in a
    power-of-two frame, scaled by length / n_fft, FFT'd, one-sided, smallest 1 % of each row's mass zeroed."""
    lengths = filter_lengths(sr, fmin, BPO)
    freqs = fmin * (2.0 ** (torch.arange(BPO, dtype=torch.float) / BPO))
    n_fft = int(2.0 ** torch.ceil(torch.log2(lengths.max())))
    B = torch.zeros(BPO, n_fft, dtype=torch.complex64)
    for r in range(BPO):
        h = torch.div(lengths[r], 2, rounding_mode="floor")
        n = torch.arange(-h, h, dtype=torch.float)
        g = torch.exp(n * 1j * 2 * torch.pi * freqs[r] / sr) * torch.hann_window(len(n))
        g = g / g.abs().sum()
        lp = (n_fft - len(n)) // 2
        B[r, lp : lp + len(n)] = g
    B = B * (lengths[:, None] / float(n_fft))
    F = torch.fft.fft(B, dim=1)[:, : n_fft // 2 + 1]
    mag = F.abs()
    order = torch.sort(mag, dim=1).values
    cum = torch.cumsum(order / mag.sum(1, keepdim=True), dim=1)
    first_kept = torch.argmin((cum < sparsity).to(torch.uint8), dim=1)
    cut = order[torch.arange(BPO), first_kept]
    return torch.where(mag >= cut[:, None], F, torch.zeros((), dtype=F.dtype)), n_fft


def cqt(y, sr, hop_length=1024, tuning=None):
    """rosa/constantq.py:13-115 (cqt -> vqt with gamma 0) -> complex64 [252, T]."""
    if tuning is None:
        tuning = estimate_tuning(y, sr, bins_per_octave=BPO)
    fmin = torch.tensor(C1).float() * 2.0 ** (torch.as_tensor(tuning) / BPO)
    top = (fmin * 2.0 ** (torch.arange(N_BINS, dtype=torch.float) / BPO))[-BPO:].min()
    sig, sr_i, hop_i = y.float(), float(sr), hop_length
    octaves = []
    for i in range(N_OCT):
        if i > 0:
            sig = dsp.resample(sig, 2, 1, method="kaiser") * np.float32(np.sqrt(2))
            sr_i /= 2.0
            hop_i //= 2
        basis, n_fft = octave_fft_basis(sr_i, top * 2.0 ** -i)
        basis = basis * np.sqrt(2 ** i)
        D = dsp.stft(sig, n_fft, hop_i, window=None)[:, :-1]
        octaves.append(basis @ D)
    T = min(o.shape[1] for o in octaves)
    V = torch.cat([o[:, :T] for o in reversed(octaves)], dim=0)
    return V / torch.sqrt(filter_lengths(float(sr), fmin, N_BINS))[:, None]


def chroma_fold_matrix(n_input=N_BINS, n_chroma=12):
    """rosa/convert.py:69-117 with fmin = C1 (MIDI 24 -> no rotation), window None."""
    merge = BPO // n_chroma
    M = torch.repeat_interleave(torch.eye(n_chroma), merge, dim=1)
    M = torch.roll(M, -(merge // 2), dims=1)
    M = M.repeat(1, math.ceil(n_input / BPO))[:, :n_input]
    roll = int(round(((12 * (math.log2(C1) - math.log2(440.0)) + 69) % 12) * (n_chroma / 12.0)))
    return torch.roll(M, roll, dims=0)


def chroma_cqt(y, sr, hop_length=1024, tuning=None, norm=True):
    """rosa/spectral.py:286-325."""
    chroma = chroma_fold_matrix() @ cqt(y, sr, hop_length, tuning).abs()
    chroma = torch.where(chroma < 0.0, torch.zeros(()), chroma)
    return chroma / chroma.max() if norm else chroma


# ----------------------------------------------------------------------------------------------
# CENS soft quantiser
# ----------------------------------------------------------------------------------------------
def quantiser_knots():
    """rosa/spectral.py:164-188."""
    p1, p2, p3, p4 = np.diff(list(reversed([0.4, 0.2, 0.1, 0.05] + [0])))
    xs = [torch.linspace(-0.1, 0.025, 101)[:-1], torch.linspace(0.025, p1, 11)[:-1], torch.linspace(p1, p1 + p2, 11)[:-1],
          torch.linspace(p1 + p2, p1 + p2 + p3, 11)[:-1], torch.linspace(p1 + p2 + p3, 0.5, 11)[:-1],
          torch.linspace(0.5, 1.1, 100)]
    ys = torch.cat((0.5 * torch.ones(100), xs[1] / p1, (xs[2] - p1) / p2 + 1, (xs[3] - p1 - p2) / p3 + 2,
                    (xs[4] - p1 - p2 - p3) / p4 + 3, 4.5 * torch.ones(100)))
    return torch.cat(xs), ys


def natural_spline(t, y):
    """Natural cubic spline (second derivative zero at both ends) in torchcubicspline's parameterisation
    (a, b, two_c, three_d) per interval; solved densely in fp64."""
    t, y = t.double().numpy(), y.double().numpy()
    n, h, dy = len(t), np.diff(t), np.diff(y)
    A, r = np.zeros((n, n)), np.zeros(n)
    for i in range(n - 1):  # each interval contributes to the derivative equations of both of its ends
        for j in (i, i + 1):
            A[j, j] += 2.0 / h[i]
            A[j, i + 1 if j == i else i] += 1.0 / h[i]
            r[j] += 3.0 * dy[i] / h[i] ** 2
    k = np.linalg.solve(A, r)
    two_c = (6 * dy / h - 4 * k[:-1] - 2 * k[1:]) / h
    three_d = (-6 * dy / h + 3 * (k[:-1] + k[1:])) / h ** 2
    return tuple(torch.from_numpy(v).float() for v in (y[:-1], k[:-1], two_c, three_d))


_QUANT = None


def spline_quantize(chroma):
    """rosa/spectral.py:192-219."""
    global _QUANT
    if _QUANT is None:
        xs, ys = quantiser_knots()
        _QUANT = (xs, *natural_spline(xs, ys))
    xs, a, b, c2, d3 = _QUANT
    idx = (torch.bucketize(chroma, xs) - 1).clamp(0, len(xs) - 2)
    f = chroma - xs[idx]
    if SPLINE_AS_WRITTEN:
        w = a[idx] + (b[idx] + (c2[idx] + d3[idx] * f) * f) * f
    else:
        w = a[idx] + (b[idx] + (0.5 * c2[idx] + d3[idx] * f / 3) * f) * f
    u = w - 0.5
    saw = u - torch.floor(u) - 0.5
    m20 = 1 / (1 + np.exp(-20)) - 0.5
    return 0.25 * (torch.floor(u) + 1 / (2 * m20) * 1 / (1 + torch.exp(-2 * 20 * saw)))


def chroma_cens(y, sr, hop_length=1024, tuning=None, win_len_smooth=41):
    """rosa/spectral.py:239-280."""
    chroma = chroma_cqt(y, sr, hop_length, tuning, norm=False)
    chroma = chroma / chroma.abs().sum(0)
    q = spline_quantize(chroma)
    if win_len_smooth:
        w = torch.hann_window(win_len_smooth + 2)
        w = w / w.sum()
        L = len(w)
        left = (L - 1) // 2
        qp = torch.nn.functional.pad(q, (left, L - 1 - left))
        q = torch.stack([(qp[:, k : k + q.shape[1]] * w[k]) for k in range(L)]).sum(0)
    return q / torch.sqrt((q * q).sum(0))


def chromagram(audio, sr):
    """features/audio.py:44-45 -> [T, 12]."""
    return chroma_cens(dsp.harmonic(audio), sr).T


def tonnetz(y, sr, chroma=None):
    """features/audio.py:48-57 -> [T, 6]; `chroma` [12, T] overrides the default chromagram(y).T."""
    if chroma is None:
        chroma = chromagram(y, sr).T
    dim_map = torch.linspace(0, 12, chroma.shape[0])
    V = torch.tensor([7.0 / 6, 7.0 / 6, 3.0 / 2, 3.0 / 2, 2.0 / 3, 2.0 / 3])[:, None] * dim_map
    V[::2] -= 0.5
    phi = torch.tensor([1, 1, 1, 1, 0.5, 0.5])[:, None] * torch.cos(torch.pi * V)
    return (phi @ (chroma / chroma.abs().sum(0))).T
