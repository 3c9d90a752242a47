"""GPU parity: tuning estimate, CQT, chroma / CENS, tonnetz and spectral contrast vs goldens and the oracle."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_clip
from ssar_oracle import cqt as ocqt
from ssar_oracle import dsp

pytestmark = pytest.mark.gpu

TOL_CQT = 1e-4     # north-star tolerance on spectra / CQT magnitudes
TOL_CHROMA = 1e-3  # north-star tolerance on chroma / tonnetz / contrast


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_chroma_chain_vs_golden():
    from ssar_b200.features import audio
    from ssar_b200.features.rosa import constantq, pitch, spectral

    g = load("chroma")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"])).cuda()
    assert float(pitch.estimate_tuning(x, sr, bins_per_octave=36)) == float(g["tuning36"])
    assert float(pitch.estimate_tuning(x, sr)) == float(g["tuning12"])
    C = constantq.cqt(x, sr, n_bins=252, bins_per_octave=36, tuning=None)
    assert tuple(C.shape) == tuple(g["cqt_re"].shape)
    ref = torch.complex(g["cqt_re"], g["cqt_im"])
    assert rel_l2(torch.view_as_real(C.cpu().resolve_conj()), torch.view_as_real(ref)) < TOL_CQT
    C0 = constantq.cqt(x, sr, n_bins=252, bins_per_octave=36, tuning=0.0)
    assert rel_l2(C0.abs(), g["cqt0_abs"]) < TOL_CQT
    assert rel_l2(spectral.chroma_cqt(x, sr), g["chroma_cqt"]) < TOL_CHROMA
    cens = spectral.chroma_cens(x, sr)
    assert tuple(cens.shape) == tuple(g["cens"].shape)
    assert rel_l2(cens, g["cens"]) < TOL_CHROMA
    assert rel_l2(audio.chromagram(x, sr), g["chromagram"]) < TOL_CHROMA
    assert rel_l2(audio.tonnetz(x, sr, chroma_fn=lambda a, s: cens), g["tonnetz_of_cens"]) < TOL_CHROMA
    assert rel_l2(audio.tonnetz(x, sr), g["tonnetz"]) < 2e-3  # harmonic(harmonic(x)) chain: two ISTFTs deep
    assert rel_l2(audio.spectral_contrast(x, sr), g["contrast"]) < TOL_CHROMA
    assert rel_l2(audio.spectral_contrast(x, sr, linear=True), g["contrast_linear"]) < TOL_CHROMA


def test_chroma_batched_other_tunings_vs_oracle():
    """Batch of clips with different estimated tunings (per-clip basis slots) vs the oracle, 6 s clips."""
    from ssar_b200 import ops
    from ssar_b200.features.rosa import spectral

    sr, n = 24576, 24576 * 6
    xs = torch.stack([synth_clip(n, sr, i) for i in (0, 7, 11)])
    idx, val = ops.estimate_tuning(xs.cuda(), sr, bins_per_octave=36)
    cens = spectral.chroma_cens(xs.cuda(), sr)
    for b in range(xs.shape[0]):
        t_ref = ocqt.estimate_tuning(xs[b], sr, bins_per_octave=36)
        assert float(val[b]) == float(t_ref), (b, float(val[b]), float(t_ref))
        assert rel_l2(cens[b], ocqt.chroma_cens(xs[b], sr)) < TOL_CHROMA


def test_contrast_full_length_properties():
    """cfg24 length (T = 5760): shape, finiteness, and equality with the oracle on a strided subset of frames."""
    from ssar_b200.features import audio

    sr, n = 24576, 24576 * 240
    x = synth_clip(n, sr, 2)
    c = audio.spectral_contrast(x.cuda(), sr)
    assert tuple(c.shape) == (5760, 7) and bool(torch.isfinite(c).all())
    ref = dsp.spectral_contrast(x, sr)
    assert rel_l2(c, ref) < TOL_CHROMA


def test_contrast_sorted_columns_equal_general_kernel():
    """Bands of <= 512 bins take the sorted-lane-column kernel; declaring a wider slab forces the general arg-min
    kernel on the same bands.  Same values summed in the same order -> bit-equal, ragged band sizes included."""
    import dataclasses

    from ssar_b200 import ops, plan

    g = torch.Generator().manual_seed(5)
    mag = torch.rand(2, 301, 1025, generator=g).cuda()
    mag[0, :, 40:60] = 0.25  # ties
    for n_fft_bins, fmin, nb in ((2048, 200.0, 6), (2048, 300.0, 6)):
        bands = plan.contrast_bands(22050.0, n_fft_bins, fmin, nb, 0.02, "cuda")
        assert bands.max_count <= 512
        fast = ops.spectral_contrast(mag, bands)
        slow = ops.spectral_contrast(mag, dataclasses.replace(bands, max_count=600))
        assert torch.equal(fast, slow)


def test_cqt_tensor_core_vs_fft_path_and_golden():
    """`ssar_cqt_tc` (time-domain tcgen05 contraction over TMA frame views with overlapping rows, fp16 3-term split) against
    the per-octave FFT + banded projection path, the reference golden and the oracle: batch with different tunings, a
    quiet clip and a loud clip (per-clip operand scaling), ragged length, edge frames (reflect padding)."""
    from ssar_b200 import ops

    g = load("chroma")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"])).cuda()[None]
    idx, _ = ops.estimate_tuning(x, sr, bins_per_octave=36)
    ref = torch.complex(g["cqt_re"], g["cqt_im"])                                      # [252, T]
    C_tc = torch.view_as_complex(ops.cqt(x, sr, tuning_idx=idx, tensor_core=True)[0]).T
    C_ff = torch.view_as_complex(ops.cqt(x, sr, tuning_idx=idx, tensor_core=False)[0]).T
    assert tuple(C_tc.shape) == tuple(ref.shape)
    e_tc = rel_l2(torch.view_as_real(C_tc.contiguous()), torch.view_as_real(ref))
    e_ff = rel_l2(torch.view_as_real(C_ff.contiguous()), torch.view_as_real(ref))
    print("cqt rel-L2 vs reference golden: tensor core", e_tc, "fft path", e_ff)
    assert e_tc < TOL_CQT and e_ff < TOL_CQT
    for o in range(7):  # every octave on its own (the high octaves carry most of the energy)
        sl = slice(252 - 36 * (o + 1), 252 - 36 * o)
        assert rel_l2(torch.view_as_real(C_tc[sl].contiguous()), torch.view_as_real(ref[sl].contiguous())) < TOL_CQT, o
    n = sr * 7 + 333
    xs = torch.stack([synth_clip(n, sr, 0), 1e-3 * synth_clip(n, sr, 7), 40.0 * synth_clip(n, sr, 11)]).cuda()
    idx, _ = ops.estimate_tuning(xs, sr, bins_per_octave=36)
    a = ops.cqt(xs, sr, tuning_idx=idx, tensor_core=True)
    b = ops.cqt(xs, sr, tuning_idx=idx, tensor_core=False)
    assert a.shape == b.shape
    for c in range(3):
        assert rel_l2(a[c], b[c]) < 2e-5, (c, rel_l2(a[c], b[c]))
        assert rel_l2(a[c, :2], b[c, :2]) < 2e-5 and rel_l2(a[c, -2:], b[c, -2:]) < 2e-5  # reflect-padded edge frames
    C_or = ocqt.cqt(xs[1].cpu(), sr)
    assert rel_l2(torch.view_as_real(torch.view_as_complex(a[1]).T.contiguous()), torch.view_as_real(C_or.contiguous())) < TOL_CQT


def test_fused_piptrack_equals_spectrogram_path():
    """estimate_tuning with the peak picking in the STFT epilogue == spectrogram to HBM + ssar_piptrack: same tuning, and
    the same multiset of (pitch, magnitude) pairs per clip bit for bit (the lists are unordered)."""
    from ssar_b200 import ops, plan

    sr, n = 24576, 24576 * 9 + 77
    xs = torch.stack([synth_clip(n, sr, i) for i in (0, 7, 11, 4)]).cuda()
    try:
        ops.FUSED_PIPTRACK = False
        i0, v0 = ops.estimate_tuning(xs, sr, bins_per_octave=36)
        ops.FUSED_PIPTRACK = True
        i1, v1 = ops.estimate_tuning(xs, sr, bins_per_octave=36)
    finally:
        ops.FUSED_PIPTRACK = True
    assert torch.equal(i0, i1) and torch.equal(v0, v1)
    k_lo, k_hi = plan.piptrack_bin_range(float(sr), 2048)
    T = n // 512
    o = ops.frames_fft(xs, 2048, 512, T, torch.hann_window, spec_power=1, piptrack=dict(k_lo=k_lo, k_hi=k_hi, threshold=0.1, sr=sr))
    S = o["spec"]
    import ctypes as C

    from ssar_b200 import _cabi as cabi
    cap = o["pip_pitch"].shape[1]
    pitch, mag = torch.empty_like(o["pip_pitch"]), torch.empty_like(o["pip_mag"])
    count = torch.empty_like(o["pip_count"])
    cabi.call("ssar_piptrack", cabi.ptr(S), cabi.i64(S.shape[2]), cabi.i32(4), cabi.i32(T), cabi.i32(S.shape[2]), cabi.i32(k_lo),
              cabi.i32(k_hi), cabi.f32(0.1), cabi.f32(float(sr)), cabi.i32(2048), cabi.i64(cap), cabi.ptr(pitch), cabi.ptr(mag),
              cabi.ptr(count))
    assert torch.equal(count, o["pip_count"]) and int(count.min()) > 0
    for c in range(4):
        m = int(count[c])
        a = torch.stack([pitch[c, :m], mag[c, :m]], 1).cpu()
        b = torch.stack([o["pip_pitch"][c, :m], o["pip_mag"][c, :m]], 1).cpu()
        key = lambda t: t[torch.argsort(t[:, 0] * 1e3 + t[:, 1], stable=True)]
        assert torch.equal(torch.sort(a[:, 0]).values, torch.sort(b[:, 0]).values)
        assert torch.equal(torch.sort(a[:, 1]).values, torch.sort(b[:, 1]).values)
