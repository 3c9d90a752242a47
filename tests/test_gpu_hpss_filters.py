"""GPU parity: HPSS (31-tap medians + soft masks), ISTFT, biquad scan, polyphase resampler."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_clip
from ssar_oracle import dsp

pytestmark = pytest.mark.gpu
TOL = 1e-4


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_hpss_masks_and_waveforms_vs_golden():
    from ssar_b200.features import audio
    from ssar_b200.features.rosa import spectral

    g = load("hpss_filters")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    xd = x.cuda()
    H, P = spectral.hpss(spectral.stft(xd), margin=8.0)
    assert rel_l2(H.abs()[::4], g["hpss_h_abs"]) < TOL and rel_l2(P.abs()[::4], g["hpss_p_abs"]) < TOL
    Ho, Po = dsp.hpss(dsp.stft(x), margin=8.0)
    assert rel_l2(torch.view_as_real(H.cpu().contiguous()), torch.view_as_real(Ho.contiguous())) < TOL
    assert rel_l2(torch.view_as_real(P.cpu().contiguous()), torch.view_as_real(Po.contiguous())) < TOL
    yh, yp = audio.harmonic(xd), audio.percussive(xd)
    assert tuple(yh.shape) == tuple(x.shape)
    assert rel_l2(yh, g["harmonic"]) < TOL and rel_l2(yp, g["percussive"]) < TOL
    # margin 1 exercises the split_zeros branch (rosa/spectral.py:158)
    H1, P1 = spectral.hpss(spectral.stft(xd), margin=1.0)
    H1o, P1o = dsp.hpss(dsp.stft(x), margin=1.0)
    assert rel_l2(H1.abs(), H1o.abs()) < TOL and rel_l2(P1.abs(), P1o.abs()) < TOL


def test_istft_roundtrip_and_batch():
    from ssar_b200.features.rosa import spectral

    xs = torch.stack([synth_clip(24576 + 123, 24576, i) for i in range(3)]).cuda()
    D = spectral.stft(xs)
    y = spectral.istft(D, length=xs.shape[-1])
    assert tuple(y.shape) == tuple(xs.shape) and rel_l2(y, xs) < 1e-5
    # even lengths take the fused transform + overlap-add kernel (n_fft 2048, hop 1024): whole hops, a ragged tail,
    # a clip shorter than one run of blocks, and the oracle's torch.istft on the same spectrum
    for n in (24576 * 3, 24576 * 2 + 1000, 4096 + 2, 1026):
        xe = torch.stack([synth_clip(n, 24576, i) for i in range(2)]).cuda()
        De = spectral.stft(xe)
        ye = spectral.istft(De, length=n)
        assert tuple(ye.shape) == (2, n) and rel_l2(ye, xe) < 1e-5, n
        ref = dsp.istft(De[1].cpu(), 2048, 1024, length=n)
        assert rel_l2(ye[1], ref) < 1e-5, n
    # 1024-point path with hop 1 (the PLP configuration, rosa/beat.py:69)
    e = synth_clip(1500, 24, 5).abs().cuda()
    De = spectral.stft(e, n_fft=1024, hop_length=1)
    ye = spectral.istft(De, n_fft=1024, hop_length=1, length=1500)
    ref = dsp.istft(dsp.stft(e.cpu(), 1024, 1), 1024, 1, length=1500)
    assert rel_l2(ye, ref) < TOL and rel_l2(ye, e) < 1e-4


def test_biquads_vs_golden_and_long_clip():
    from ssar_b200.features import processing

    g = load("hpss_filters")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    xd = x.cuda()
    assert rel_l2(processing.low_pass(xd, sr), g["low"]) < TOL
    assert rel_l2(processing.high_pass(xd, sr), g["high"]) < TOL
    assert rel_l2(processing.mid_pass(xd, sr), g["mid"]) < TOL
    # long clip (many chunks per scan thread), batched, odd length, includes clamping
    xs = torch.stack([synth_clip(24576 * 60 + 37, 24576, i) for i in (6, 7)]) * 1.5
    for fn, ofn in [(processing.low_pass, dsp.low_pass), (processing.high_pass, dsp.high_pass)]:
        y = fn(xs.cuda(), 24576)
        assert float(y.abs().max()) <= 1.0
        for b in range(2):
            assert rel_l2(y[b], ofn(xs[b], 24576)) < TOL
    # 44.1 kHz, 200 Hz low-pass: pole radius closer to 1 (slow decay across chunk boundaries)
    x44 = synth_clip(44100 * 20 + 5, 44100, 8)
    assert rel_l2(processing.low_pass(x44.cuda(), 44100, 60), dsp.low_pass(x44, 44100, 60)) < TOL


def test_resample_vs_golden():
    from ssar_b200.resampling import resample

    g = load("resample")
    x = synth_clip(int(g["n"]), 44100, int(g["clip_index"]))
    y = resample(x.cuda(), 44100, 24576)
    assert tuple(y.shape) == tuple(g["to24576"].shape)
    assert rel_l2(y, g["to24576"]) < 1e-5
    y2 = resample(g["to24576"].cuda(), 24576, 12288, "kaiser_window")
    assert tuple(y2.shape) == tuple(g["half_kaiser"].shape) and rel_l2(y2, g["half_kaiser"]) < 1e-5
    xs = torch.stack([synth_clip(50001, 48000, i) for i in (1, 2)])
    yb = resample(xs.cuda(), 48000, 24576)
    for b in range(2):
        assert rel_l2(yb[b], dsp.resample(xs[b], 48000, 24576)) < 1e-5


def test_biquad_rms_equals_rms_of_biquad():
    """ops.biquad_rms (framed RMS from the filter's own output pass) == frames_rms(biquad(x)) — features/audio.py:31-37 on
    features/processing.py:141-150 — including the reflected first frame; the optional signal is biquad's, bit for bit."""
    import torch

    from ssar_b200 import ops, plan
    from ssar_b200.synth import synth_clip

    sr = 24576
    x = torch.stack([synth_clip(sr * 20 + 777, sr, i) for i in (2, 5)]).cuda()
    T = x.shape[1] // 1024
    for kind, fc in (("low", 200), ("high", 4000)):
        c = plan.biquad_coeffs(kind, sr, fc)
        y = ops.biquad(x, c, clamp=True)
        ref = ops.frames_rms(y, 2048, 1024, T)
        got, y2 = ops.biquad_rms(x, c, want_signal=True, n_frames=T)
        assert torch.equal(y, y2)
        assert got.shape == ref.shape and rel_l2(got, ref) < 1e-6
        assert float((got[:, 0] - ref[:, 0]).abs().max()) < 1e-6 * float(ref.abs().max())
        assert rel_l2(ops.biquad_rms(x, c, n_frames=T), ref) < 1e-6


def test_biquad_pair_equals_two_single_filters():
    """ssar_biquad_pair (two filters over one input per pass) == ssar_biquad / ssar_biquad_ss per filter, bit for bit:
    signals, framed RMS, ragged length (tile tail), batch of clips."""
    from ssar_b200 import ops, plan

    sr = 24576
    xs = torch.stack([synth_clip(sr * 7 + 333, sr, i) for i in (2, 6, 9)]).cuda()
    lo_c, hi_c = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
    T = xs.shape[1] // 1024
    (y_lo, r_lo), (y_hi, r_hi) = ops.biquad_pair(xs, lo_c, hi_c, want_signal=(True, True), want_rms=(True, True), n_frames=T)
    assert torch.equal(y_lo, ops.biquad(xs, lo_c)) and torch.equal(y_hi, ops.biquad(xs, hi_c))
    assert torch.equal(r_lo, ops.biquad_rms(xs, lo_c, n_frames=T)) and torch.equal(r_hi, ops.biquad_rms(xs, hi_c, n_frames=T))
    (n0, r0), (y1, n1) = ops.biquad_pair(xs, lo_c, hi_c, want_signal=(False, True), want_rms=(True, False), n_frames=T)
    assert n0 is None and n1 is None and torch.equal(r0, r_lo) and torch.equal(y1, y_hi)


def test_biquad_chain_low_high_mid_from_one_read():
    """ssar_biquad_chain: low / high bit-equal to the single-filter launches; mid = low_pass(high_pass(x)) within 1e-6 of
    the composed launches (its input history at chunk starts is the high-pass state, not the stored samples) and of the
    oracle; framed RMS of all three; 70 s clips (> 32 spans: the look-back crosses group boundaries), odd length, loud
    input (clamping between the chained filters); and the same bits on every run."""
    from ssar_b200 import ops, plan

    sr = 24576
    xs = (torch.stack([synth_clip(sr * 70 + 4321, sr, i) for i in (3, 8)]) * 6.0).cuda()
    lo_c, hi_c = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
    T = xs.shape[1] // 1024
    out = ops.biquad_chain(xs, lo_c, hi_c, lo_c, want_signal=(True, True, True), want_rms=(True, True, True), n_frames=T)
    (y_lo, r_lo), (y_hi, r_hi), (y_mid, r_mid) = out
    assert torch.equal(y_lo, ops.biquad(xs, lo_c)) and torch.equal(y_hi, ops.biquad(xs, hi_c))
    assert torch.equal(r_lo, ops.biquad_rms(xs, lo_c, n_frames=T)) and torch.equal(r_hi, ops.biquad_rms(xs, hi_c, n_frames=T))
    mid_ref = ops.biquad(y_hi, lo_c)
    assert rel_l2(y_mid, mid_ref) < 1e-6
    assert rel_l2(r_mid, ops.frames_rms(y_mid, 2048, 1024, T)) < 1e-6
    assert float(y_hi.abs().max()) == 1.0  # the chained filter saw a clamped signal
    for b in range(2):
        assert rel_l2(y_mid[b], dsp.mid_pass(xs[b].cpu(), sr)) < TOL
    again = ops.biquad_chain(xs, lo_c, hi_c, lo_c, want_signal=(True, True, True), want_rms=(True, True, True), n_frames=T)
    for (a, ra), (b, rb) in zip(out, again):
        assert torch.equal(a, b) and torch.equal(ra, rb)
    only_rms = ops.biquad_chain(xs, lo_c, hi_c, lo_c, want_signal=(False, False, False), want_rms=(True, True, True), n_frames=T)
    for (_, ra), (y, rb) in zip(out, only_rms):
        assert y is None and torch.equal(ra, rb)
