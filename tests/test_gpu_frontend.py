"""GPU parity: fused frame-FFT kernel family vs the CPU oracle and the reference-generated goldens."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_clip
from ssar_oracle import dsp

pytestmark = pytest.mark.gpu

TOL_SPEC = 1e-4   # north-star tolerance on spectra / mel (rel-L2, fp32)
TOL_ONSET = 1e-3  # north-star tolerance on onset


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_stft_and_spectrogram_vs_golden():
    from ssar_b200.features.rosa import spectral

    g = load("spectral_short")
    x = synth_clip(int(g["n"]), int(g["sr"]), 0).cuda()
    D = spectral.stft(x)
    assert tuple(D.shape) == tuple(g["stft_re"].shape)  # [1025, 1 + N//hop]: frame indexing exact
    assert rel_l2(D.real, g["stft_re"]) < TOL_SPEC and rel_l2(D.imag, g["stft_im"]) < TOL_SPEC
    S1 = spectral.spectrogram(x, power=1)
    assert tuple(S1.shape) == tuple(g["spec_p1"].shape)
    assert rel_l2(S1, g["spec_p1"]) < TOL_SPEC
    assert rel_l2(spectral.spectrogram(x, hop_length=512, power=2), g["spec_p2_hop512"]) < TOL_SPEC
    R = spectral.stft(x, n_fft=1024, hop_length=256, window=None)
    assert tuple(R.shape) == tuple(g["stft_rect1024_re"].shape)
    assert rel_l2(R.real, g["stft_rect1024_re"]) < TOL_SPEC and rel_l2(R.imag, g["stft_rect1024_im"]) < TOL_SPEC


def test_frame_features_vs_golden():
    from ssar_b200.features import audio
    from ssar_b200.features.rosa import beat, spectral

    g = load("frame_features")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"])).cuda()
    assert rel_l2(spectral.melspectrogram(x, sr), g["mel"]) < TOL_SPEC
    assert rel_l2(spectral.melspectrogram(x, sr, hop_length=512), g["mel_hop512"]) < TOL_SPEC
    o = beat.onset_strength(x, sr)
    assert tuple(o.shape) == tuple(g["onset"].shape)
    assert rel_l2(o, g["onset"]) < TOL_ONSET
    assert torch.all(o[:2] == 0)  # pad_width = 2 at hop 1024: bit-exact frame offset
    assert rel_l2(beat.onset_strength(x, sr, hop_length=512), g["onset_hop512"]) < TOL_ONSET
    med = lambda *a, **k: torch.median(*a, **k).values
    assert rel_l2(beat.onset_strength(x, sr, aggregate=med), g["onset_median"]) < TOL_ONSET
    r = audio.rms(x, sr)
    assert tuple(r.shape) == tuple(g["rms"].shape)
    assert rel_l2(r, g["rms"]) < TOL_SPEC
    assert rel_l2(audio.rms(x, sr, hop_length=512), g["rms_hop512"]) < TOL_SPEC
    assert rel_l2(audio.mfcc(x, sr), g["mfcc"]) < TOL_ONSET
    assert rel_l2(audio.spectral_flatness(x, sr), g["flatness"]) < TOL_ONSET


def test_config2_batched_vs_oracle_and_golden():
    """BASELINE config 2 shape (44.1 kHz, n_fft 2048, hop 512), batched, ragged clip length."""
    from ssar_b200.pipeline import spectral_frontend

    g = load("frame_features_44k")
    sr, n = int(g["sr"]), int(g["n"])
    xs = torch.stack([synth_clip(n, sr, i) for i in (2, 5, 9)])
    out = spectral_frontend(xs.cuda(), sr, 2048, 512)
    T = n // 512
    assert tuple(out["spec"].shape) == (3, T, 1025) and tuple(out["mel"].shape) == (3, T, 128)
    assert tuple(out["onset"].shape) == (3, T) and tuple(out["rms"].shape) == (3, T)
    # clip 0 of the batch is the golden clip
    assert rel_l2(out["spec"][0].T[::8], g["spec_p2"]) < TOL_SPEC
    assert rel_l2(out["mel"][0].T, g["mel"]) < TOL_SPEC
    assert rel_l2(out["onset"][0], g["onset"]) < TOL_ONSET
    assert rel_l2(out["rms"][0], g["rms"][:, 0]) < TOL_SPEC
    for b in range(3):
        x = xs[b]
        assert rel_l2(out["spec"][b].T, dsp.spectrogram(x, 2048, 512, power=2)) < TOL_SPEC
        assert rel_l2(out["mel"][b].T, dsp.melspectrogram(x, sr, 2048, 512)) < TOL_SPEC
        assert rel_l2(out["onset"][b], dsp.onset_strength(x, sr, 512, 2048)) < TOL_ONSET
        assert rel_l2(out["rms"][b], dsp.rms(x, sr, 2048, 512)[:, 0]) < TOL_SPEC
        assert torch.all(out["onset"][b, :3] == 0)  # pad_width = 3 at hop 512


@pytest.mark.parametrize("n", [2048 + 1, 4096, 4096 + 511, 4096 + 1023, 5 * 1024])
def test_frame_count_edges(n):
    from ssar_b200.features.rosa import spectral

    x = synth_clip(n, 24576, 7)
    D = spectral.stft(x.cuda())
    R = dsp.stft(x)
    assert tuple(D.shape) == tuple(R.shape) == (1025, 1 + n // 1024)
    assert rel_l2(torch.view_as_real(D.cpu().contiguous()), torch.view_as_real(R.contiguous())) < TOL_SPEC


def test_known_answers():
    """impulse -> flat spectrum in the frame that contains it; bin-centred tone -> energy in that bin; linearity."""
    from ssar_b200.features.rosa import spectral

    n, sr = 16384, 24576
    x = torch.zeros(n)
    x[5000] = 1.0
    S = spectral.spectrogram(x.cuda(), power=1).cpu()  # [1025, 16]
    w = torch.hann_window(2048)
    for t in range(16):
        pos = 5000 - (t * 1024 - 1024)
        expect = float(w[pos]) if 0 <= pos < 2048 else 0.0
        assert torch.allclose(S[:, t], torch.full((1025,), expect), atol=2e-6), t
    k = 100
    tone = torch.sin(2 * torch.pi * (k * sr / 2048) * torch.arange(n) / sr)
    St = spectral.spectrogram(tone.cuda(), power=2).cpu()
    assert int(St[:, 5].argmax()) == k
    a, b = synth_clip(n, sr, 1).cuda(), synth_clip(n, sr, 2).cuda()
    lhs = spectral.stft(2.0 * a - 0.5 * b)
    rhs = 2.0 * spectral.stft(a) - 0.5 * spectral.stft(b)
    assert rel_l2(torch.view_as_real(lhs.contiguous()), torch.view_as_real(rhs.contiguous())) < 1e-5


def test_parseval_full_size():
    """Size-independent property at the BASELINE config-2 clip length: windowed-frame energy == spectrum energy."""
    from ssar_b200 import ops

    sr, n = 44100, 44100 * 240
    from ssar_b200.synth import synth_batch_device

    x = synth_batch_device(2, n, sr, "cuda")
    T = n // 512
    out = ops.frames_fft(x, 2048, 512, T, torch.hann_window, spec_power=2, want_rms=True)
    spec = out["spec"]
    assert tuple(spec.shape) == (2, T, 1025)
    # Parseval for a real frame: sum_n (x w)^2 = (|X0|^2 + 2 sum_{1..N/2-1} |Xk|^2 + |X_{N/2}|^2) / N
    lhs = (spec[..., 0] + 2 * spec[..., 1:-1].sum(-1) + spec[..., -1]) / 2048
    t_idx = torch.tensor([10, 1000, T // 2, T - 10], device="cuda")
    w = torch.hann_window(2048, device="cuda")
    for t in t_idx.tolist():
        seg = x[:, t * 512 - 1024 : t * 512 + 1024] * w
        assert torch.allclose(lhs[:, t], (seg * seg).sum(-1), rtol=2e-4)
        assert torch.allclose(out["rms"][:, t], x[:, t * 512 - 1024 : t * 512 + 1024].pow(2).mean(-1).sqrt(), rtol=1e-4)


def test_lane_interleaved_banks_equal_packed_banks():
    """The filterbank epilogue streams lane-interleaved, zero-padded weights (plan.lane_banks); dropping that form sends
    the same launch down the packed per-band loop.  Zero-weight terms are fmaf(0, finite, acc): bit-equal, for both
    transform sizes, two banks per launch, band counts that are not a multiple of 32, and with the running max."""
    import dataclasses

    from ssar_b200 import ops, plan

    x = torch.stack([synth_clip(22050 * 3, 22050, i) for i in (1, 5)]).cuda()
    for n_fft, hop, n_mels in ((2048, 512, 128), (1024, 256, 40)):
        a = plan.mel_bank(22050.0, n_fft, n_mels, 0.0, None, "cuda")
        b = plan.mel_bank(22050.0, n_fft, 20, 300.0, 8000.0, "cuda")
        assert a.lanes is not None and b.lanes is not None
        n_frames = 1 + x.shape[1] // hop
        fast = ops.frames_fft(x, n_fft, hop, n_frames, banks=((a, True, True), (b, True, False)))
        packed = [dataclasses.replace(k, lanes=None) for k in (a, b)]
        slow = ops.frames_fft(x, n_fft, hop, n_frames, banks=((packed[0], True, True), (packed[1], True, False)))
        for key in ("bank0", "bank1", "bank0_max"):
            assert torch.equal(fast[key], slow[key]), (n_fft, key)


def test_time_segments_with_halos_equal_whole_track():
    """Long tracks cut into time segments with n_fft / 2 halos (frames_fft `segment`, pipeline.spectral_frontend_segmented):
    every output equals the whole-track call bit for bit — device-resident and pinned-host waveforms, segment lengths that
    do and do not divide the frame count, a ragged track length."""
    from ssar_b200 import pipeline

    sr = 44100
    xs = torch.stack([synth_clip(sr * 21 + 311, sr, i) for i in (1, 4)])
    whole = pipeline.spectral_frontend(xs.cuda(), sr, 2048, 512)
    for seg_frames, src in ((256, xs.cuda()), (401, xs.pin_memory()), (5000, xs.cuda())):
        seg = pipeline.spectral_frontend_segmented(src, sr, 2048, 512, seg_frames=seg_frames)
        torch.cuda.synchronize()
        for k in ("spec", "mel", "onset", "rms"):
            assert torch.equal(seg[k], whole[k]), (seg_frames, k, rel_l2(seg[k], whole[k]))


def test_sub_every_hop512_launch_serves_hop1024_consumers():
    """A transform at hop 512 with sub_every = 2 hands out the even frames' complex spectrum and RMS: bit-equal to the hop 1024
    launch (rosa/spectral.py:10-21, features/audio.py:31-37), and ops.estimate_tuning's by-products are those tensors while
    its tuning is unchanged."""
    from ssar_b200 import ops

    sr, n = 24576, 24576 * 11
    xs = torch.stack([synth_clip(n, sr, i) for i in (2, 5, 7)]).cuda()
    T = n // 1024
    ref = ops.frames_fft(xs, 2048, 1024, T + 1, torch.hann_window, want_complex=True, want_rms=True, aux_frames=T)
    got = ops.frames_fft(xs, 2048, 512, 2 * T + 1, torch.hann_window, want_complex=True, want_rms=True, aux_frames=2 * T,
                         sub_every=2)
    assert got["complex"].shape == ref["complex"].shape and torch.equal(got["complex"], ref["complex"])
    assert got["rms"].shape == ref["rms"].shape and torch.equal(got["rms"], ref["rms"])
    by = {}
    i1, v1 = ops.estimate_tuning(xs, sr, bins_per_octave=36, stft_out=by)
    i0, v0 = ops.estimate_tuning(xs, sr, bins_per_octave=36)
    assert torch.equal(i0, i1) and torch.equal(v0, v1)
    assert torch.equal(by["complex"], ref["complex"]) and torch.equal(by["rms"], ref["rms"])
