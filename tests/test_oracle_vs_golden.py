"""CPU: pins the oracle (oracle/ssar_oracle) against fixtures produced by the unmodified reference."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_clip
from ssar_oracle import dsp

TOL = 2e-6  # fp32 reordering only: the oracle restates the same fp32 algorithm


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_spectral_short():
    g = load("spectral_short")
    x = synth_clip(int(g["n"]), int(g["sr"]), 0)
    D = dsp.stft(x)
    assert D.shape == g["stft_re"].shape  # frame count 1 + N//hop is exact
    assert rel_l2(D.real, g["stft_re"]) < TOL and rel_l2(D.imag, g["stft_im"]) < TOL
    assert dsp.spectrogram(x, power=1).shape == g["spec_p1"].shape
    assert rel_l2(dsp.spectrogram(x, power=1), g["spec_p1"]) < TOL
    assert rel_l2(dsp.spectrogram(x, hop_length=512, power=2), g["spec_p2_hop512"]) < TOL
    R = dsp.stft(x, n_fft=1024, hop_length=256, window=None)
    assert rel_l2(R.real, g["stft_rect1024_re"]) < TOL and rel_l2(R.imag, g["stft_rect1024_im"]) < TOL
    assert torch.equal(dsp.mel_basis(int(g["sr"]), 2048).sum(1), g["mel_basis_sum"])
    assert torch.equal(dsp.mel_basis(int(g["sr"]), 2048, fmax=11025.0).sum(1), g["mel_basis_11025_sum"])


def test_frame_features():
    g = load("frame_features")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    assert rel_l2(dsp.melspectrogram(x, sr), g["mel"]) < TOL
    assert rel_l2(dsp.melspectrogram(x, sr, hop_length=512), g["mel_hop512"]) < TOL
    for key, kw in [("onset", {}), ("onset_hop512", dict(hop_length=512)), ("onset_median", dict(aggregate="median"))]:
        o = dsp.onset_strength(x, sr, **kw)
        assert o.shape == g[key].shape
        assert rel_l2(o, g[key]) < 1e-5, key
    assert rel_l2(dsp.rms(x), g["rms"]) < TOL
    assert rel_l2(dsp.rms(x, hop_length=512), g["rms_hop512"]) < TOL
    assert rel_l2(dsp.mfcc(x, sr), g["mfcc"]) < 1e-5
    assert rel_l2(dsp.spectral_flatness(x), g["flatness"]) < 1e-5


def test_frame_features_44k():
    g = load("frame_features_44k")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    assert rel_l2(dsp.spectrogram(x, 2048, 512, power=2)[::8], g["spec_p2"]) < TOL
    assert rel_l2(dsp.melspectrogram(x, sr, 2048, 512), g["mel"]) < TOL
    assert rel_l2(dsp.onset_strength(x, sr, 512, 2048), g["onset"]) < 1e-5
    assert rel_l2(dsp.rms(x, sr, 2048, 512), g["rms"]) < TOL


def test_istft_roundtrip():
    x = synth_clip(24576 + 123, 24576, 3)
    y = dsp.istft(dsp.stft(x), length=len(x))
    assert y.shape == x.shape and rel_l2(y, x) < 1e-6
    assert rel_l2(y, torch.istft(torch.stft(x, 2048, 1024, window=torch.hann_window(2048), return_complex=True),
                                 2048, 1024, window=torch.hann_window(2048), length=len(x))) < 1e-6


def test_hpss_and_filters():
    g = load("hpss_filters")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    H, P = dsp.hpss(dsp.stft(x), margin=8.0)
    assert rel_l2(H.abs()[::4], g["hpss_h_abs"]) < TOL and rel_l2(P.abs()[::4], g["hpss_p_abs"]) < TOL
    assert rel_l2(dsp.harmonic(x), g["harmonic"]) < 5e-6
    assert rel_l2(dsp.percussive(x), g["percussive"]) < 5e-6
    assert rel_l2(dsp.low_pass(x, sr), g["low"]) < 1e-5
    assert rel_l2(dsp.high_pass(x, sr), g["high"]) < 1e-5
    assert rel_l2(dsp.mid_pass(x, sr), g["mid"]) < 1e-5


def test_resample():
    g = load("resample")
    x = synth_clip(int(g["n"]), 44100, int(g["clip_index"]))
    y = dsp.resample(x, 44100, 24576)
    assert y.shape == g["to24576"].shape  # ceil(2048 * N / 3675): exact length
    assert rel_l2(y, g["to24576"]) < TOL
    y2 = dsp.resample(g["to24576"], 24576, 12288, method="kaiser")
    assert y2.shape == g["half_kaiser"].shape and rel_l2(y2, g["half_kaiser"]) < TOL
