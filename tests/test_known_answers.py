"""Known-answer cases SURVEY 8(c) asks for: DC, white noise, a click train (onset timing / pad_width).  Each fact is
checked on the CPU oracle (always) and on the CUDA path (`-m gpu`), so neither depends on the other for these."""
import math

import pytest
import torch

from conftest import rel_l2

SR = 24576


def click_train(seconds=10, every=12288):
    n = SR * seconds
    x = torch.zeros(n)
    clicks = list(range(every, n - 2048, every))
    x[clicks] = 0.8
    x += 1e-4 * torch.randn(n, generator=torch.Generator().manual_seed(0))  # never exactly silent
    return x, [c // 1024 for c in clicks]


def peaks_above_half_max(on):
    on = on.flatten().cpu()
    return [i for i in range(1, len(on) - 1) if on[i] > on[i - 1] and on[i] >= on[i + 1] and on[i] > 0.5 * on.max()]


def check_dc(spec, rms_v):
    # periodic Hann of 2048 points: sum = 1024, first harmonic -512 -> |X0| = c * 1024, |X1| = c * 512, nothing above
    c = 0.5
    assert torch.allclose(spec[0], torch.full_like(spec[0], c * 1024), rtol=1e-5)
    assert torch.allclose(spec[1], torch.full_like(spec[1], c * 512), rtol=1e-5)
    assert float(spec[2:].abs().max()) < 1e-3
    assert torch.allclose(rms_v.flatten(), torch.full_like(rms_v.flatten(), c), rtol=1e-6)


def check_flatness_white(fl):
    # power spectrum of Gaussian noise: exponentially distributed bins, geometric / arithmetic mean -> exp(-gamma)
    assert abs(float(fl.mean()) - math.exp(-0.5772156649)) < 0.01 and float(fl.std()) < 0.04


def test_oracle_known_answers():
    from ssar_oracle import dsp

    x, frames = click_train()
    on = dsp.onset_strength(x, SR)
    # a click on a frame centre fills exactly that frame (Hann is zero one hop away); lag 1 + pad_width 2 - 1 (diff)
    assert len(on) == len(x) // 1024 and peaks_above_half_max(on) == [f + 1 for f in frames]
    dc = torch.full((SR * 2,), 0.5)
    check_dc(dsp.spectrogram(dc, 2048, 1024, power=1), dsp.rms(dc, SR))
    g = 0.1 * torch.randn(SR * 20, generator=torch.Generator().manual_seed(1))
    check_flatness_white(dsp.spectral_flatness(g))


@pytest.mark.gpu
def test_gpu_known_answers():
    from ssar_b200.features import audio
    from ssar_b200.features.rosa import beat, spectral
    from ssar_oracle import dsp

    x, frames = click_train()
    on = beat.onset_strength(x.cuda(), SR)
    assert on.numel() == len(x) // 1024 and peaks_above_half_max(on) == [f + 1 for f in frames]
    assert rel_l2(on, dsp.onset_strength(x, SR)) < 1e-3
    dc = torch.full((SR * 2,), 0.5)
    check_dc(spectral.spectrogram(dc.cuda(), 2048, 1024, power=1).cpu(), audio.rms(dc.cuda(), SR).cpu())
    g = 0.1 * torch.randn(SR * 20, generator=torch.Generator().manual_seed(1))
    fl = audio.spectral_flatness(g.cuda(), SR).cpu()
    check_flatness_white(fl)
    assert rel_l2(fl, dsp.spectral_flatness(g)) < 1e-4
