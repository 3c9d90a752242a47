"""GPU: the latent / noise builders of ssar/features/generate.py:155-230 and the fixed decoder of ssar/optimize.py:103-142
against the CPU oracle (builders.py: restated, unpinned — the primitives underneath are pinned).  rel-L2 <= 1e-4."""
import pytest
import torch

from conftest import rel_l2
from ssar_b200.synth import synth_features, synth_patch_inputs

pytestmark = pytest.mark.gpu
T = 720


def _inputs():
    chroma = synth_features(T, 12, 21).abs() + 0.01
    ons = synth_features(T, 1, 22).abs()
    _, _, palette = synth_patch_inputs(8, 1)
    g = torch.Generator("cpu").manual_seed(9)
    return chroma, ons, palette, g


def test_generate_builders_vs_oracle():
    from ssar_b200.features import generate as gen
    from ssar_oracle import builders

    chroma, ons, palette, g = _inputs()
    z32 = torch.randn(T, 1, 32, 32, generator=g)
    z64 = torch.randn(T, 1, 64, 64, generator=g)
    lat, nz = gen.high_chroma_correlation(chroma.cuda(), palette[:12].cuda(), 24, z32.cuda())
    rl, rn = builders.high_chroma_correlation(chroma, palette[:12], 24, z32)
    assert lat.shape == (T, 18, 512) and rel_l2(lat, rl) < 1e-4 and rel_l2(nz, rn) < 1e-4
    lat, nz = gen.high_onset_correlation(ons.cuda(), palette[:2].cuda(), 24, z64.cuda())
    rl, rn = builders.high_onset_correlation(ons, palette[:2], 24, z64)
    assert rel_l2(lat, rl) < 1e-4 and rel_l2(nz, rn) < 1e-4
    lat, nz = gen.high_both_correlation(chroma.cuda(), ons.cuda(), palette[:14].cuda(), 24, z64.cuda())
    rl, rn = builders.high_both_correlation(chroma, ons, palette[:14], 24, z64)
    assert rel_l2(lat, rl) < 1e-4 and rel_l2(nz, rn) < 1e-4


def test_optimize_fixed_decoder_vs_oracle():
    from ssar_b200.optimize import FixedLatentNoiseDecoder
    from ssar_oracle import builders

    _, _, palette, g = _inputs()
    S, G, H, n_noise = 3, 3, 3, 4
    x = synth_features(T, S * G * H + 2 * n_noise, 31)
    draws = [torch.randn(T, 2 ** (i + 2), 2 ** (i + 2), generator=g) for i in range(n_noise)]
    dec = FixedLatentNoiseDecoder(palette[:27].cuda(), S, G, H)
    lat, noise = dec(x.cuda(), noise_draws=[d.cuda() for d in draws])
    rl, rn = builders.optimize_fixed_decoder(palette[:27], x, S, G, H, draws)
    assert lat.shape == (T, 18, 512) and rel_l2(lat, rl) < 1e-4
    for a, b in zip(noise, rn):
        assert a.shape == b.shape and rel_l2(a, b) < 1e-5
    with pytest.raises(AssertionError):
        FixedLatentNoiseDecoder(palette[:20].cuda(), S, G, H)
