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


# ---------------------------------------------------------------------------------------------------------------
# against fixtures made by executing the reference's own (unimportable-module) function text: tests/golden/widened.npz
# ---------------------------------------------------------------------------------------------------------------
def test_widened_rows_vs_reference_golden():
    import random

    import numpy as np

    from ssar_b200 import stylegan2
    from ssar_b200.features import generate as gen
    from ssar_b200.optimize import FixedLatentNoiseDecoder
    from ssar_b200.supervised.latent_augmenter import LatentAugmenter
    from ssar_b200.synth import projection_matrix, synth_clip
    from test_oracle_widened import PYR_SIZES, augmenter_ws, builder_draws, load, pyr_sub

    g = load("widened")
    R = projection_matrix()
    SR = 24576
    # generate.py:106-136
    x = synth_clip(int(g["gen_n"]), SR, int(g["gen_clip_index"])).cuda()
    raw = gen.audio2features(x, SR, clamp=False, smooth=False)
    assert tuple(raw.shape) == (384, 59)
    bad = {j: rel_l2(raw[:, j], g["gen_a2f_raw"][:, j]) for j in range(59)}
    bad = {j: v for j, v in bad.items() if v > 2e-3}
    assert not bad, bad
    assert rel_l2(gen.audio2features(x, SR), g["gen_a2f"]) < 1e-3
    fb = gen.audio2features(torch.stack([x, x]), SR)
    assert tuple(fb.shape) == (2, 384, 59) and rel_l2(fb[1], g["gen_a2f"]) < 1e-3
    # generate.py:155-230
    _, _, palette = synth_patch_inputs(8, 1)
    pal = palette.cuda()
    chroma, ons = g["gen_chroma"].cuda(), g["gen_onsets"].cuda()
    Tg = chroma.shape[0]
    cases = (("high_chroma_correlation", 12, 32, lambda z: gen.high_chroma_correlation(chroma, pal[:12], 24, z)),
             ("high_onset_correlation", 2, 64, lambda z: gen.high_onset_correlation(ons, pal[:2], 24, z)),
             ("high_both_correlation", 14, 64, lambda z: gen.high_both_correlation(chroma, ons, pal[:14], 24, z)))
    for name, k, size, fn in cases:
        lat, noise = fn(builder_draws(k, Tg, size).cuda())
        assert rel_l2(lat.cpu() @ R, g[f"{name}_proj"]) < 1e-4, name
        assert rel_l2(noise[::16, :, ::4, ::4], g[f"{name}_noise"]) < 1e-4, name
    # optimize.py:103-142
    Td = 720
    xd = synth_features(Td, 27 + 8, 31)
    torch.manual_seed(0)
    draws = [torch.randn(Td, 2 ** (i + 2), 2 ** (i + 2)) for i in range(4)]
    lat, noise = FixedLatentNoiseDecoder(pal[:27], 3, 3, 3)(xd.cuda(), noise_draws=[d.cuda() for d in draws])
    assert rel_l2(lat.cpu() @ R, g["opt_proj"]) < 1e-4
    for i, nz in enumerate(noise):
        assert rel_l2(nz[:, ::2, ::2] if nz.shape[-1] > 8 else nz, g[f"opt_noise_{i}"]) < 1e-5, i
    # maua stylegan2.py:193-210, :86-96
    for tag, hw in (("16", (16, 16)), ("32", (32, 32)), ("64", (64, 64)), ("12x20", (12, 20))):
        nz = torch.randn(40, 1, *hw, generator=torch.Generator("cpu").manual_seed(77)).cuda()
        pyr = stylegan2.make_noise_pyramid(nz)
        assert list(pyr) == [f"noise{l}" for l in range(9)]
        for l, (h, w) in enumerate(PYR_SIZES):
            p = pyr[f"noise{l}"]
            assert tuple(p.shape) == (40, 1, h, w) and rel_l2(pyr_sub(p), g[f"pyr{tag}_{l}"]) < 1e-5, (tag, l)
    nz = torch.randn(6, 1, 16, 16, generator=torch.Generator("cpu").manual_seed(78)).cuda()
    assert rel_l2(stylegan2.resize_noise(nz, [(64, 64)])[0], g["resize_16_to_64"]) < 1e-5
    # supervised/latent_augmenter.py:68-105
    ws = augmenter_ws().cuda()
    feats = synth_features(384, 59, 33).cuda()
    for seed in (0, 1, 2, 3):
        la = LatentAugmenter(ws, n_patches=5 if seed < 3 else 9)
        np.random.seed(seed)
        random.seed(seed)
        res, off = la.random_patch(feats)
        assert tuple(res.shape) == (384, 18, 512) and tuple(off.shape) == (1, 1, 512)
        assert rel_l2(off, g[f"aug_off_{seed}"]) < 1e-4 and rel_l2(res.cpu() @ R, g[f"aug_res_proj_{seed}"]) < 1e-4, seed
        assert rel_l2(res[::128], g[f"aug_res_frames_{seed}"]) < 1e-4
