"""CPU: pins the oracle's restatements of functions whose reference MODULES cannot be imported — generate.py's
audio2features variant and builders, optimize.py's decoder, LatentAugmenter.random_patch, make_noise_pyramid — against
fixtures made by executing the reference's unmodified function text (tests/golden/make_golden.py::g_widened)."""
import random

import numpy as np
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import projection_matrix, synth_clip, synth_features, synth_patch_inputs

SR = 24576


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def pyr_sub(t):
    return t[::8] if t.shape[-1] <= 32 else t[::8, :, ::2, ::2]


PYR_SIZES = [(s, s) for s in (8, 8, 16, 16, 32, 32, 64, 64, 128)]  # synthesis layers 1..9 of a 1024 px StyleGAN2


def augmenter_ws():
    return torch.randn(256, 18, 512, generator=torch.Generator("cpu").manual_seed(4242)).repeat(64, 1, 1)


def builder_draws(k, T, size):
    torch.manual_seed(100 + k)
    torch.randn(k, 512)  # the mapper's z (generate.py:166)
    return torch.randn(T, 1, size, size)


def test_generate_audio2features_variant():
    from ssar_oracle import features

    g = load("widened")
    x = synth_clip(int(g["gen_n"]), SR, int(g["gen_clip_index"]))
    raw = features.audio2features_generate(x, SR, clamp=False, smooth=False)
    assert raw.shape == g["gen_a2f_raw"].shape == (384, 59)
    for j in range(59):
        assert rel_l2(raw[:, j], g["gen_a2f_raw"][:, j]) < 1e-4, j
    assert rel_l2(features.audio2features_generate(x, SR), g["gen_a2f"]) < 1e-5


def test_builders_and_optimize_decoder():
    from ssar_oracle import builders

    g = load("widened")
    R = projection_matrix()
    _, _, palette = synth_patch_inputs(8, 1)
    chroma, ons = g["gen_chroma"], g["gen_onsets"]
    T = chroma.shape[0]
    cases = (("high_chroma_correlation", 12, 32, lambda z: builders.high_chroma_correlation(chroma, palette[:12], 24, z)),
             ("high_onset_correlation", 2, 64, lambda z: builders.high_onset_correlation(ons, palette[:2], 24, z)),
             ("high_both_correlation", 14, 64, lambda z: builders.high_both_correlation(chroma, ons, palette[:14], 24, z)))
    for name, k, size, fn in cases:
        lat, noise = fn(builder_draws(k, T, size))
        assert rel_l2(lat @ R, g[f"{name}_proj"]) < 1e-5, name
        assert rel_l2(noise[::16, :, ::4, ::4], g[f"{name}_noise"]) < 1e-5, name
        assert rel_l2(noise.std(dim=(1, 2, 3)), g[f"{name}_noise_std"]) < 1e-5, name
    T = 720
    x = synth_features(T, 27 + 8, 31)
    torch.manual_seed(0)
    draws = [torch.randn(T, 2 ** (i + 2), 2 ** (i + 2)) for i in range(4)]
    lat, noise = builders.optimize_fixed_decoder(palette[:27], x, 3, 3, 3, draws)
    assert rel_l2(lat @ R, g["opt_proj"]) < 1e-5
    for i, nz in enumerate(noise):
        assert rel_l2(nz[:, ::2, ::2] if nz.shape[-1] > 8 else nz, g[f"opt_noise_{i}"]) < 1e-5, i


def test_noise_pyramid():
    from ssar_oracle import maua

    g = load("widened")
    for tag, hw in (("16", (16, 16)), ("32", (32, 32)), ("64", (64, 64)), ("12x20", (12, 20))):
        nz = torch.randn(40, 1, *hw, generator=torch.Generator("cpu").manual_seed(77))
        for l, p in enumerate(maua.make_noise_pyramid(nz, PYR_SIZES)):
            assert rel_l2(pyr_sub(p), g[f"pyr{tag}_{l}"]) < 1e-6, (tag, l)


def test_latent_augmenter():
    from ssar_oracle import traj

    g = load("widened")
    R = projection_matrix()
    ws = augmenter_ws()
    feats = synth_features(384, 59, 33)
    for seed in (0, 1, 2, 3):
        la = traj.LatentAugmenter(ws, n_patches=5 if seed < 3 else 9)
        np.random.seed(seed)
        random.seed(seed)
        res, off = la.random_patch(feats.clone())
        assert rel_l2(off, g[f"aug_off_{seed}"]) < 1e-5 and rel_l2(res @ R, g[f"aug_res_proj_{seed}"]) < 1e-5, seed
        assert rel_l2(res[::128], g[f"aug_res_frames_{seed}"]) < 1e-5
