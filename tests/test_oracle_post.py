"""CPU: pins oracle/ssar_oracle/post.py against fixtures produced by the unmodified reference (make_golden.py post)."""
import contextlib
import io

import numpy as np
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_features
from ssar_oracle import post

TOL = 2e-6


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def inputs():
    return (synth_features(700, 7, 0), synth_features(64, 6, 1).reshape(64, 2, 3),
            synth_features(48, 8, 2).reshape(48, 2, 2, 2), synth_features(10, 3, 3))


def test_gaussian_filter_shapes_and_values():
    g = load("post")
    x, x3, x4, short = inputs()
    cases = {
        "gauss_s2": post.gaussian_filter(x, 2.4), "gauss_s10_reflect": post.gaussian_filter(x, 10, "reflect"),
        "gauss_1d": post.gaussian_filter(x[:, 0], 3), "gauss_col1": post.gaussian_filter(x[:, :1], 3),
        "gauss_3d": post.gaussian_filter(x3, 1), "gauss_4d": post.gaussian_filter(x4, 2),
        "gauss_short": post.gaussian_filter(short, 5.0),
    }
    for k, v in cases.items():
        assert v.shape == g[k].shape, k
        assert rel_l2(v, g[k]) < TOL, k


def test_normalise_quantiles_clamps():
    g = load("post")
    x = inputs()[0]
    assert rel_l2(post.normalize(x), g["normalize"]) < TOL
    assert rel_l2(post.standardize(x), g["standardize"]) < TOL
    assert float(post.quantile_midpoint(x, 0.25)) == float(g["q25"])
    assert float(post.quantile_midpoint(x[:, 3], 0.975)) == float(g["q975"])
    assert torch.equal(post.spectral_flux(x), g["flux"])
    assert rel_l2(post.onset_envelope(post.spectral_flux(x)), g["onset_env"]) < TOL
    for k, v in {
        "clamp_peaks": post.clamp_peaks_percentile(x, 97.5), "clamp_peaks_1d": post.clamp_peaks_percentile(x[:, 2], 90),
        "clamp_upper": post.clamp_upper_percentile(x, 95), "clamp_lower": post.clamp_lower_percentile(x, 10),
    }.items():
        assert v.shape == g[k].shape and torch.equal(v, g[k]), k  # order statistics: bit-exact
    assert rel_l2(post.emphasize(x, 2, 75), g["emphasize"]) < TOL
    assert rel_l2(post.emphasize(x[:, :1], 10, 50), g["emphasize_10_50"]) < TOL


def test_overlapping_slices_and_window_stats():
    """supervised/data.py:255-258, 317-319 against the reference's own function (tests/golden/slices.npz)."""
    import numpy as np
    from ssar_b200.synth import synth_features

    g = {k: np.asarray(v) for k, v in np.load(GOLDEN / "slices.npz").items()}
    for T, L in ((1000, 192), (768, 192), (777, 96), (5760, 192)):
        x = synth_features(T, 59, 50 + T % 7)
        w = post.overlapping_slices(x, L)
        assert w.shape == (int(g[f"n_{T}_{L}"]), L, 59)
        for k, r in enumerate(g[f"first_row_{T}_{L}"]):
            assert torch.equal(w[k], x[int(r) : int(r) + L])
    tracks = [synth_features(T, 59, 60 + i) for i, T in enumerate((1000, 2111, 777))]
    allw = torch.cat([post.overlapping_slices(f, 192) for f in tracks]).numpy()
    assert np.allclose(np.mean(allw, axis=(0, 1)), g["win_mean"], rtol=1e-6) and np.allclose(np.std(allw, axis=(0, 1)), g["win_std"], rtol=1e-6)


def test_video_twins():
    """video.py:11-31, 58-73 restated in oracle/video.py against the reference's own function text."""
    from video_synth import synth_video
    from ssar_oracle import video as ovid

    g = load("video")
    for tag, kw in (("a", dict()), ("b", dict(T=129, H=33, W=65, seed=6)), ("c", dict(T=2, H=16, W=16, seed=7))):
        v = synth_video(**kw)
        assert torch.equal(ovid.channel_hist(v, 0), g[f"red_{tag}"]) and torch.equal(ovid.channel_hist(v, 1, 20), g[f"green_{tag}"])
        assert torch.equal(ovid.rgb_hist(v), g[f"rgb_{tag}"])
        assert rel_l2(ovid.visual_variance(v), g[f"var_{tag}"]) < 1e-6 and rel_l2(ovid.absdiff(v), g[f"absdiff_{tag}"]) < 1e-6
