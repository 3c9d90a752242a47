"""GPU parity: post-processing kernels (csrc/post.cu) vs the oracle and the reference-generated goldens."""
import contextlib
import io

import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_features
from ssar_oracle import post

pytestmark = pytest.mark.gpu
TOL = 1e-5  # fp32 elementwise / short sums


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def test_post_vs_golden():
    from ssar_b200.features import processing as P

    g = load("post")
    x = synth_features(700, 7, 0).cuda()
    x3 = synth_features(64, 6, 1).reshape(64, 2, 3).cuda()
    x4 = synth_features(48, 8, 2).reshape(48, 2, 2, 2).cuda()
    short = synth_features(10, 3, 3).cuda()
    cases = {
        "gauss_s2": P.gaussian_filter(x, 2.4), "gauss_s10_reflect": P.gaussian_filter(x, 10, mode="reflect"),
        "gauss_1d": P.gaussian_filter(x[:, 0], 3), "gauss_col1": P.gaussian_filter(x[:, :1], 3),
        "gauss_3d": P.gaussian_filter(x3, 1), "gauss_4d": P.gaussian_filter(x4, 2),
        "gauss_short": quiet(P.gaussian_filter, short, 5.0),
        "normalize": P.normalize(x), "standardize": P.standardize(x),
        "flux": P.spectral_flux(x), "onset_env": P.onset_envelope(P.spectral_flux(x)),
        "emphasize": P.emphasize(x, 2, 75), "emphasize_10_50": P.emphasize(x[:, :1], 10, 50),
    }
    for k, v in cases.items():
        assert tuple(v.shape) == tuple(g[k].shape), k
        assert rel_l2(v, g[k]) < TOL, k
    # order statistics and clamps are selections of input values: bit-exact
    assert float(P.quantile(x, 0.25)) == float(g["q25"])
    assert float(P.quantile(x[:, 3], 0.975)) == float(g["q975"])
    for k, v in {
        "clamp_peaks": P.clamp_peaks_percentile(x, 97.5), "clamp_peaks_1d": P.clamp_peaks_percentile(x[:, 2], 90),
        "clamp_upper": P.clamp_upper_percentile(x, 95), "clamp_lower": P.clamp_lower_percentile(x, 10),
    }.items():
        assert tuple(v.shape) == tuple(g[k].shape), k
        assert rel_l2(v, g[k]) < 1e-7, k


def test_post_batched_vs_oracle_full_size():
    """cfg24 size: [B, 5760, 59] batched clamps + smoothing equal the per-clip oracle; stats equal numpy."""
    from ssar_b200 import ops

    B, T, C = 3, 5760, 59
    xs = torch.stack([synth_features(T, C, 10 + i) for i in range(B)])
    xd = xs.cuda()
    thr, cnt = ops.col_quantile(xd, 0.975, "peaks", "linear", want_count=True)
    assert int(cnt.min()) > 0
    y = ops.clamp_cols(xd, None, thr)
    y = ops.clamp_cols(y, ops.col_quantile(y, 0.10), None)
    z = ops.gauss1d(y, 2.4, "circular", batched=True)
    for b in range(B):
        ref = post.clamp_lower_percentile(post.clamp_peaks_percentile(xs[b], 97.5), 10)
        assert torch.equal(y[b].cpu(), ref)
        assert rel_l2(z[b], post.gaussian_filter(ref, 2.4)) < TOL
    st = ops.feature_stats(z).cpu().numpy()
    mean, std = post.feature_stats(z.cpu())
    assert np.all(st[0] == B * T)
    np.testing.assert_allclose(st[1] / st[0], mean, rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(np.sqrt(st[2] / st[0] - (st[1] / st[0]) ** 2), std, rtol=1e-7, atol=1e-9)


def test_clamp_peaks_raises_without_peaks():
    from ssar_b200.features import processing as P

    with pytest.raises(RuntimeError):
        P.clamp_peaks_percentile(torch.arange(100, dtype=torch.float32).cuda(), 97.5)


def test_overlapping_slices_and_window_stats_vs_golden():
    """supervised/data.py:255-258 (one gather launch) and :317-319 (mean / std over all windows: frames count once per
    window that holds them) against the reference's own outputs."""
    import numpy as np
    from ssar_b200.supervised import data
    from ssar_b200.synth import synth_features

    g = {k: np.asarray(v) for k, v in np.load(GOLDEN / "slices.npz").items()}
    for T, L in ((1000, 192), (768, 192), (777, 96), (5760, 192)):
        x = synth_features(T, 59, 50 + T % 7)
        w = data.overlapping_slices(x.cuda(), L).cpu()
        assert tuple(w.shape) == (int(g[f"n_{T}_{L}"]), L, 59)
        for k, r in enumerate(g[f"first_row_{T}_{L}"]):
            assert torch.equal(w[k], x[int(r) : int(r) + L])
    lat = torch.randn(700, 18, 512, generator=torch.Generator().manual_seed(2)).cuda()
    wl = data.overlapping_slices(lat, 192)
    assert tuple(wl.shape) == (11, 192, 18, 512)  # starts 0 / 48 / 96 / 144 -> 3 + 3 + 3 + 2 windows
    assert torch.equal(wl[5], lat[48 + 384 : 48 + 576]) and torch.equal(wl[10], lat[144 + 192 : 144 + 384])
    tracks = [synth_features(T, 59, 60 + i).cuda() for i, T in enumerate((1000, 2111, 777))]
    mean, std = data.window_mean_std(tracks, 192)
    assert np.allclose(mean.cpu().numpy(), g["win_mean"], rtol=1e-5, atol=1e-5) and np.allclose(std.cpu().numpy(), g["win_std"], rtol=1e-5)  # (the golden is numpy fp32 sums)
    with pytest.raises(RuntimeError):
        data.overlapping_slices(torch.zeros(191, 59).cuda(), 192)


def test_video_twins_vs_golden():
    """features/video.py:11-31 (histogram COUNTS are integers: exact), :58-59, :62-73 against the reference's outputs."""
    from video_synth import synth_video
    from ssar_b200.features import video as vid

    g = load("video")
    for tag, kw in (("a", dict()), ("b", dict(T=129, H=33, W=65, seed=6)), ("c", dict(T=2, H=16, W=16, seed=7))):
        v = synth_video(**kw).cuda()
        assert torch.equal(vid.redogram(v).cpu(), g[f"red_{tag}"]), tag
        assert torch.equal(vid.greenogram(v, 20).cpu(), g[f"green_{tag}"]) and torch.equal(vid.blueogram(v).cpu(), g[f"blue_{tag}"])
        assert torch.equal(vid.rgb_hist(v).cpu(), g[f"rgb_{tag}"])
        assert rel_l2(vid.visual_variance(v), g[f"var_{tag}"]) < 1e-5 and rel_l2(vid.absdiff(v), g[f"absdiff_{tag}"]) < 1e-5
        assert tuple(vid.absdiff(v).shape) == (v.shape[0], 1)
