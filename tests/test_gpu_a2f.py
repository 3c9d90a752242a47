"""GPU parity: plp / onsets / drop_strength and the full audio2features [T, 59] vs reference-generated goldens."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_clip

pytestmark = pytest.mark.gpu


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_plp_onsets_drop_vs_golden():
    from ssar_b200.features import audio
    from ssar_b200.features.rosa import beat

    g = load("a2f")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"])).cuda()
    p = beat.plp(x, sr)
    assert tuple(p.shape) == tuple(g["plp"].shape)
    assert rel_l2(p, g["plp"]) < 1e-3
    assert rel_l2(audio.pulse(x, sr), g["pulse"]) < 1e-3
    assert rel_l2(audio.onsets(x, sr), g["onsets"]) < 1e-3
    d = audio.drop_strength(x, sr)
    assert tuple(d.shape) == tuple(g["drop_strength"].shape) and rel_l2(d, g["drop_strength"]) < 1e-4


def test_plp_exact_given_envelope():
    """The PLP kernels reproduce the oracle to 1e-5 when fed the oracle's own onset envelope."""
    from ssar_b200 import ops
    from ssar_oracle import dsp

    g = load("a2f")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    env = dsp.onset_strength(x, sr, aggregate="median")
    assert rel_l2(ops.plp_from_onset(env.cuda()[None], sr)[0], dsp.plp(x, sr)) < 1e-5


# per-column tolerance on the un-clamped, un-smoothed features: spectra-derived columns 1e-3 (north star: onset /
# chroma class), everything measured as rel-L2 per column against the unmodified reference's output
def test_audio2features_vs_golden():
    from ssar_b200.supervised.data import FEATURE_NAMES, audio2features

    g = load("a2f")
    x = synth_clip(int(g["n44"]), 44100, int(g["clip_index44"])).cuda()
    raw = audio2features(x, 44100, 24, clamp=False, smooth=False)
    assert tuple(raw.shape) == (1080, 59) and len(FEATURE_NAMES) == 59
    worst = {}
    for j, name in enumerate(FEATURE_NAMES):
        worst[name] = rel_l2(raw[:, j], g["features_raw"][:, j])
    # pulse: plp keeps ONE arg-max tempogram bin per frame; a near-tie flips under the 1e-5 fp32 noise of the onset
    # envelope (reference noise floor), which moves the pulse by O(1e-3).  The kernel itself is exact given the same
    # envelope (test_plp_exact_given_envelope, 1e-5); through the chain the column is held to 1e-2.
    tol = {name: 2e-3 for name in FEATURE_NAMES}
    tol["pulse"] = 1e-2
    bad = {k: v for k, v in worst.items() if v > tol[k]}
    assert not bad, bad
    f = audio2features(x, 44100, 24)
    assert tuple(f.shape) == (1080, 59)
    assert rel_l2(f, g["features"]) < 1e-3
    keep = [j for j, n in enumerate(FEATURE_NAMES) if n != "pulse"]
    assert rel_l2(f[:, keep], g["features"][:, keep]) < 1e-4
    # stereo input is averaged to mono first (supervised/data.py:61-62)
    f2 = audio2features(torch.stack([x, x]), 44100, 24)
    assert rel_l2(f2, f) < 1e-6


def test_audio2features_batch_matches_single():
    from ssar_b200 import pipeline

    xs = torch.stack([synth_clip(44100 * 45, 44100, i) for i in (3, 9)]).cuda()
    fb = pipeline.audio2features_batch(xs, 44100, 24)
    for b in range(2):
        fs = pipeline.audio2features_batch(xs[b : b + 1], 44100, 24)[0]
        assert rel_l2(fb[b], fs) < 1e-6


def test_feature_graph_replay_equals_eager():
    """The CUDA-graph replay of the feature path (pipeline.FeatureGraph: device-side choice of each clip's tuned CQT
    basis, no read-back) gives the eager call's features bit for bit, replay after replay with new inputs."""
    from ssar_b200 import pipeline

    g = pipeline.FeatureGraph.get(2, 44100 * 45, 44100, 24, "cuda:0")
    for ids in ((3, 9), (5, 1)):
        xs = torch.stack([synth_clip(44100 * 45, 44100, i) for i in ids]).cuda()
        checks = []
        f, pf = g.run(xs, deferred_checks=checks)
        for c in checks:
            c()
        fe = pipeline.audio2features_batch(xs, 44100, 24)
        assert torch.equal(f, fe)
        pe = pipeline.patch_features_batch(fe)
        assert all(torch.equal(pf[k], pe[k]) for k in pe)


def test_features_and_latents_graph_vs_serial():
    """The threaded pipeline (feature graphs, batched sampler, side streams) == the serially queued one."""
    from ssar_b200 import pipeline
    from ssar_b200.synth import synth_patch_inputs

    xs = torch.stack([synth_clip(44100 * 45, 44100, i) for i in (3, 9, 4)]).cuda()
    _, segs, palette = synth_patch_inputs(1080, 0)
    palette = palette.cuda()
    f0, l0, _ = pipeline.features_and_latents(xs, 44100, 24, palette, segs, chunk=2)  # first sight of the shape: eager
    fa, la, _ = pipeline.features_and_latents(xs, 44100, 24, palette, segs, chunk=2)  # recurring shape: graph replays
    assert any(k[0] == 2 and k[1] == xs.shape[1] for k in pipeline.FeatureGraph._cache)
    fb, lb, _ = pipeline.features_and_latents(xs, 44100, 24, palette, segs, chunk=2, serial=True)
    torch.cuda.synchronize()
    assert torch.equal(fa, fb) and torch.equal(la, lb) and torch.equal(f0, fb) and torch.equal(l0, lb)
    # pinned host input: staged chunk by chunk
    fc, lc, _ = pipeline.features_and_latents(xs.cpu().pin_memory(), 44100, 24, palette, segs, chunk=2)
    torch.cuda.synchronize()
    assert torch.equal(fa, fc) and torch.equal(la, lc)
    # again, with other audio: the two staging buffers are reused across chunks and across calls
    ys = torch.stack([synth_clip(44100 * 45, 44100, i) for i in (5, 1, 8)])
    fd, ld, _ = pipeline.features_and_latents(ys.pin_memory(), 44100, 24, palette, segs, chunk=2)
    fe, le, _ = pipeline.features_and_latents(ys.cuda(), 44100, 24, palette, segs, chunk=2, serial=True)
    torch.cuda.synchronize()
    assert torch.equal(fd, fe) and torch.equal(ld, le)


def test_features_and_latents_tapered_chunks():
    """B >= 3 chunks of pinned host audio: the threaded pipeline runs half chunks first and last (two feature-graph
    shapes); same results as the serially queued pipeline on whole chunks."""
    from ssar_b200 import pipeline
    from ssar_b200.synth import synth_patch_inputs

    xs = torch.stack([synth_clip(44100 * 45, 44100, i) for i in range(12)])
    _, segs, palette = synth_patch_inputs(1080, 0)
    palette = palette.cuda()
    fs, ls, _ = pipeline.features_and_latents(xs.cuda(), 44100, 24, palette, segs, chunk=4, serial=True)
    for audio in (xs.cuda(), xs.pin_memory(), xs.pin_memory(), xs.pin_memory()):  # whole chunks; tapered: eager, capture, replay
        f, l, _ = pipeline.features_and_latents(audio, 44100, 24, palette, segs, chunk=4)
        torch.cuda.synchronize()
        assert torch.equal(f, fs) and torch.equal(l, ls)
