"""Ragged batches (SURVEY 8d cfg 5 / 8e: tracks of different durations in ONE set of launches): clips padded with zeros to
a common capacity + per-clip lengths as device data give, for every clip, the bits of a per-clip call."""
import pytest
import torch

from conftest import rel_l2
from ssar_b200.synth import synth_clip

pytestmark = pytest.mark.gpu

SR44, FPS = 44100, 24
SECS = (50.0, 61.37, 47.713, 55.5)  # all >= 43 s: the PLP window is 1024 frames for every clip


def ragged_batch(secs=SECS, pad=0.0):
    from ssar_b200.ragged import ClipLengths

    clips = [synth_clip(int(SR44 * s), SR44, 20 + i) for i, s in enumerate(secs)]
    cap = max(len(c) for c in clips) + int(pad * SR44)
    x = torch.zeros(len(clips), cap)
    for b, c in enumerate(clips):
        x[b, : len(c)] = c
    lens = ClipLengths(len(clips), "cuda").update([len(c) for c in clips], SR44, FPS * 1024)
    return clips, x.cuda(), lens


def test_ragged_features_equal_per_clip_calls():
    from ssar_b200 import pipeline

    clips, x, lens = ragged_batch(pad=3.3)  # capacity beyond the longest clip: every clip has a zero tail
    f, pf = pipeline.features_and_patch_inputs(x, SR44, FPS, lens=lens)
    for b, c in enumerate(clips):
        fs, ps = pipeline.features_and_patch_inputs(c.cuda()[None], SR44, FPS)
        T = fs.shape[1]
        assert T == lens.host_T[b]
        assert torch.equal(f[b, :T], fs[0]), (b, rel_l2(f[b, :T], fs[0]))
        for k in ps:
            assert torch.equal(pf[k][b, :T], ps[k][0]), (b, k, rel_l2(pf[k][b, :T], ps[k][0]))


def test_ragged_stages_equal_per_clip_calls():
    """Stage by stage (easier to localise than the whole chain): resample, STFT + HPSS + ISTFT, CQT, PLP, post-processing."""
    from ssar_b200 import ops, pipeline
    from ssar_b200.resampling import resample

    clips, x44, lens = ragged_batch()
    sr = FPS * 1024
    x = resample(x44, SR44, sr, clip_n_in=lens.n_in)
    singles = [resample(c.cuda()[None], SR44, sr) for c in clips]
    for b, s1 in enumerate(singles):
        n = s1.shape[1]
        assert n == lens.host_n[b] and torch.equal(x[b, :n], s1[0]) and float(x[b, n:].abs().max() if n < x.shape[1] else 0) == 0
    harm, perc = pipeline.hpss_pair(x, lens=lens)
    C = ops.cqt(harm, sr, 1024, lens=lens)
    on, on_med = pipeline.onset_pair(perc, sr, want_median=True, lens=lens)
    pulse = ops.plp_from_onset(on_med, sr, 1024, lens=lens)
    for b, s1 in enumerate(singles):
        n, T = s1.shape[1], s1.shape[1] // 1024
        h1, p1 = pipeline.hpss_pair(s1)
        assert torch.equal(harm[b, :n], h1[0]) and torch.equal(perc[b, :n], p1[0]), b
        assert float(harm[b, n:].abs().max() if n < harm.shape[1] else 0) == 0
        C1 = ops.cqt(h1, sr, 1024)
        assert torch.equal(C[b, :T], C1[0, :T]), (b, rel_l2(C[b, :T], C1[0, :T]))
        o1, m1 = pipeline.onset_pair(p1, sr, want_median=True)
        assert torch.equal(on[b, :T], o1[0]) and torch.equal(on_med[b, :T], m1[0]), b
        assert torch.equal(pulse[b, :T], ops.plp_from_onset(m1, sr, 1024)[0]), b
    g = torch.Generator().manual_seed(4)
    Ts = [700, 1000, 555]
    feats = torch.randn(3, 1000, 59, generator=g).cuda()
    from ssar_b200.ragged import ClipLengths
    L = ClipLengths(3, "cuda")
    L.update([t * 1024 for t in Ts], 1, 1)
    fin = pipeline.finish_features(feats, FPS, lens=L)
    stats = ops.feature_stats(feats, clip_frames=L.T)
    ref_stats = torch.zeros_like(stats)
    for b, T in enumerate(Ts):
        assert torch.equal(fin[b, :T], pipeline.finish_features(feats[b : b + 1, :T].contiguous(), FPS)[0]), b
        ops.feature_stats(feats[b, :T].contiguous(), ref_stats)
    assert torch.equal(stats[0], ref_stats[0]) and rel_l2(stats, ref_stats) < 1e-12


def test_ragged_pipeline_graph_equals_per_track_calls():
    """features_and_latents_ragged (longest-first chunks, capacity buckets, replayed ragged graph) == one
    features_and_latents call per track, features and latents bit for bit; the dataset statistic over valid frames only."""
    from ssar_b200 import ops, pipeline
    from ssar_b200.synth import synth_patch_inputs

    secs = (50.0, 61.37, 47.713, 55.5, 44.2)
    clips = [synth_clip(int(SR44 * s), SR44, 30 + i).cuda() for i, s in enumerate(secs)]
    Tmax = max(len(c) for c in clips) * FPS // SR44 + 2
    _, segs, palette = synth_patch_inputs(Tmax, 0)
    palette = palette.cuda()
    segs = {k: v.cuda() for k, v in segs.items()}
    ref = []
    for i, c in enumerate(clips):
        T = (-(-len(c) * 2048 // 3675)) // 1024
        f, lat, _ = pipeline.features_and_latents(c[None], SR44, FPS, palette, {k: v[:T] for k, v in segs.items()}, seed=42 + i,
                                                  serial=True)
        ref.append((f[0], lat[0]))
    for use_graphs in (False, True, True):  # eager ragged, graph capture, graph replay
        stats = torch.zeros(3, 59, dtype=torch.float64, device="cuda")
        feats, lats = pipeline.features_and_latents_ragged(clips, SR44, FPS, palette, segs, chunk=2, bucket_seconds=10.0,
                                                           use_graphs=use_graphs, stats=stats)
        torch.cuda.synchronize()
        ref_stats = torch.zeros_like(stats)
        for i, (f, lat) in enumerate(ref):
            assert torch.equal(feats[i], f), (use_graphs, i, rel_l2(feats[i], f))
            assert torch.equal(lats[i], lat), (use_graphs, i, rel_l2(lats[i], lat))
            ops.feature_stats(f, ref_stats)
        assert torch.equal(stats[0], ref_stats[0]) and rel_l2(stats, ref_stats) < 1e-12
