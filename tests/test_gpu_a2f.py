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
    tol["pulse"] = 5e-3  # measured 2.2e-3 (0 arg-max flips: test_pulse_argmax_flips_counted)
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
        fe2, pe = pipeline.features_and_patch_inputs(xs, 44100, 24)
        assert torch.equal(fe2, fe) and all(torch.equal(pf[k], pe[k]) for k in pe)


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
    # software-pipelined host batches (stage_batch): batch k + 1 is in flight while batch k is computed on; three batches
    # through the two device buffers, each result compared after the buffer it read from has been re-staged
    xp, yp = xs.cpu().pin_memory(), ys.pin_memory()
    h0 = pipeline.stage_batch(xp, "cuda:0")
    h1 = pipeline.stage_batch(yp, "cuda:0")
    g0 = pipeline.features_and_latents(h0, 44100, 24, palette, segs, chunk=2)
    h2 = pipeline.stage_batch(xp, "cuda:0")  # reuses h0's buffer: must wait for g0's kernels
    g1 = pipeline.features_and_latents(h1, 44100, 24, palette, segs, chunk=2)
    g2 = pipeline.features_and_latents(h2, 44100, 24, palette, segs, chunk=2)
    torch.cuda.synchronize()
    assert h0.tensor.data_ptr() == h2.tensor.data_ptr() != h1.tensor.data_ptr()
    assert torch.equal(g0[0], fa) and torch.equal(g0[1], la) and torch.equal(g1[0], fe) and torch.equal(g1[1], le)
    assert torch.equal(g2[0], fa) and torch.equal(g2[1], la)


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


def test_audio2features_short_clips_vs_golden():
    """Clips shorter than the 1024-frame tempogram window (rosa/beat.py:46-47: n_fft = T): T = 240, 384, 1008 — the
    band-limited direct DFT / inverse kernels stand in for the radix FFT."""
    from ssar_b200.supervised.data import FEATURE_NAMES, audio2features

    g = load("a2f_short")
    for secs, ci in ((10, 11), (16, 12), (42, 13)):
        x = synth_clip(44100 * secs, 44100, ci).cuda()
        raw = audio2features(x, 44100, 24, clamp=False, smooth=False)
        ref = g[f"features_raw_{secs}s"]
        assert tuple(raw.shape) == tuple(ref.shape) == (24 * secs, 59)
        worst = {n: rel_l2(raw[:, j], ref[:, j]) for j, n in enumerate(FEATURE_NAMES)}
        bad = {k: v for k, v in worst.items() if v > 2e-3}
        assert not bad, (secs, bad)
        f = audio2features(x, 44100, 24)
        keep = [j for j, n in enumerate(FEATURE_NAMES) if n != "pulse"]
        assert rel_l2(f[:, keep], g[f"features_{secs}s"][:, keep]) < 1e-4
        assert rel_l2(f, g[f"features_{secs}s"]) < 1e-3


def test_plp_short_even_and_odd_vs_golden():
    """plp on T = 300 (even window) and T = 333 (odd window: torch.stft / istft with an odd n_fft) from the golden
    median-aggregated onset envelope: exact given the envelope (1e-5), 1e-3 through the GPU's own envelope."""
    from ssar_b200 import ops
    from ssar_b200.features.rosa import beat

    g = load("a2f_short")
    for T, ci in ((300, 14), (333, 15)):
        env = g[f"onset_median_T{T}"].cuda()
        p = ops.plp_from_onset(env[None], 24576)[0]
        assert tuple(p.shape) == (T,) and rel_l2(p, g[f"plp_T{T}"]) < 1e-5, T
        x = synth_clip(1024 * T + 17, 24576, ci).cuda()
        assert rel_l2(beat.plp(x, 24576), g[f"plp_T{T}"]) < 1e-3, T


def test_frames_dft_and_band_inverse_vs_torch():
    """ssar_frames_dft / ssar_istft_band against torch.stft / torch.istft for window lengths that are not powers of two."""
    from ssar_b200 import ops

    gen = torch.Generator().manual_seed(3)
    for n_fft, n in ((240, 240), (333, 333), (1008, 1008), (384, 500)):
        x = torch.randn(2, n, generator=gen)
        win = torch.hann_window(n_fft)
        ref = torch.stft(x, n_fft, 1, window=win, center=True, pad_mode="reflect", return_complex=True)  # [B, F, frames]
        F, n_frames = n_fft // 2 + 1, 1 + n + 2 * (n_fft // 2) - n_fft  # an odd window gives n frames, an even one n + 1
        assert ref.shape[2] == n_frames
        got = ops.frames_dft(x.cuda(), n_fft, 1, n_frames, torch.hann_window)
        assert tuple(got.shape) == (2, n_frames, F, 2)
        assert rel_l2(torch.view_as_complex(got.cpu()).transpose(1, 2).abs(), ref.abs()) < 1e-5
        assert rel_l2(torch.view_as_real(torch.view_as_complex(got.cpu()).transpose(1, 2)), torch.view_as_real(ref)) < 1e-5
        k_lo, k_hi = F // 5, F // 2
        band = ref.clone()
        band[:, :k_lo] = 0
        band[:, k_hi:] = 0
        y_ref = torch.istft(band, n_fft, 1, window=win, center=True, length=n)
        y = ops.istft_band(got[:, :, k_lo:k_hi].contiguous(), n_fft, 1, torch.hann_window, k_lo, n)
        assert rel_l2(y, y_ref) < 1e-5, n_fft


def test_pulse_argmax_flips_counted():
    """Why the pulse column is held to 1e-2 through the chain: plp keeps ONE arg-max tempogram bin per frame, and a
    near-tie flips under the 1e-5 fp32 noise of the onset envelope.  Count the frames whose kept bin differs between the
    GPU's and the oracle's envelope on the golden clip: a handful out of 1081, and the pulse is exact (1e-5) elsewhere."""
    from ssar_b200 import ops, pipeline
    from ssar_b200.features.audio import percussive
    from ssar_b200.resampling import resample
    from ssar_oracle import dsp

    g = load("a2f")
    x = synth_clip(int(g["n44"]), 44100, int(g["clip_index44"]))
    sr = 24 * 1024

    def kept_bins(env):
        G = dsp.stft(env.cpu().float(), min(len(env), 1024), 1)
        bpm = torch.linspace(0, float(sr * 60 / 1024.0) / 2, G.shape[0])
        mag = torch.where(((bpm < 60) | (bpm > 180))[:, None], torch.zeros(()), G.abs())
        return mag.argmax(0)

    perc_gpu = percussive(resample(x.cuda()[None], 44100, sr))
    env_gpu = pipeline.onset_pair(perc_gpu, sr, want_median=True)[1][0]
    env_ref = dsp.onset_strength(dsp.percussive(dsp.resample(x, 44100, sr)), sr, aggregate="median")
    assert rel_l2(env_gpu, env_ref) < 1e-3
    flips = int((kept_bins(env_gpu) != kept_bins(env_ref)).sum())
    print("plp arg-max flips on the golden clip:", flips, "of", len(env_ref) + 1)
    assert flips <= 2
    # and given the GPU envelope, the CUDA plp equals the oracle's arithmetic on that same envelope
    p_gpu = ops.plp_from_onset(env_gpu[None], sr)[0]
    G = dsp.stft(env_gpu.cpu(), 1024, 1)
    bpm = torch.linspace(0, float(sr * 60 / 1024.0) / 2, 513)
    G = torch.where(((bpm < 60) | (bpm > 180))[:, None], torch.zeros((), dtype=G.dtype), G)
    st = torch.log1p(1e6 * G.abs())
    G = torch.where(st < st.max(0, keepdim=True).values, torch.zeros((), dtype=G.dtype), G)
    G = G / (torch.finfo(torch.float32).tiny ** 0.5 + G.abs().max(0, keepdim=True).values)
    p = dsp.istft(G, 1024, 1, length=len(env_gpu)).clamp(min=0)
    p = (p - p.min()) / ((p - p.min()).max() + 1e-8)
    assert rel_l2(p_gpu, p) < 1e-4


def test_mir_features_vs_golden():
    """random/mir.py:14-22,26,44: salience_weighted and the normalised Patch inputs, per-function mirror and batched
    pipeline, against the unmodified retrieve_music_information's feature dict."""
    from ssar_b200 import pipeline
    from ssar_b200.random import mir
    from ssar_b200.synth import synth_features

    g = load("mir")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"])).cuda()
    env = synth_features(500, 3, 21).abs().cuda()
    assert rel_l2(mir.salience_weighted(env), g["salience_2d"]) < 1e-5
    assert rel_l2(mir.salience_weighted(env[:, 0]), g["salience_1d"]) < 1e-5
    names = ["chromagram", "tonnetz", "mfcc", "spectral_contrast", "spectral_flatness", "rms", "drop_strength", "onsets"]
    got = mir.retrieve_music_features(x, sr)
    assert list(got) == names
    batched = pipeline.music_features_batch(torch.stack([x, x.flip(0)]), sr)
    assert list(batched) == names
    # tonnetz and mfcc are signed: mir.py's salience weight (g5 / g80)^2 divides by a smoothed signal that crosses zero,
    # which amplifies the 1e-7 fp32 noise of the raw feature to 1e-4 .. 1e-2 (ill-conditioning of the reference's formula;
    # scripts/diag_mir.py prints raw and weighted errors side by side); the other seven features are held to 1e-4
    tol = {k: 1e-4 for k in names}
    tol["tonnetz"] = tol["mfcc"] = 5e-2  # the two signed features
    for k in names:
        ref = g[f"feat_{k}"]
        assert tuple(got[k].shape) == tuple(ref.shape) and rel_l2(got[k], ref) < tol[k], (k, rel_l2(got[k], ref))
        assert tuple(batched[k].shape) == (2,) + tuple(ref.shape)
        assert rel_l2(batched[k][0], ref) < tol[k], (k, rel_l2(batched[k][0], ref))


def test_features_and_patch_inputs_match_their_single_purpose_calls():
    """The shared pass (one STFT per signal, both chroma chains) == audio2features_batch and music_features_batch."""
    from ssar_b200 import pipeline
    from ssar_b200.resampling import resample

    xs = torch.stack([synth_clip(44100 * 20, 44100, i) for i in (3, 9)]).cuda()
    f, pf = pipeline.features_and_patch_inputs(xs, 44100, 24)
    assert torch.equal(f, pipeline.audio2features_batch(xs, 44100, 24))
    mf = pipeline.music_features_batch(resample(xs, 44100, 24 * 1024), 24 * 1024)
    assert all(torch.equal(pf[k], mf[k]) for k in mf)


def test_pcm16_input_equals_normalised_float_input():
    """int16 PCM in (device or pinned host, eager / graph / serial) == the same audio as x / 32768 fp32 (torchaudio.load's
    normalisation), bit for bit."""
    from ssar_b200 import ops, pipeline
    from ssar_b200.synth import synth_patch_inputs

    xs = torch.stack([synth_clip(44100 * 45, 44100, i) for i in (3, 9, 4, 7)])
    pcm = (xs * 32767).round().clamp(-32768, 32767).to(torch.int16)
    xf = pcm.float() / 32768
    assert torch.equal(ops.pcm16_to_float(pcm.cuda()).cpu(), xf)
    assert torch.equal(ops.pcm16_to_float(pcm.cuda()[:, 1:777]).cpu(), xf[:, 1:777])  # unaligned / ragged tail path
    _, segs, palette = synth_patch_inputs(1080, 0)
    palette = palette.cuda()
    f0, l0, _ = pipeline.features_and_latents(xf.cuda(), 44100, 24, palette, segs, chunk=2, serial=True)
    for audio in (pcm.cuda(), pcm.pin_memory(), pcm.pin_memory()):  # eager, capture, replay
        f, l, _ = pipeline.features_and_latents(audio, 44100, 24, palette, segs, chunk=2)
        torch.cuda.synchronize()
        assert torch.equal(f, f0) and torch.equal(l, l0)
    f, l, _ = pipeline.features_and_latents(pcm.cuda(), 44100, 24, palette, segs, chunk=2, serial=True)
    assert torch.equal(f, f0) and torch.equal(l, l0)
