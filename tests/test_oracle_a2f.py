"""CPU: pins the oracle's plp / onsets / drop_strength / audio2features against reference-generated fixtures."""
import numpy as np
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_clip
from ssar_oracle import dsp, features


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_plp_onsets_drop():
    g = load("a2f")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    p = dsp.plp(x, sr)
    assert p.shape == g["plp"].shape and rel_l2(p, g["plp"]) < 1e-5
    assert rel_l2(dsp.plp(dsp.percussive(x), sr).unsqueeze(-1), g["pulse"]) < 1e-4
    assert rel_l2(features.onsets(x, sr), g["onsets"]) < 1e-5
    assert rel_l2(features.drop_strength(x, sr), g["drop_strength"]) < 1e-5


def test_audio2features():
    g = load("a2f")
    x = synth_clip(int(g["n44"]), 44100, int(g["clip_index44"]))
    raw = features.audio2features(x, 44100, 24, clamp=False, smooth=False)
    assert raw.shape == g["features_raw"].shape == (1080, 59)
    for j in range(59):
        assert rel_l2(raw[:, j], g["features_raw"][:, j]) < 1e-4, j
    f = features.audio2features(x, 44100, 24)
    assert rel_l2(f, g["features"]) < 1e-5


def test_audio2features_short_clips():
    """Clips shorter than the 1024-frame tempogram window (rosa/beat.py:46-47: n_fft = T): T = 240, 384, 1008."""
    g = load("a2f_short")
    for secs, ci in ((10, 11), (16, 12), (42, 13)):
        x = synth_clip(44100 * secs, 44100, ci)
        raw = features.audio2features(x, 44100, 24, clamp=False, smooth=False)
        ref = g[f"features_raw_{secs}s"]
        assert raw.shape == ref.shape == (24 * secs, 59)
        for j in range(59):
            assert rel_l2(raw[:, j], ref[:, j]) < 1e-4, (secs, j)
        assert rel_l2(features.audio2features(x, 44100, 24), g[f"features_{secs}s"]) < 1e-5


def test_plp_short_even_and_odd():
    g = load("a2f_short")
    for T, ci in ((300, 14), (333, 15)):
        x = synth_clip(1024 * T + 17, 24576, ci)
        assert rel_l2(dsp.onset_strength(x, 24576, aggregate="median"), g[f"onset_median_T{T}"]) < 1e-5
        p = dsp.plp(x, 24576)
        assert p.shape == g[f"plp_T{T}"].shape == (T,) and rel_l2(p, g[f"plp_T{T}"]) < 1e-4, T


def test_mir_features():
    """random/mir.py:26,44 (the unmodified retrieve_music_information ran with stand-in tempo / segmentations)."""
    from ssar_oracle import post

    g = load("mir")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    env = __import__("ssar_b200.synth", fromlist=["synth_features"]).synth_features(500, 3, 21).abs()
    assert rel_l2(features.salience_weighted(env), g["salience_2d"]) < 1e-6
    assert rel_l2(features.salience_weighted(env[:, 0]), g["salience_1d"]) < 1e-6
    got = features.retrieve_music_features(x, sr)
    assert list(got) == ["chromagram", "tonnetz", "mfcc", "spectral_contrast", "spectral_flatness", "rms", "drop_strength",
                         "onsets"]
    for k, v in got.items():
        ref = g[f"feat_{k}"]
        assert v.shape == ref.shape and rel_l2(v, ref) < 1e-4, (k, rel_l2(v, ref))
