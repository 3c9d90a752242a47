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
