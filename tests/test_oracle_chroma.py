"""CPU: pins oracle/ssar_oracle/cqt.py and dsp.spectral_contrast against reference-generated fixtures (chroma.npz)."""
import numpy as np
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import synth_clip
from ssar_oracle import cqt as ocqt
from ssar_oracle import dsp


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_tuning_cqt_chroma_tonnetz():
    g = load("chroma")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    assert float(ocqt.estimate_tuning(x, sr, bins_per_octave=36)) == float(g["tuning36"])
    assert float(ocqt.estimate_tuning(x, sr)) == float(g["tuning12"])
    C = ocqt.cqt(x, sr)
    assert C.shape == g["cqt_re"].shape  # frame count is exact
    assert rel_l2(C.real, g["cqt_re"]) < 5e-6 and rel_l2(C.imag, g["cqt_im"]) < 5e-6
    assert rel_l2(ocqt.cqt(x, sr, tuning=0.0).abs(), g["cqt0_abs"]) < 5e-6
    assert rel_l2(ocqt.chroma_cqt(x, sr), g["chroma_cqt"]) < 5e-6
    cens = ocqt.chroma_cens(x, sr)
    assert rel_l2(cens, g["cens"]) < 2e-5
    assert rel_l2(ocqt.chromagram(x, sr), g["chromagram"]) < 5e-5
    assert rel_l2(ocqt.tonnetz(x, sr, chroma=cens), g["tonnetz_of_cens"]) < 5e-5
    assert rel_l2(ocqt.tonnetz(x, sr), g["tonnetz"]) < 2e-4


def test_spectral_contrast():
    g = load("chroma")
    sr = int(g["sr"])
    x = synth_clip(int(g["n"]), sr, int(g["clip_index"]))
    c = dsp.spectral_contrast(x, sr)
    assert c.shape == g["contrast"].shape and rel_l2(c, g["contrast"]) < 5e-6
    assert rel_l2(dsp.spectral_contrast(x, sr, linear=True), g["contrast_linear"]) < 5e-6
