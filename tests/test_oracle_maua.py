"""CPU: pins oracle/ssar_oracle/maua.py against fixtures produced by the reference's own audioreactive/latent.py and
signal.py imported unmodified (make_golden.py maua)."""
import numpy as np
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import projection_matrix, synth_features, synth_patch_inputs
from ssar_oracle import maua

T = 600
CASES = {"a": (0, 5, 1), "b": (5, 8, 3), "c": (8, 20, 2)}


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def inputs():
    _, _, palette = synth_patch_inputs(T, 0)
    env = synth_features(T, 1, 11).abs().reshape(T)
    env = env / env.max()
    envs = synth_features(T, 7, 12).abs() + 0.05
    return palette, env, envs, synth_features(T, 5, 13)


def test_loops():
    g, R = load("maua"), projection_matrix()
    palette, _, _, _ = inputs()
    for name, (k0, k1, loops) in CASES.items():
        sl = maua.slerp_loops(palette[k0:k1], T, loops)
        assert sl.shape == (T, 18, 512)
        assert rel_l2(sl @ R, g[f"slerp_{name}_proj"]) < 2e-6 and rel_l2(sl[::97], g[f"slerp_{name}_frames"]) < 2e-6
        assert rel_l2(maua.spline_loops(palette[k0:k1], T, loops) @ R, g[f"spline_{name}_proj"]) < 1e-5
    assert rel_l2(maua.tempo_loops(palette[:4], T, 24, 96.0, type="slerp") @ R, g["tempo_slerp_proj"]) < 2e-6


def test_weighted_select_gauss():
    g, R = load("maua"), projection_matrix()
    palette, env, envs, x = inputs()
    assert rel_l2(maua.single_weighted(palette[0], palette[1], env) @ R, g["single_proj"]) < 2e-6
    assert rel_l2(maua.multi_weighted(palette[:7], envs.clone()) @ R, g["multi_proj"]) < 2e-6
    assert rel_l2(maua.multi_weighted(palette[:3], envs.clone()) @ R, g["multi_wrap_proj"]) < 2e-6
    sm = maua.select_modulo(palette[:6], env.clone(), smooth=2)
    assert rel_l2(sm @ R, g["select_proj"]) < 2e-6 and rel_l2(sm[::97], g["select_frames"]) < 2e-6
    assert rel_l2(maua.gaussian_filter(x, 3.0), g["gauss_causal_none"]) < 2e-6
    assert rel_l2(maua.gaussian_filter(x, 3.0, causal=0), g["gauss_causal_0"]) < 2e-6
    assert rel_l2(maua.gaussian_filter(x, 3.0, causal=0.5), g["gauss_causal_half"]) < 2e-6
    assert rel_l2(maua.normalize(x), g["normalize"]) < 1e-7
