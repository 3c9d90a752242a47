"""GPU: ssar_b200.audioreactive (maua audioreactive/latent.py + signal.py mirror) against the reference-generated
fixtures and the CPU oracle.  Tolerances (rel-L2): trajectories 1e-4 (the mixing GEMM's 3-term fp16 split: ~1e-5)."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import projection_matrix, synth_features, synth_patch_inputs

pytestmark = pytest.mark.gpu
T = 600
CASES = {"a": (0, 5, 1), "b": (5, 8, 3), "c": (8, 20, 2)}


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def test_slerp_and_spline_loops_vs_reference():
    from ssar_b200 import audioreactive as ar

    g, R = load("maua"), projection_matrix()
    _, _, palette = synth_patch_inputs(T, 0)
    pal = palette.cuda()
    for name, (k0, k1, loops) in CASES.items():
        sl = ar.slerp_loops(pal[k0:k1], T, loops).cpu()
        assert sl.shape == (T, 18, 512)
        assert rel_l2(sl @ R, g[f"slerp_{name}_proj"]) < 1e-5, name
        assert rel_l2(sl[::97], g[f"slerp_{name}_frames"]) < 1e-5, name
        sp = ar.spline_loops(pal[k0:k1], T, loops).cpu()
        assert rel_l2(sp @ R, g[f"spline_{name}_proj"]) < 1e-4, name
    ts = ar.tempo_loops(pal[:4], T, 24, 96.0, type="slerp").cpu()
    assert rel_l2(ts @ R, g["tempo_slerp_proj"]) < 1e-5


def test_slerp_loops_full_size_vs_oracle():
    from ssar_b200 import audioreactive as ar
    from ssar_oracle import maua

    _, _, palette = synth_patch_inputs(8, 3)
    out = ar.slerp_loops(palette[:9].cuda(), 5760, 4).cpu()
    ref = maua.slerp_loops(palette[:9], 5760, 4)
    assert out.shape == (5760, 18, 512) and rel_l2(out, ref) < 1e-5
    # every frame is a blend of two unit vectors
    assert float(out.norm(dim=-1).max()) <= 1.0 + 1e-5


def test_weighted_select_gauss_vs_reference():
    from ssar_b200 import audioreactive as ar

    g, R = load("maua"), projection_matrix()
    _, _, palette = synth_patch_inputs(T, 0)
    pal = palette.cuda()
    env = synth_features(T, 1, 11).abs().reshape(T)
    env = (env / env.max()).cuda()
    envs = (synth_features(T, 7, 12).abs() + 0.05).cuda()
    x = synth_features(T, 5, 13).cuda()
    assert rel_l2(ar.single_weighted(pal[0], pal[1], env).cpu() @ R, g["single_proj"]) < 1e-4
    e1 = envs.clone()
    assert rel_l2(ar.multi_weighted(pal[:7], e1).cpu() @ R, g["multi_proj"]) < 1e-4
    assert torch.allclose(e1.sum(1), torch.ones(T, device="cuda"), atol=1e-5)  # normalised in place, like the reference
    assert rel_l2(ar.multi_weighted(pal[:3], envs.clone()).cpu() @ R, g["multi_wrap_proj"]) < 1e-4
    sm = ar.select_modulo(pal[:6], env.clone(), smooth=2).cpu()
    assert rel_l2(sm @ R, g["select_proj"]) < 1e-4 and rel_l2(sm[::97], g["select_frames"]) < 1e-4
    assert rel_l2(ar.gaussian_filter(x, 3.0), g["gauss_causal_none"]) < 1e-5
    assert rel_l2(ar.gaussian_filter(x, 3.0, causal=0), g["gauss_causal_0"]) < 1e-5
    assert rel_l2(ar.gaussian_filter(x, 3.0, causal=0.5), g["gauss_causal_half"]) < 1e-5
    assert rel_l2(ar.normalize(x), g["normalize"]) < 1e-6
