"""GPU parity: trajectory kernels (mix GEMM, weight merges, noise programs, fixed decoder) vs reference goldens."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import projection_matrix, synth_features, synth_patch_inputs

pytestmark = pytest.mark.gpu
T = 600
TOL = 1e-4  # north-star fp32 tolerance class for sums of latents (summation order differs from the reference)


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def sub(nz):
    return nz if nz.shape[-1] <= 64 else nz[:, ::8, ::8]


def cuda_inputs():
    feats, segs, palette = synth_patch_inputs(T, 0)
    return {k: v.cuda() for k, v in feats.items()}, segs, palette.cuda()


def test_patch_forward_vs_golden():
    from ssar_b200.random.patch import Patch

    g = load("traj")
    R = projection_matrix().cuda()
    for seed in (42, 7):
        feats, segs, palette = cuda_inputs()
        patch = Patch(feats, segs, 120.0, fps=24, seed=seed, device="cpu")  # CPU generator: same draws as the golden run
        assert [len(patch.latent_patches), len(patch.noise_patches)] == g[f"n_sub_{seed}"].tolist()
        lat, noise = patch.forward(palette, downscale_factor=4)
        assert tuple(lat.shape) == (T, 18, 512) and len(noise) == 17
        assert rel_l2(lat @ R, g[f"lat_proj_{seed}"]) < TOL
        assert rel_l2(lat[::120], g[f"lat_frames_{seed}"]) < TOL
        for j in (0, 3, 8, 12, 16):
            nz = noise[j].forward(10, 7)
            assert rel_l2(sub(nz), g[f"noise_{seed}_{j}"]) < TOL, j
        tail = noise[6].forward(T - 3, 32)
        assert tuple(tail.shape) == tuple(g[f"noise_tail_{seed}"].shape)
        assert rel_l2(tail, g[f"noise_tail_{seed}"]) < TOL


def test_spline_latent_patch_decoder_vs_golden():
    from ssar_b200.models.latent_n_noise2 import FixedLatentNoiseDecoder
    from ssar_b200.random import latent

    g = load("traj")
    R = projection_matrix().cuda()
    feats, segs, palette = cuda_inputs()
    assert rel_l2(latent.spline_loop_latents(palette[:8], T, 1) @ R, g["spline_proj"]) < TOL
    assert rel_l2(latent.spline_loop_latents(palette[8:13], T, 2.37) @ R, g["spline_loops_proj"]) < TOL
    rng = torch.Generator("cpu").manual_seed(5)
    lat = latent.spline_loop_latents(palette[:6], T)
    ptr0 = lat.data_ptr()
    for kind, merge, depth, sf in (("segmentation", "average", "low", "mfcc"), ("feature", "modulate", "midhigh", "chromagram"),
                                   ("feature", "modulate", "all", "onsets"), ("loop", "overwrite", "high", "tonnetz")):
        lat = latent.latent_patch(rng, lat, palette, segs, feats, 120.0, 24, kind, 6, 8, sf, 0.8, "rms", 0.9, merge, depth)
    assert lat.data_ptr() == ptr0  # in place, like the reference
    assert rel_l2(lat @ R, g["latent_patch_proj"]) < TOL
    x = (synth_features(2 * 300, 17, 77).abs().reshape(2, 300, 17) + 0.1).cuda()
    torch.manual_seed(0)
    draws = [torch.randn(2, 300, s, s).cuda() for s in (4, 8, 16, 32)]
    dec = FixedLatentNoiseDecoder(palette[:9], hidden_size=3, n_latent_split=3, n_noise=4)
    dl, dn = dec.forward(x, noise_draws=draws)
    assert tuple(dl.shape) == (2, 300, 18, 512)
    assert rel_l2(dl @ R, g["dec_proj"]) < TOL
    for i, n in enumerate(dn):
        assert tuple(n.shape) == (2, 300, 4 * 2 ** i, 4 * 2 ** i)
        assert rel_l2(n[:, ::10], g[f"dec_noise_{i}"]) < TOL, i


def test_patch_full_size_vs_oracle_frames():
    """cfg 4 size (T = 5760): the single mixing GEMM equals the step-by-step oracle on a strided set of frames, and the
    general noise-tree fallback (composite right child) equals direct evaluation."""
    from ssar_b200.random import noise as N
    from ssar_b200.random.patch import Patch
    from ssar_oracle import traj

    Tf = 5760
    feats, segs, palette = synth_patch_inputs(Tf, 1)
    lat_o, noise_o = traj.Patch(feats, segs, 120.0, seed=11).forward(palette, downscale_factor=4)
    patch = Patch({k: v.cuda() for k, v in feats.items()}, segs, 120.0, seed=11, device="cpu")
    lat, noise = patch.forward(palette.cuda(), downscale_factor=4)
    assert tuple(lat.shape) == (Tf, 18, 512)
    assert rel_l2(lat[::97], lat_o[::97]) < TOL
    for j in (2, 9, 14):
        assert rel_l2(noise[j].forward(4000, 32), noise_o[j](4000, 32)) < TOL, j
    # composite right child -> BUFFER fallback
    rng = torch.Generator("cpu").manual_seed(1)
    a = N.Loop(rng, Tf, (16, 16), 2, 3.0, device="cuda")
    b = N.ScaleBias(N.Multiply(rng, Tf, (16, 16), feats["mfcc"].cuda()), 0.5, 0.1)
    tree = N.Average(a, b)
    direct = (a.forward(5, 9) + b.forward(5, 9)) / 2
    assert rel_l2(tree.forward(5, 9), direct) < 1e-6


def _eager_configs(features, segmentations, seed, device, min_subpatches=5, max_subpatches=20):
    """patch.py:41-66 written with the eager helpers (`random_choice` = real Tensor.multinomial, tensor `skewnorm`)."""
    from ssar_b200.random.mir import ALLFEATS, UNITFEATS
    from ssar_b200.random.patch import random_choice, skewnorm

    rng = torch.Generator(device)
    rng.manual_seed(seed)
    head = [torch.randint(3, 15, size=(), generator=rng, device=rng.device).item(),
            1 + 9 * torch.rand((), generator=rng, device=rng.device).item(),
            random_choice(rng, [1, 2, 4, 8, 16, 32, 64])]
    ks = np.unique([k for (_, k) in segmentations]).tolist()
    depth = (["low", "mid", "high", "lowmid", "midhigh", "all"], [3, 3, 3, 2, 2, 1])

    def common(kinds, with_segments):
        d = dict(patch_type=random_choice(rng, kinds))
        if with_segments:
            d["segments"] = random_choice(rng, ks)
        d.update(loop_bars=random_choice(rng, [4, 8, 16, 32], weights=[2, 2, 2, 1]), seq_feat=random_choice(rng, ALLFEATS),
                 seq_feat_weight=skewnorm(rng, a=5, loc=0.666, scale=0.5).item(), mod_feat=random_choice(rng, UNITFEATS),
                 mod_feat_weight=skewnorm(rng, a=5, loc=0.666, scale=0.5).item(),
                 merge_type=random_choice(rng, ["average", "modulate"], weights=[1, 3]),
                 merge_depth=random_choice(rng, depth[0], weights=depth[1]))
        return d

    lat = [common(["segmentation", "feature", "loop"], True)
           for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device))]
    noi = []
    for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device)):
        d = common(["blend", "multiply", "loop"], False)
        d["noise_mean"] = torch.randn((), generator=rng, device=rng.device).item() * 0.5
        d["noise_std"] = skewnorm(rng, a=5, loc=0.666, scale=0.5).item()
        noi.append(d)
    return head, lat, noi, rng.get_state()


def test_deferred_sampler_equals_eager_draws_cuda_generator():
    """The deferred sampler (one exponential_ per choice, arithmetic after one bulk copy) reproduces the eager
    multinomial / skewnorm draws of patch.py:12-32 on a CUDA (Philox) generator value for value."""
    from ssar_b200.random.patch import Patch

    from ssar_b200.random import patch as patch_mod

    feats, segs, _ = synth_patch_inputs(64, 0)
    for graphs in (True, False):  # CUDA-graph replay of the draw sequence / eager deferred draws
        patch_mod.USE_SAMPLER_GRAPHS = graphs
        try:
            for seed in range(12):
                p = Patch(feats, segs, 120.0, seed=seed, device="cuda")
                head, lat, noi, state = _eager_configs(feats, segs, seed, "cuda")
                assert [p.num_base_latents, p.sigma_base_noise, p.loops_base_noise] == head, (graphs, seed)
                assert p.latent_patches == lat and p.noise_patches == noi, (graphs, seed)
                assert torch.equal(p.rng.get_state(), state), (graphs, seed)  # later draws continue identically
        finally:
            patch_mod.USE_SAMPLER_GRAPHS = True


def test_subpatch_kernel_vs_separate_steps():
    """ssar_subpatch_merge / ssar_onehot_gauss vs the unfused sequence (scale, gauss1d, index_add, weights_merge)."""
    from ssar_b200 import ops

    g = torch.Generator("cpu").manual_seed(3)
    Tn, G, P = 601, 3, 180
    w0 = torch.randn(Tn, G, P, generator=g).cuda()
    mod = torch.rand(Tn, 1, generator=g).cuda()
    for K, lerp, use_mod, (g0, g1) in [(12, False, True, (0, 2)), (2, True, False, (1, 3)), (20, False, False, (0, 3)),
                                       (2, True, True, (2, 3))]:
        sel = torch.randperm(P, generator=g)[:K].cuda()
        src = torch.rand(Tn, 1 if lerp else K, generator=g).cuda()
        scale, ms = 0.83, 1.21
        f = scale * src
        S = ops.gauss1d(torch.cat([f, 1 - f], 1) if lerp else f, 1, "circular")
        D = torch.zeros(Tn, P, device="cuda").index_add_(1, sel, S)
        if use_mod:
            m = (ms * mod).reshape(Tn)
            ref = ops.weights_merge(w0.clone(), D, g0, g1, (1 - m).contiguous(), m.contiguous())
        else:
            ref = ops.weights_merge(w0.clone(), D, g0, g1, 0.5, 0.5)
        got = ops.subpatch_merge(w0.clone(), g0, g1, sel, src, src_scale=scale, lerp=lerp, sigma=1,
                                 mod=mod if use_mod else None, mod_scale=ms, alpha=0.5, beta=0.5)
        assert torch.equal(got, ref), (K, lerp, use_mod, float((got - ref).abs().max()))
    labels = torch.randint(0, 6, (Tn,), generator=g).cuda()
    onehot = torch.zeros(Tn, 6, device="cuda")
    onehot[torch.arange(Tn, device="cuda"), labels] = 1.0
    assert torch.equal(ops.onehot_gauss(labels, 6, 5), ops.gauss1d(onehot, 5, "circular"))
    # overwrite mode never reads the old weights (NaN-poisoned buffer)
    sel = torch.arange(5, device="cuda")
    src = torch.rand(Tn, 5, generator=g).cuda()
    out = ops.subpatch_merge(torch.full((Tn, G, P), float("nan"), device="cuda"), 0, G, sel, src, sigma=0, alpha=0.0, beta=1.0)
    assert torch.equal(out[:, 1, :5], src) and float(out[:, :, 5:].abs().max()) == 0.0


def test_tensor_core_mix_vs_fp32_kernel_and_fp64():
    """ssar_mix_gemm_tc (tcgen05, fp16 3-term split, TMA in / TMA out) against the CUDA-core fp32 kernel and an fp64
    product: full-size render, ragged row counts (TMA clips the last row block), one group, K not a multiple of 64."""
    from ssar_b200 import ops

    g = torch.Generator("cpu").manual_seed(11)
    try:
        for Tn, G, K in [(5760, 3, 180), (333, 3, 192), (1000, 1, 64), (129, 3, 7), (4097, 2, 100)]:
            W = (torch.rand(Tn, G, K, generator=g) * (torch.rand(Tn, G, K, generator=g) < 0.5)).cuda()
            P = torch.randn(K, 18, 512, generator=g).cuda()
            ops.USE_TENSOR_CORE_MIX = False
            ref32 = ops.mix_gemm(W, P)
            ops.USE_TENSOR_CORE_MIX = True
            launches = __import__("ssar_b200._cabi", fromlist=["x"]).launch_count
            out = torch.full((Tn + 3, 18, 512), float("nan"), device="cuda")  # guard rows: the kernel must not touch them
            got = ops.mix_gemm(W, P, out=out[:Tn])
            assert torch.isnan(out[Tn:]).all()
            ref64 = torch.einsum("tgk,kgn->tgn", W.double(), P.reshape(K, G, -1).double()).reshape(Tn, 18, 512)
            assert rel_l2(got.double(), ref64) < 5e-6, (Tn, G, K)
            assert rel_l2(got, ref32) < 5e-6, (Tn, G, K)
    finally:
        ops.USE_TENSOR_CORE_MIX = True


def test_tensor_core_mix_operand_ranges_and_presplit_basis():
    """Operands far outside fp16's range (palette x 1e5: fp16(value) would be inf; palette x 1e-6: the lo halves would be
    subnormal) keep fp32 accuracy through the per-operand power-of-two scaling; a basis split once
    (`ops.SplitBasis`) gives the same bits as the per-call split."""
    from ssar_b200 import ops

    g = torch.Generator("cpu").manual_seed(12)
    Tn, G, K = 1024, 3, 180
    W = torch.rand(Tn, G, K, generator=g).cuda()
    P = torch.randn(K, 18, 512, generator=g).cuda()
    for ws, ps in ((1.0, 1e5), (1.0, 1e-6), (3e4, 1e3), (1e-5, 1e-5)):
        got = ops.mix_gemm(W * ws, P * ps)
        ref64 = torch.einsum("tgk,kgn->tgn", (W * ws).double(), (P * ps).reshape(K, G, -1).double()).reshape(Tn, 18, 512)
        assert bool(torch.isfinite(got).all()) and rel_l2(got.double(), ref64) < 2e-6, (ws, ps, rel_l2(got.double(), ref64))
    sp = ops.SplitBasis(P)
    assert sp.block is not None
    assert torch.equal(ops.mix_gemm(W, P, basis_split=sp), ops.mix_gemm(W, P))
    W2 = torch.rand(300, G, K, generator=g).cuda()
    assert torch.equal(ops.mix_gemm(W2, P, basis_split=sp), ops.mix_gemm(W2, P))


def test_spline_loop_weights_kernel_vs_host_tables():
    """ssar_spline_loop_weights (device linspace + bucketize + fp64 cubic) against plan.spline_loop_weights_cpu (the host
    tables the goldens of spline_loop_latents were checked with): the device linspace may sit an ulp of t away from
    torch's CPU linspace (t reaches n_loops = 7 in one case), the spline is continuous -> 1e-5."""
    from ssar_b200 import plan

    for K, T, nl in ((8, 600, 1), (5, 777, 2.37), (12, 5760, 0.013), (3, 1000, 7), (20, 8641, 1)):
        plan.spline_loop_weights.cache_clear()
        w = plan.spline_loop_weights(K, T, nl, "cuda:0")
        ref = plan.spline_loop_weights_cpu(K, T, nl)
        assert tuple(w.shape) == (T, K) and float((w.cpu() - ref).abs().max()) < 1e-5, (K, T, nl, float((w.cpu() - ref).abs().max()))
        assert float((w.sum(1) - 1).abs().max()) < 1e-5  # cardinal splines sum to one
