"""CPU: pins oracle/ssar_oracle/traj.py against fixtures produced by the unmodified reference (make_golden.py traj)."""
import numpy as np
import torch

from conftest import GOLDEN, rel_l2
from ssar_b200.synth import projection_matrix, synth_features, synth_patch_inputs
from ssar_oracle import traj

T = 600


def load(name):
    return {k: torch.from_numpy(np.asarray(v)) for k, v in np.load(GOLDEN / f"{name}.npz").items()}


def sub(nz):
    return nz if nz.shape[-1] <= 64 else nz[:, ::8, ::8]


def test_patch_forward():
    g = load("traj")
    R = projection_matrix()
    for seed in (42, 7):
        feats, segs, palette = synth_patch_inputs(T, 0)
        patch = traj.Patch(feats, segs, 120.0, fps=24, seed=seed)
        assert [len(patch.latent_patches), len(patch.noise_patches)] == g[f"n_sub_{seed}"].tolist()
        lat, noise = patch.forward(palette, downscale_factor=4)
        assert lat.shape == (T, 18, 512)
        assert rel_l2(lat @ R, g[f"lat_proj_{seed}"]) < 2e-6
        assert rel_l2(lat[::120], g[f"lat_frames_{seed}"]) < 2e-6
        for j in (0, 3, 8, 12, 16):
            assert rel_l2(sub(noise[j](10, 7)), g[f"noise_{seed}_{j}"]) < 2e-6, j
        tail = noise[6](T - 3, 32)
        assert tail.shape == g[f"noise_tail_{seed}"].shape and rel_l2(tail, g[f"noise_tail_{seed}"]) < 2e-6


def test_spline_latent_patch_decoder():
    g = load("traj")
    R = projection_matrix()
    feats, segs, palette = synth_patch_inputs(T, 0)
    assert rel_l2(traj.spline_loop_latents(palette[:8], T, 1) @ R, g["spline_proj"]) < 2e-6
    assert rel_l2(traj.spline_loop_latents(palette[8:13], T, 2.37) @ R, g["spline_loops_proj"]) < 2e-6
    rng = torch.Generator("cpu").manual_seed(5)
    lat = traj.spline_loop_latents(palette[:6], T)
    for kind, merge, depth, sf in (("segmentation", "average", "low", "mfcc"), ("feature", "modulate", "midhigh", "chromagram"),
                                   ("feature", "modulate", "all", "onsets"), ("loop", "overwrite", "high", "tonnetz")):
        lat = traj.latent_patch(rng, lat, palette, segs, feats, 120.0, 24, kind, 6, 8, sf, 0.8, "rms", 0.9, merge, depth)
    assert rel_l2(lat @ R, g["latent_patch_proj"]) < 2e-6
    x = synth_features(2 * 300, 17, 77).abs().reshape(2, 300, 17) + 0.1
    torch.manual_seed(0)
    draws = [torch.randn(2, 300, s, s) for s in (4, 8, 16, 32)]
    dl, dn = traj.fixed_decoder(palette[:9], x, 3, 3, draws)
    assert rel_l2(dl @ R, g["dec_proj"]) < 2e-6
    for i, n in enumerate(dn):
        assert rel_l2(n[:, ::10], g[f"dec_noise_{i}"]) < 2e-6, i


def test_product_sampler_equals_oracle_sampler_cpu_generator():
    """ssar_b200's deferred Patch sampler (multinomial fast path restated as one exponential_ draw per choice) yields the
    oracle's configs — which use the real Tensor.multinomial like patch.py:12-20 — value for value (CPU generator)."""
    import torch
    from ssar_b200.random.patch import Patch as ProductPatch
    from ssar_oracle import traj

    T = 64
    dims = dict(chromagram=12, tonnetz=6, mfcc=20, spectral_contrast=7, rms=1, drop_strength=1, onsets=1, spectral_flatness=1)
    feats = {k: torch.rand(T, c) for k, c in dims.items()}
    segs = {(f, k): torch.randint(0, k, (T,)) for f in feats for k in (2, 4, 6, 8, 12, 16)}
    for seed in range(40):
        a, b = ProductPatch(feats, segs, 120.0, seed=seed, device="cpu"), traj.Patch(feats, segs, 120.0, seed=seed)
        assert (a.num_base_latents, a.sigma_base_noise, a.loops_base_noise) == (b.num_base_latents, b.sigma_base_noise, b.loops_base_noise)
        assert len(a.latent_patches) == len(b.latent_patches) and len(a.noise_patches) == len(b.noise_patches)
        for x, y in zip(a.latent_patches + a.noise_patches, b.latent_patches + b.noise_patches):
            assert all(x[k] == v for k, v in y.items())
