"""Batched generator draws (csrc/philox_draws.cu) against the per-call torch draws they replace — bit for bit, values
and generator offsets (ssar/random/patch.py:12-66, 109-131; ssar/random/noise.py:15,31,46 issue one torch call per
draw on the seeded generator)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda", 0)


def test_normal_segments_equal_torch_randn(dev):
    from ssar_b200 import ops

    shapes = [(3, 1, 1), (3, 2, 2), (3, 4, 4), (7, 8, 8), (2, 12, 16, 16), (20, 32, 32), (3, 64, 64), (2, 6, 128, 128),
              (3, 256, 256), (2, 20, 256, 256), (1,), (255,), (256,), (257,), (303104,), (303105,), (4 * 303104 + 1,)]
    for seed, off in [(42, 0), (7, 4), (2 ** 40 + 3, 1000), (0, 2 ** 33)]:
        g = torch.Generator(dev)
        g.manual_seed(seed)
        g.set_offset(off)
        ref = [torch.randn(s, generator=g, device=dev) for s in shapes]
        g2 = torch.Generator(dev)
        g2.manual_seed(seed)
        g2.set_offset(off)
        got = ops.philox_normal_segments(shapes, g2)
        assert g2.get_offset() == g.get_offset()
        for s, a, b in zip(shapes, ref, got):
            assert a.shape == b.shape and torch.equal(a, b), (seed, off, s)
    # more segments than one launch carries
    many = [(3, 4, 4)] * 250
    g = torch.Generator(dev)
    g.manual_seed(5)
    ref = [torch.randn(s, generator=g, device=dev) for s in many]
    g2 = torch.Generator(dev)
    g2.manual_seed(5)
    got = ops.philox_normal_segments(many, g2)
    assert g2.get_offset() == g.get_offset() and all(torch.equal(a, b) for a, b in zip(ref, got))
    assert ops.philox_batched_ok(dev)


def test_patch_sampler_kernel_equals_eager_draws(dev):
    from ssar_b200.random import patch as P

    ks = [2, 4, 6, 8, 12, 16]
    seeds = list(range(40)) + [2 ** 31 + 11, 2 ** 45 + 1, 123456789]
    got = P._KernelSampler.get(dev).sample_many(seeds, ks, 5, 20)
    for seed, g in zip(seeds, got):
        rng = torch.Generator(dev)
        rng.manual_seed(seed)
        ref = P._eager_sample(rng, ks, 5, 20)
        assert ref == g[:5], seed
        assert rng.get_offset() == g[5], seed
    assert P._KernelSampler.get(dev).ok(ks)
    # other ks / sub-patch ranges
    got = P._KernelSampler.get(dev).sample_many([3, 4], [3, 5], 1, 4)
    for seed, g in zip([3, 4], got):
        rng = torch.Generator(dev)
        rng.manual_seed(seed)
        assert P._eager_sample(rng, [3, 5], 1, 4) == g[:5]


def test_patch_paths_agree(dev):
    """Patch(...) / Patch.forward with the batched draws and the one-launch sub-patch stack == the per-call CUDA draws and
    one launch per sub-patch (configs, latents bit for bit, bases, generator state)."""
    from ssar_b200.random import patch as P
    from ssar_b200.synth import synth_patch_inputs

    T = 600
    feats, segs, palette = synth_patch_inputs(T, 0)
    feats = {k: v.to(dev) for k, v in feats.items()}
    palette = palette.to(dev)

    def run(kernel, batched, seed):
        P.USE_SAMPLER_KERNEL, P.USE_BATCHED_DRAWS, P.USE_LATENT_STACK = kernel, batched, batched
        try:
            p = P.Patch(feats, segs, 120.0, fps=24, seed=seed, device=dev)
            off0 = p.rng.get_offset()
            lat, noise = p.forward(palette, downscale_factor=8)
            return p, off0, lat, noise, p.rng.get_offset()
        finally:
            P.USE_SAMPLER_KERNEL = P.USE_BATCHED_DRAWS = P.USE_LATENT_STACK = True

    for seed in (42, 7, 1001, 5, 6, 8, 9, 10):
        a = run(True, True, seed)
        b = run(False, False, seed)
        assert a[0].latent_patches == b[0].latent_patches and a[0].noise_patches == b[0].noise_patches
        assert (a[0].num_base_latents, a[0].sigma_base_noise, a[0].loops_base_noise) == \
               (b[0].num_base_latents, b[0].sigma_base_noise, b[0].loops_base_noise)
        assert a[1] == b[1] and a[4] == b[4]
        assert torch.equal(a[2], b[2])
        for na, nb in zip(a[3], b[3]):
            assert len(na.steps) == len(nb.steps)
            for sa, sb in zip(na.steps, nb.steps):
                if sa[1] is not None:
                    assert torch.equal(sa[1], sb[1])
            # same program, same bases: equal up to the order of the kernel's atomic RMS accumulation
            assert torch.allclose(na.forward(3, 5), nb.forward(3, 5), rtol=1e-5, atol=1e-6, equal_nan=True)
    # sample_batch == one Patch per seed
    batch = P.Patch.sample_batch([42, 43, 44], segs, 120.0, fps=24, device=dev)
    for s, pb in zip([42, 43, 44], batch):
        P.USE_SAMPLER_KERNEL = False
        try:
            one = P.Patch(None, segs, 120.0, fps=24, seed=s, device=dev)
        finally:
            P.USE_SAMPLER_KERNEL = True
        assert pb.latent_patches == one.latent_patches and pb.noise_patches == one.noise_patches
        assert pb.rng.get_offset() == one.rng.get_offset()


def test_randperm_equals_torch(dev):
    from ssar_b200 import ops

    for n in (180, 7, 64, 1000):
        stride = ops.philox_randperm_stride(dev, n)
        assert stride is not None, n
        for seed, off in [(42, 0), (2 ** 35 + 9, 184 * 7)]:
            g = torch.Generator(dev)
            g.manual_seed(seed)
            g.set_offset(off)
            ref = torch.stack([torch.randperm(n, generator=g, device=dev) for _ in range(300)])
            g2 = torch.Generator(dev)
            g2.manual_seed(seed)
            g2.set_offset(off)
            got = ops.philox_randperm(g2, n, 300, stride)
            assert g2.get_offset() == g.get_offset()
            assert torch.equal(ref, got), (n, seed)
