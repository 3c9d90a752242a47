"""Mirror of ssar/random/patch.py — the seeded random-patch sampler and its forward pass.

All random draws stay in torch host code in the reference's order (patch.py:41-66, 109-131; noise.py:15,31,46) so a
given seed yields the same patch; the arithmetic runs in the trajectory kernels."""
from __future__ import annotations

import json

import numpy as np
import torch

from . import DEVICE
from .latent import LatentWeights, latent_patch, merge_coefficients, sequence_weights, spline_loop_latents  # noqa: F401
from .mir import ALLFEATS, UNITFEATS
from .noise import (LazyNoiseProgram, LazyNoisePrograms, Loop, base_noise_programs, noise_draw_shapes,  # noqa: F401
                    noise_patch, noise_patch_programs)
from .. import ops, plan


def random_choice(rng, options, weights=None, n=1, replacement=False):
    """patch.py:12-20."""
    if weights is None:
        probabilities = torch.ones(len(options), device=rng.device) / len(options)
    else:
        probabilities = torch.tensor(weights, device=rng.device) / np.sum(weights)
    idx = probabilities.multinomial(num_samples=n, replacement=replacement, generator=rng)
    return options[idx]


def skewnorm(rng, a, loc, scale, size=()):
    """patch.py:23-32."""
    u0 = torch.randn(size, generator=rng, device=rng.device)
    v = torch.randn(size, generator=rng, device=rng.device)
    d = a / np.sqrt(1 + a ** 2)
    u1 = d * u0 + v * np.sqrt(1 - d ** 2)
    return loc + scale * torch.where(u0 >= 0, u1, -u1)


class _Deferred:
    """A draw whose value is still on the generator's device; `Patch._resolve` replaces it after ONE bulk copy."""

    __slots__ = ("kind", "slot", "aux")

    def __init__(self, kind, slot, aux=None):
        self.kind, self.slot, self.aux = kind, slot, aux


class _Sampler:
    """Issues exactly the generator calls of patch.py:12-32 / :41-66 in the reference's order, but leaves the drawn
    values on the generator's device and converts them to Python objects with one device-to-host copy at the end
    (a CUDA generator otherwise costs one stream synchronisation per draw: ~400 per Patch)."""

    def __init__(self, rng, caches=None):
        self.rng = rng
        self.floats, self._n_floats = [], 0
        self._uniform, self._weighted, self._host_probs = caches if caches is not None else ({}, {}, {})

    def _probs(self, n_options, weights):
        if weights is None:
            p = self._uniform.get(n_options)
            if p is None:
                p = self._uniform[n_options] = torch.ones(n_options, device=self.rng.device) / n_options
            return p
        key = tuple(weights)
        p = self._weighted.get(key)
        if p is None:
            p = self._weighted[key] = torch.tensor(weights, device=self.rng.device) / np.sum(weights)
        return p

    def choice(self, options, weights=None):
        """random_choice (patch.py:12-20) with n = 1.  `Tensor.multinomial(1)` draws q ~ Exp(1) with ONE
        `exponential_` call on a tensor shaped like the probabilities and returns argmax(p / q) (ATen
        native/Distributions.cpp multinomial_out, the no-replacement fast path; its other ~10 ops are input checks).
        Only that generator call is issued here; the fp32 division and the arg-max are done after the bulk copy.
        tests: the sampler equals the eager `random_choice` path draw for draw (CPU and CUDA generators)."""
        p = self._probs(len(options), weights)
        self.floats.append(torch.empty_like(p).exponential_(1, generator=self.rng))
        self._n_floats += len(options)
        return _Deferred("choice", self._n_floats - len(options), (options, p, weights))

    def randn(self, scale=1.0):
        self.floats.append(torch.randn((), generator=self.rng, device=self.rng.device).reshape(1))
        self._n_floats += 1
        return _Deferred("randn", self._n_floats - 1, scale)

    def skewnorm(self, a, loc, scale):
        """skewnorm (patch.py:23-32): the two normal draws now, the fp32 arithmetic after the copy."""
        self.floats.append(torch.randn((), generator=self.rng, device=self.rng.device).reshape(1))
        self.floats.append(torch.randn((), generator=self.rng, device=self.rng.device).reshape(1))
        self._n_floats += 2
        return _Deferred("skewnorm", self._n_floats - 2, (a, loc, scale))

    def flat(self):
        """All draws so far as one fp32 vector on the generator's device (slot order)."""
        return torch.cat(self.floats)

    def resolve(self, tree, floats=None):
        """tree of `_Deferred` -> Python values; floats: host copy of `flat()` (fetched here when None)."""
        if floats is None:
            floats = self.flat().cpu().numpy() if self.floats else None
        f = np.float32

        def host_probs(p, weights, n_options):
            key = (n_options, None if weights is None else tuple(weights))
            h = self._host_probs.get(key)
            if h is None:
                if p is None:  # dry sampler: the same fp32 expressions as `_probs`, evaluated by torch on the host
                    p = torch.ones(n_options) / n_options if weights is None else torch.tensor(weights) / np.sum(weights)
                h = self._host_probs[key] = p.cpu().numpy()
            return h

        def value(d):
            if d.kind == "choice":
                options, p, weights = d.aux
                q = floats[d.slot : d.slot + len(options)]
                return options[int(np.argmax(host_probs(p, weights, len(options)) / q))]  # fp32 IEEE division, first maximum
            if d.kind == "randn":
                return float(floats[d.slot]) * d.aux
            a, loc, scale = d.aux
            u0, v = f(floats[d.slot]), f(floats[d.slot + 1])
            dd = a / np.sqrt(1 + a ** 2)
            # fp32 op by op, like the tensor expression d * u0 + v * sqrt(1 - d^2); loc + scale * where(u0 >= 0, u1, -u1)
            u1 = f(f(dd) * u0) + f(v * f(np.sqrt(1 - dd ** 2)))
            return float(f(loc) + f(f(scale) * (u1 if u0 >= 0 else -u1)))

        def walk(o):
            if isinstance(o, _Deferred):
                return value(o)
            if isinstance(o, dict):
                return {k: walk(v) for k, v in o.items()}
            if isinstance(o, list):
                return [walk(v) for v in o]
            return o

        return walk(tree)


def _resolve_many(sampler, tree, F):
    """`sampler.resolve(tree, F[i])` for every row of F [n, n_floats] (fp32), vectorised over the rows: `tree` is a flat
    dict key -> `_Deferred` (the config trees of patch.py:77-106).  Same fp32 expressions per value as `resolve` (IEEE
    division + first maximum for the choices, op-by-op fp32 for skewnorm), so the results are identical; the per-value
    Python of `resolve` was the largest host cost of a step (~9 000 values per 32 Patches)."""
    n = F.shape[0]
    if n == 0:
        return []
    f = np.float32
    cols = []
    for key, d in tree.items():
        if d.kind == "choice":
            options, p, weights = d.aux
            hkey = (len(options), None if weights is None else tuple(weights))
            h = sampler._host_probs.get(hkey)
            if h is None:
                pp = p
                if pp is None:
                    pp = (torch.ones(len(options)) / len(options) if weights is None
                          else torch.tensor(weights) / np.sum(weights))
                h = sampler._host_probs[hkey] = pp.cpu().numpy()
            idx = np.argmax(h[None, :] / F[:, d.slot : d.slot + len(options)], axis=1)
            cols.append([options[i] for i in idx.tolist()])
        elif d.kind == "randn":
            cols.append([float(v) * d.aux for v in F[:, d.slot].tolist()])
        else:
            a, loc, scale = d.aux
            u0, v = F[:, d.slot], F[:, d.slot + 1]
            dd = a / np.sqrt(1 + a ** 2)
            u1 = (f(dd) * u0).astype(f) + (v * f(np.sqrt(1 - dd ** 2))).astype(f)
            r = f(loc) + (f(scale) * np.where(u0 >= 0, u1, -u1)).astype(f)
            cols.append(r.astype(f).astype(np.float64).tolist())
    keys = list(tree.keys())
    return [dict(zip(keys, vals)) for vals in zip(*cols)]


class _DrySampler(_Sampler):
    """Records the draw program of a config (which torch call, on how many elements) without drawing: the program
    `ssar_patch_sample` executes on the device, and the deferred tree `resolve` turns the drawn floats into."""

    def __init__(self):
        super().__init__(None, ({}, {}, {}))
        self.ops = []

    def choice(self, options, weights=None):
        n = len(options)
        self.ops.append((1, n))  # SSAR_DRAW_EXPONENTIAL
        self._n_floats += n
        return _Deferred("choice", self._n_floats - n, (options, None, weights))

    def randn(self, scale=1.0):
        self.ops.append((0, 1))  # SSAR_DRAW_RANDN
        self._n_floats += 1
        return _Deferred("randn", self._n_floats - 1, scale)

    def skewnorm(self, a, loc, scale):
        self.ops += [(0, 1), (0, 1)]
        self._n_floats += 2
        return _Deferred("skewnorm", self._n_floats - 2, (a, loc, scale))


DEPTHS = ["low", "mid", "high", "lowmid", "midhigh", "all"]


def _latent_config(choice, skew, ks):
    """patch.py:77-91 (draw order is part of the result)."""
    return dict(
        patch_type=choice(["segmentation", "feature", "loop"]),
        segments=choice(ks),
        loop_bars=choice([4, 8, 16, 32], weights=[2, 2, 2, 1]),
        seq_feat=choice(ALLFEATS),
        seq_feat_weight=skew(5, 0.666, 0.5),
        mod_feat=choice(UNITFEATS),
        mod_feat_weight=skew(5, 0.666, 0.5),
        merge_type=choice(["average", "modulate"], weights=[1, 3]),
        merge_depth=choice(DEPTHS, weights=[3, 3, 3, 2, 2, 1]),
    )


def _noise_config(choice, skew, randn):
    """patch.py:93-106 (draw order is part of the result)."""
    return dict(
        patch_type=choice(["blend", "multiply", "loop"]),
        loop_bars=choice([4, 8, 16, 32], weights=[2, 2, 2, 1]),
        seq_feat=choice(ALLFEATS),
        seq_feat_weight=skew(5, 0.666, 0.5),
        mod_feat=choice(UNITFEATS),
        mod_feat_weight=skew(5, 0.666, 0.5),
        merge_type=choice(["average", "modulate"], weights=[1, 3]),
        merge_depth=choice(DEPTHS, weights=[3, 3, 3, 2, 2, 1]),
        noise_mean=randn(0.5),
        noise_std=skew(5, 0.666, 0.5),
    )


class _GraphSampler:
    """CUDA-generator fast path of the Patch sampler.  The draws of n consecutive latent (or noise) sub-patch configs are
    a fixed sequence of ~11 n (14 n) tiny generator kernels; they are captured ONCE per n as a CUDA graph whose Philox
    seed / offset are re-read from the generator at every replay (`CUDAGraph.register_generator_state`), so a replay
    issues exactly the draws the eager sequence would issue from the generator's current state — one graph launch instead
    of ~250 kernel launches + ~500 Python-level torch calls per Patch.  One shared generator per device is re-seeded per
    Patch; the Patch's own generator is then set to the shared generator's final state."""

    _instances = {}
    _guard = __import__("threading").Lock()

    @classmethod
    def get(cls, device):
        device = torch.device(device)
        idx = device.index if device.index is not None else torch.cuda.current_device()
        with cls._guard:
            inst = cls._instances.get(idx)
            if inst is None:
                inst = cls._instances[idx] = cls(torch.device("cuda", idx))
            return inst

    def __init__(self, device):
        self.device = device
        self.rng = torch.Generator(device)
        self.lock = __import__("threading").RLock()
        self.caches = ({}, {}, {})
        self.blocks = {}
        self.capture_stream = None

    def _build(self, kind, n, ks):
        sm = _Sampler(self.rng, self.caches)
        if kind == "latent":
            tree = [_latent_config(sm.choice, sm.skewnorm, ks) for _ in range(n)]
        else:
            tree = [_noise_config(sm.choice, sm.skewnorm, sm.randn) for _ in range(n)]
        return sm, tree, sm.flat()

    def block(self, kind, n, ks):
        """(graph, static flat output, deferred tree, sampler) for n consecutive configs; captured on first use with the
        generator state preserved."""
        key = (kind, int(n), tuple(ks))
        blk = self.blocks.get(key)
        if blk is None:
            if n == 0:
                blk = self.blocks[key] = (None, None, [], None)
                return blk
            state = self.rng.get_state()
            if self.capture_stream is None:
                self.capture_stream = torch.cuda.Stream(self.device)
            cs = self.capture_stream
            cs.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(cs):
                self._build(kind, n, ks)  # eager warm-up: fills the probability-tensor caches outside the graph
            cs.synchronize()
            graph = torch.cuda.CUDAGraph()
            graph.register_generator_state(self.rng)
            with torch.cuda.graph(graph, stream=cs, capture_error_mode="thread_local"):
                sm, tree, flat = self._build(kind, n, ks)
            self.rng.set_state(state)
            blk = self.blocks[key] = (graph, flat, tree, sm)
        return blk

    def sample(self, seed, ks, min_subpatches, max_subpatches):
        """patch.py:41-66 -> (num_base_latents, sigma_base_noise, loops_base_noise, latent configs, noise configs, final
        generator state)."""
        rng, dev = self.rng, self.device
        with self.lock:
            rng.manual_seed(seed)
            num_base = torch.randint(3, 15, size=(), generator=rng, device=dev)
            sigma_u = torch.rand((), generator=rng, device=dev)
            head = _Sampler(rng, self.caches)
            loops = head.choice([1, 2, 4, 8, 16, 32, 64])
            n_lat = int(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=dev))
            g_l, flat_l, tree_l, sm_l = self.block("latent", n_lat, ks)
            if g_l is not None:
                g_l.replay()
            n_noi = int(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=dev))
            g_n, flat_n, tree_n, sm_n = self.block("noise", n_noi, ks)
            if g_n is not None:
                g_n.replay()
            parts = [head.flat(), num_base.float().reshape(1), sigma_u.reshape(1)]
            parts += [t for t in (flat_l, flat_n) if t is not None]
            host = torch.cat(parts).cpu().numpy()  # the one bulk copy (static graph outputs are read before the next replay)
            state = rng.get_state()
        n_head = head._n_floats
        o = n_head + 2
        loops_v = head.resolve(loops, host[:n_head])
        lat = sm_l.resolve(tree_l, host[o : o + flat_l.numel()]) if flat_l is not None else []
        o += flat_l.numel() if flat_l is not None else 0
        noi = sm_n.resolve(tree_n, host[o : o + flat_n.numel()]) if flat_n is not None else []
        return int(host[n_head]), 1 + 9 * float(host[n_head + 1]), loops_v, lat, noi, state


class _KernelSampler:
    """CUDA-generator path of the Patch sampler as ONE kernel for a whole batch of seeds (`ssar_patch_sample`): the
    device walks the draw programs of patch.py:41-66 / 77-106 with the Philox streams the torch calls would use, the
    host reads everything back with one copy and resolves the configs (`_Sampler.resolve`).  Checked once per device
    against the eager torch sequence (`ok`); tests/test_gpu_philox.py checks it seed by seed."""

    _instances = {}
    _guard = __import__("threading").Lock()

    @classmethod
    def get(cls, device):
        device = torch.device(device)
        idx = device.index if device.index is not None else torch.cuda.current_device()
        with cls._guard:
            inst = cls._instances.get(idx)
            if inst is None:
                inst = cls._instances[idx] = cls(torch.device("cuda", idx))
            return inst

    def __init__(self, device):
        self.device = device
        self.programs = {}
        self._ok = None

    def program(self, ks):
        key = tuple(ks)
        prog = self.programs.get(key)
        if prog is None:
            head, lat, noi = _DrySampler(), _DrySampler(), _DrySampler()
            head_tree = head.choice([1, 2, 4, 8, 16, 32, 64])
            lat_tree = _latent_config(lat.choice, lat.skewnorm, list(ks))
            noi_tree = _noise_config(noi.choice, noi.skewnorm, noi.randn)
            ops_all = head.ops + lat.ops + noi.ops
            dev_prog = torch.tensor([k for k, _ in ops_all] + [n for _, n in ops_all], dtype=torch.int32).to(self.device)
            prog = self.programs[key] = (dev_prog, (head, head_tree), (lat, lat_tree), (noi, noi_tree))
        return prog

    def sample_many(self, seeds, ks, min_subpatches, max_subpatches):
        """-> list of (num_base_latents, sigma_base_noise, loops_base_noise, latent configs, noise configs, end offset)."""
        dev_prog, (head, head_tree), (lat, lat_tree), (noi, noi_tree) = self.program(ks)
        cap = 2 + head._n_floats + (max_subpatches - 1) * (lat._n_floats + noi._n_floats)
        with torch.cuda.device(self.device):
            seeds_d = torch.tensor([int(s) for s in seeds], dtype=torch.int64).to(self.device)
            out, counts, end = ops.patch_sample(seeds_d, dev_prog, len(head.ops), len(lat.ops), len(noi.ops),
                                                min_subpatches, max_subpatches, cap)
            # ONE read-back for the whole batch
            packed = torch.cat([out.view(-1), counts.view(-1).float(), end.float()]).cpu().numpy()
        B = len(seeds)
        floats = packed[: B * cap].reshape(B, cap)
        cnt = packed[B * cap : B * cap + 3 * B].reshape(B, 3).astype(np.int64)
        ends = packed[B * cap + 3 * B :].astype(np.int64)
        # every config of the batch resolved at once (vectorised over all sub-patches of all seeds)
        n_lat, n_noi = cnt[:, 0], cnt[:, 1]
        o_lat = 2 + head._n_floats
        o_noi = o_lat + n_lat * lat._n_floats
        assert np.all(o_noi + n_noi * noi._n_floats == cnt[:, 2]) and np.all(cnt[:, 2] <= cap)

        def gather(counts, first, width):
            rows = np.repeat(np.arange(B), counts)
            within = np.concatenate([np.arange(c) for c in counts]) if counts.sum() else np.zeros(0, dtype=np.int64)
            starts = np.repeat(first, counts) + within * width
            return floats[rows[:, None], starts[:, None] + np.arange(width)[None, :]]

        heads = _resolve_many(head, {"loops": head_tree}, floats[:, 2 : 2 + head._n_floats])
        lat_all = _resolve_many(lat, lat_tree, gather(n_lat, np.full(B, o_lat), lat._n_floats))
        noi_all = _resolve_many(noi, noi_tree, gather(n_noi, o_noi, noi._n_floats))
        res, il, io = [], 0, 0
        for b in range(B):
            f = floats[b]
            lats, nois = lat_all[il : il + n_lat[b]], noi_all[io : io + n_noi[b]]
            il += n_lat[b]
            io += n_noi[b]
            res.append((int(f[0]), 1 + 9 * float(f[1]), heads[b]["loops"], lats, nois, int(ends[b])))
        return res

    def ok(self, segmentations_ks):
        """One-time check per device: the kernel's draws for two seeds equal the eager deferred sampler's."""
        if self._ok is None:
            try:
                ks = list(segmentations_ks)
                got = self.sample_many([1234567, 42], ks, 5, 20)
                good = True
                for seed, g in zip([1234567, 42], got):
                    rng = torch.Generator(self.device)
                    rng.manual_seed(seed)
                    ref = _eager_sample(rng, ks, 5, 20)
                    good = good and ref == g[:5] and rng.get_offset() == g[5]
                self._ok = bool(good)
            except Exception:  # pragma: no cover - a torch build without get_offset etc.
                self._ok = False
            if not self._ok:
                import warnings

                warnings.warn("ssar_b200: ssar_patch_sample differs from this torch build's eager sampler draws; "
                              "using the per-call sampler")
        return self._ok


def _eager_sample(rng, ks, min_subpatches, max_subpatches):
    """patch.py:41-66 with one torch call per draw (deferred read-back) -> (num_base, sigma, loops, latent, noise)."""
    sm = _Sampler(rng)
    num_base_latents = torch.randint(3, 15, size=(), generator=rng, device=rng.device)
    sigma_u = torch.rand((), generator=rng, device=rng.device)
    loops_base_noise = sm.choice([1, 2, 4, 8, 16, 32, 64])
    latent_patches = [
        _latent_config(sm.choice, sm.skewnorm, ks)
        for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device))
    ]
    noise_patches = [
        _noise_config(sm.choice, sm.skewnorm, sm.randn)
        for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device))
    ]
    loops, lats, nois = sm.resolve([loops_base_noise, latent_patches, noise_patches])
    return int(num_base_latents.item()), 1 + 9 * sigma_u.item(), loops, lats, nois


USE_SAMPLER_GRAPHS = True  # CUDA generators only; False = the eager deferred sampler
USE_BATCHED_DRAWS = True   # CUDA generators only: noise bases of a Patch in one Philox launch (ssar_philox_normal_segments)
USE_LAZY_NOISE_PROGRAMS = True  # with batched draws: per-layer noise op lists are built on first use (bases drawn eagerly)
USE_LATENT_STACK = True    # Patch.forward: all latent sub-patches of a clip in one launch (ssar_latent_stack)
USE_SAMPLER_KERNEL = True  # CUDA generators only: the config sampler as one kernel per batch of seeds (ssar_patch_sample)


def sampler_kernel_ready(device, segmentations) -> bool:
    """True when Patch sampling on `device` goes through `ssar_patch_sample` (checked against the eager draws once)."""
    if not USE_SAMPLER_KERNEL:
        return False
    ks = np.unique([k for (_, k) in segmentations]).tolist()
    return _KernelSampler.get(device).ok(ks)


def precapture_sampler_graphs(device, segmentations, min_subpatches=5, max_subpatches=20):
    """Captures every sampler graph a Patch on `device` can need (one per sub-patch count) — call before starting other
    host threads that launch CUDA work, so no capture happens while they run."""
    gs = _GraphSampler.get(device)
    ks = np.unique([k for (_, k) in segmentations]).tolist()
    with gs.lock:
        for n in range(min_subpatches, max_subpatches):
            gs.block("latent", n, ks)
            gs.block("noise", n, ks)


class Patch:
    """patch.py:35-175.  `device`: where the generator lives (as in the reference); `compute_device`: where the
    trajectory tensors live (defaults to the palette's device at forward time)."""

    def __init__(self, features, segmentations, tempo, fps=24, seed=42, min_subpatches=5, max_subpatches=20,
                 device=DEVICE, presampled=None):
        rng = torch.Generator(device)
        rng.manual_seed(seed)
        self.seed = seed
        self.rng = rng
        self.fps = fps
        self.tempo = tempo
        self.features = features
        self.length = features[list(features.keys())[0]].shape[0] if features is not None else None
        self.segmentations = segmentations
        self.ks = np.unique([k for (_, k) in segmentations]).tolist()
        if presampled is None and rng.device.type == "cuda" and USE_SAMPLER_KERNEL:
            ksamp = _KernelSampler.get(rng.device)
            if ksamp.ok(self.ks):
                presampled = ksamp.sample_many([seed], self.ks, min_subpatches, max_subpatches)[0]
        if presampled is not None:
            (self.num_base_latents, self.sigma_base_noise, self.loops_base_noise, self.latent_patches, self.noise_patches,
             end_offset) = presampled
            if rng.device.type == "cuda":
                rng.set_offset(end_offset)
            self._sampler = None
            return
        if rng.device.type == "cuda" and USE_SAMPLER_GRAPHS:
            (self.num_base_latents, self.sigma_base_noise, self.loops_base_noise, self.latent_patches, self.noise_patches,
             state) = _GraphSampler.get(rng.device).sample(seed, self.ks, min_subpatches, max_subpatches)
            rng.set_state(state)
            self._sampler = None
            return
        sm = self._sampler = _Sampler(rng)
        num_base_latents = torch.randint(3, 15, size=(), generator=rng, device=rng.device)
        sigma_u = torch.rand((), generator=rng, device=rng.device)
        loops_base_noise = sm.choice([1, 2, 4, 8, 16, 32, 64])
        latent_patches = [
            self.random_latent_patch()
            for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device))
        ]
        noise_patches = [
            self.random_noise_patch()
            for _ in range(torch.randint(min_subpatches, max_subpatches, size=(), generator=rng, device=rng.device))
        ]
        self.loops_base_noise, self.latent_patches, self.noise_patches = sm.resolve(
            [loops_base_noise, latent_patches, noise_patches])
        self.num_base_latents = num_base_latents.item()
        self.sigma_base_noise = 1 + 9 * sigma_u.item()
        self._sampler = None

    @classmethod
    def sample_batch(cls, seeds, segmentations, tempo, fps=24, min_subpatches=5, max_subpatches=20, device=DEVICE):
        """[Patch(None, segmentations, tempo, fps, seed=s, device=device) for s in seeds] with the config draws of all
        seeds made by ONE kernel and read back with ONE copy (CUDA generators; per-seed construction otherwise)."""
        device = torch.device(device)
        pres = [None] * len(seeds)
        if device.type == "cuda" and USE_SAMPLER_KERNEL and len(seeds):
            ks = np.unique([k for (_, k) in segmentations]).tolist()
            ksamp = _KernelSampler.get(device)
            if ksamp.ok(ks):
                pres = ksamp.sample_many(list(seeds), ks, min_subpatches, max_subpatches)
        return [cls(None, segmentations, tempo, fps=fps, seed=int(s), min_subpatches=min_subpatches,
                    max_subpatches=max_subpatches, device=device, presampled=p) for s, p in zip(seeds, pres)]

    def bind(self, features):
        """Attach the feature dict of a patch sampled with features=None (batched pipelines sample the configs — which
        need no feature values — while the GPU is still computing the features)."""
        self.features = features
        self.length = features[list(features.keys())[0]].shape[0]
        return self

    def update_intensity(self, val):
        """patch.py:68-75 (including the write of noise_std into latent_patches[p], as in the reference)."""
        for p in range(len(self.latent_patches)):
            self.latent_patches[p]["seq_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
            self.latent_patches[p]["mod_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
        for p in range(len(self.noise_patches)):
            self.noise_patches[p]["seq_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
            self.noise_patches[p]["mod_feat_weight"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()
            self.latent_patches[p]["noise_std"] = skewnorm(self.rng, a=5, loc=val, scale=0.5).item()

    def _draws(self):
        """Draw source for the config samplers: the deferred sampler while __init__ runs, eager draws afterwards."""
        sm = getattr(self, "_sampler", None)
        if sm is not None:
            return sm.choice, (lambda a, loc, scale: sm.skewnorm(a, loc, scale)), sm.randn
        rng = self.rng
        return ((lambda options, weights=None: random_choice(rng, options, weights)),
                (lambda a, loc, scale: skewnorm(rng, a=a, loc=loc, scale=scale).item()),
                (lambda scale=1.0: torch.randn((), generator=rng, device=rng.device).item() * scale))

    def random_latent_patch(self):
        """patch.py:77-91 (same draws in the same order)."""
        choice, skew, _ = self._draws()
        return _latent_config(choice, skew, self.ks)

    def random_noise_patch(self):
        """patch.py:93-106 (same draws in the same order)."""
        choice, skew, randn = self._draws()
        return _noise_config(choice, skew, randn)

    def forward(self, latent_palette, downscale_factor=1, aspect_ratio=1, latents_out=None, palette_split=None):
        """patch.py:108-133 -> (latents [T, 18, 512], list of 17 lazy noise modules).
        latents_out: optional preallocated contiguous [T, 18, 512] tensor to render into (batch pipelines).
        palette_split: optional `ops.SplitBasis(latent_palette)` shared by every Patch rendered against this palette."""
        self.rng.manual_seed(self.seed)
        dev = latent_palette.device
        T, P = self.length, len(latent_palette)
        feats = {}
        for k, v in self.features.items():
            v = v.to(dev)
            if v.dtype != torch.float32 or v.dim() != 2 or v.stride(1) != 1:
                v = v.float().reshape(T, -1).contiguous()
            feats[k] = v

        perms = None
        if self.rng.device.type == "cuda" and USE_BATCHED_DRAWS:
            stride = ops.philox_randperm_stride(self.rng.device, P)
            if stride is not None:
                # the base permutation and the one of every latent sub-patch (latent.py:37) in one launch
                perms = ops.philox_randperm(self.rng, P, 1 + len(self.latent_patches), stride)
        base_perm = perms[0] if perms is not None else torch.randperm(P, generator=self.rng, device=self.rng.device)
        base_selection = base_perm[: self.num_base_latents]
        acc = LatentWeights(T, P, latent_palette.shape[1], 6, dev)
        acc.deferred = USE_LATENT_STACK  # the whole sub-patch stack as one ssar_latent_stack launch
        acc.set_base(base_selection, plan.spline_loop_weights(len(base_selection), T, 1, str(dev)))
        for j, sp in enumerate(self.latent_patches):
            acc.apply_subpatch(self.rng, self.segmentations, feats, self.tempo, self.fps,
                               permutation=perms[j + 1] if perms is not None else None, **sp)
        latents = acc.render(latent_palette, out=latents_out, palette_split=palette_split)

        sizes = [(round(aspect_ratio * size / downscale_factor), round(size / downscale_factor))
                 for size in [4, 8, 8, 16, 16, 32, 32, 64, 64, 128, 128, 256, 256, 512, 512, 1024, 1024]]
        draws = None
        if (self.rng.device.type == "cuda" and USE_BATCHED_DRAWS and USE_LAZY_NOISE_PROGRAMS
                and ops.philox_batched_ok(self.rng.device)):
            # every basis drawn now (one launch, generator advanced as the reference's calls would); op lists on demand
            views_fn = ops.philox_normal_segments_deferred(noise_draw_shapes(sizes, self.noise_patches, feats), self.rng)
            owner = LazyNoisePrograms(views_fn, self.length, sizes, self.loops_base_noise, self.sigma_base_noise, dev,
                                      self.noise_patches, feats, self.tempo, self.fps)
            return latents, [LazyNoiseProgram(owner, j, self.length, sizes[j], dev) for j in range(len(sizes))]
        if self.rng.device.type == "cuda" and USE_BATCHED_DRAWS and ops.philox_batched_ok(self.rng.device):
            # every basis draw of this patch in one launch: same Philox streams as the per-layer torch.randn calls
            draws = iter(ops.philox_normal_segments(noise_draw_shapes(sizes, self.noise_patches, feats), self.rng))
        noise = base_noise_programs(self.rng, self.length, sizes, self.loops_base_noise, self.sigma_base_noise, dev,
                                    draws=draws)
        for sp in self.noise_patches:
            noise = noise_patch_programs(self.rng, noise, feats, self.tempo, self.fps, draws=draws, **sp)
        return latents, noise

    __call__ = forward

    def save(self, path):
        with open(path, mode="w") as f:
            f.write(json.dumps(dict(
                seed=self.seed, latent_patches=self.latent_patches, noise_patches=self.noise_patches,
                num_base_latents=self.num_base_latents, sigma_base_noise=self.sigma_base_noise,
                loops_base_noise=self.loops_base_noise,
            )))

    @staticmethod
    def load(path, features, segmentations, tempo, fps, device):
        patch = Patch(features, segmentations, tempo, fps, device=device)
        with open(path, mode="r") as f:
            for key, val in json.loads(f.read()).items():
                setattr(patch, key, val)
        return patch
