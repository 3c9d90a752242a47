"""Mirror of ssar/random/noise.py — lazy noise modules; `forward(i, b)` compiles the module tree into a short program
evaluated by ONE kernel launch per call (csrc/traj_noise.cu)."""
from __future__ import annotations

import torch

from .. import ops, plan
from .._cabi import NoiseOp

LOOP, BLEND, MULTIPLY, BUFFER, AVERAGE, MODULATE, SCALEBIAS = range(7)
_LEAVES = (LOOP, BLEND, MULTIPLY, BUFFER)


def _op(kind, noise=None, mod=None, M=0, p0=0.0, p1=0.0, to_acc=0):
    o = NoiseOp()
    o.kind, o.M, o.to_acc = kind, M, to_acc
    o.noise = noise.data_ptr() if noise is not None else None
    o.mod = mod.data_ptr() if mod is not None else None
    o.p0, o.p1 = float(p0), float(p1)
    return o


class Noise:
    """noise.py:4-8.  Plain object (not an nn.Module): buffers are ordinary tensors moved with `.to(device)`."""

    _buffers = ()

    def __init__(self, length, size):
        self.length = length
        self.size = size

    def _own_to(self, device):
        for name in self._buffers:
            t = getattr(self, name)
            if t.device != torch.device(device) or t.dtype != torch.float32 or not t.is_contiguous():
                setattr(self, name, t.to(device).float().contiguous())
        return self

    def to(self, device):
        self._own_to(device)
        self.__dict__.pop("_child_device", None)
        for child in ("left", "right", "base"):
            if hasattr(self, child):
                getattr(self, child).to(device)
        return self

    @property
    def device(self):
        for name in self._buffers:
            return getattr(self, name).device
        d = self.__dict__.get("_child_device")
        if d is None:
            for child in ("left", "right", "base"):
                if hasattr(self, child):
                    d = self._child_device = getattr(self, child).device
                    break
            else:
                raise RuntimeError("noise module without buffers")
        return d

    # ---- program compilation -------------------------------------------------------------------
    def _leaf(self, to_acc, i, b, keep):
        raise NotImplementedError

    def _emit(self, prog, i, b, keep):
        """Appends ops that leave this module's value in `acc`."""
        prog.append(self._leaf(1, i, b, keep))

    def _emit_cur(self, prog, i, b, keep):
        """Appends ONE op that leaves this module's value in `cur` (composite modules are pre-rendered to a buffer)."""
        if type(self)._emit is Noise._emit:
            prog.append(self._leaf(0, i, b, keep))
        else:
            buf = self.forward(i, b).reshape(b, -1)
            keep.append(buf)
            prog.append(_op(BUFFER, noise=buf))

    def forward(self, i, b):
        b = max(0, min(int(b), self.length - int(i)))
        prog, keep = [], []
        self._emit(prog, int(i), b, keep)
        return ops.noise_eval(prog, int(i), b, int(self.size[0]), int(self.size[1]), self.device)

    __call__ = forward


class Blend(Noise):
    """noise.py:11-24."""

    _buffers = ("noise", "modulator")

    def __init__(self, rng, length, size, modulator):
        super().__init__(length, size)
        self.noise = torch.randn((2, modulator.shape[1], size[0], size[1]), generator=rng, device=rng.device)
        self.modulator = modulator
        self._own_to(modulator.device)

    def _leaf(self, to_acc, i, b, keep):
        return _op(BLEND, self.noise, self.modulator, M=self.modulator.shape[1], to_acc=to_acc)


class Multiply(Noise):
    """noise.py:27-39."""

    _buffers = ("noise", "modulator")

    def __init__(self, rng, length, size, modulator):
        super().__init__(length, size)
        self.noise = torch.randn((modulator.shape[1], size[0], size[1]), generator=rng, device=rng.device)
        self.modulator = modulator
        self._own_to(modulator.device)

    def _leaf(self, to_acc, i, b, keep):
        return _op(MULTIPLY, self.noise, self.modulator, M=self.modulator.shape[1], to_acc=to_acc)


class Loop(Noise):
    """noise.py:42-53."""

    _buffers = ("noise", "idx")

    def __init__(self, rng, length, size, n_loops=1, sigma=5, device=None):
        super().__init__(length, size)
        self.sigma = sigma
        self.noise = torch.randn((3, size[0], size[1]), generator=rng, device=rng.device)
        if device is not None:
            self.idx = plan.loop_phase(n_loops, int(length), str(device))  # cached, read-only
            self._own_to(device)
        else:
            self.idx = torch.linspace(0, n_loops * 2 * torch.pi, length)

    def _leaf(self, to_acc, i, b, keep):
        return _op(LOOP, self.noise, self.idx, p0=self.sigma / 50, to_acc=to_acc)


class Average(Noise):
    """noise.py:56-63."""

    def __init__(self, left, right):
        super().__init__(left.length, left.size)
        self.left, self.right = left, right

    def _emit(self, prog, i, b, keep):
        self.left._emit(prog, i, b, keep)
        self.right._emit_cur(prog, i, b, keep)
        prog.append(_op(AVERAGE))


class Modulate(Noise):
    """noise.py:66-75."""

    _buffers = ("modulator",)

    def __init__(self, left, right, modulator, _mean=None):
        super().__init__(left.length, left.size)
        self.left, self.right = left, right
        # _mean: modulator.mean(1) computed once by noise_patch for all the layers it wraps
        self.modulator = modulator.float().mean(1) if _mean is None else _mean
        if _mean is None:
            self._own_to(left.device)

    def _emit(self, prog, i, b, keep):
        self.left._emit(prog, i, b, keep)
        self.right._emit_cur(prog, i, b, keep)
        prog.append(_op(MODULATE, mod=self.modulator))


class ScaleBias(Noise):
    """noise.py:78-86."""

    def __init__(self, base, scale, bias):
        super().__init__(base.length, base.size)
        self.base, self.scale, self.bias = base, scale, bias

    def _emit(self, prog, i, b, keep):
        self.base._emit(prog, i, b, keep)
        prog.append(_op(SCALEBIAS, p0=self.scale, p1=self.bias))


NOISE_LAYER_RANGES = {"low": range(0, 6), "mid": range(6, 12), "high": range(12, 17), "lowmid": range(0, 12),
                      "midhigh": range(6, 17), "all": range(0, 17)}


def noise_patch(rng, noise, features, tempo, fps, patch_type, loop_bars, seq_feat, seq_feat_weight, mod_feat,
                mod_feat_weight, merge_type, merge_depth, noise_mean, noise_std):
    """ssar/random/noise.py:89-140 (same RNG draw order: one basis draw per wrapped layer, in layer order)."""
    layers = NOISE_LAYER_RANGES[merge_depth]
    if len(layers) == 0:
        return noise
    # everything that does not depend on the layer is computed once per sub-patch (the reference recomputes the scaled
    # features and the modulator mean for each wrapped layer: same values)
    dev = noise[layers[0]].device
    feature = (seq_feat_weight * features[seq_feat]).to(dev)
    if feature.dtype != torch.float32 or not feature.is_contiguous():
        feature = feature.float().contiguous()
    length = len(feature)
    if merge_type == "modulate":
        modulator = (mod_feat_weight * features[mod_feat]).to(dev)
        mod_mean = modulator.float().mean(1).contiguous()
    if patch_type == "loop":
        n_loops = length / fps / 60 / tempo / 4 / loop_bars
    for n in layers:
        if noise[n].device != dev:  # mixed-device module lists: fall back to per-layer placement
            return _noise_patch_per_layer(rng, noise, features, tempo, fps, patch_type, loop_bars, seq_feat, seq_feat_weight,
                                          mod_feat, mod_feat_weight, merge_type, layers, noise_mean, noise_std, start=n)
        if patch_type == "blend":
            new_noise = Blend(rng=rng, length=length, size=noise[n].size, modulator=feature)
        elif patch_type == "multiply":
            new_noise = Multiply(rng=rng, length=length, size=noise[n].size, modulator=feature)
        elif patch_type == "loop":
            new_noise = Loop(rng=rng, length=length, size=noise[n].size, n_loops=n_loops, device=dev)
        if merge_type == "average":
            noise[n] = Average(left=noise[n], right=new_noise)
        elif merge_type == "modulate":
            noise[n] = Modulate(left=noise[n], right=new_noise, modulator=modulator, _mean=mod_mean)
        else:
            noise[n] = new_noise
        noise[n] = ScaleBias(noise[n], scale=noise_std, bias=noise_mean)
    return noise


def _noise_patch_per_layer(rng, noise, features, tempo, fps, patch_type, loop_bars, seq_feat, seq_feat_weight, mod_feat,
                           mod_feat_weight, merge_type, layers, noise_mean, noise_std, start):
    """noise.py:89-140 layer by layer (each layer on its own device), from layer `start` of `layers` on."""
    feature = seq_feat_weight * features[seq_feat]
    for n in layers:
        if n < start:
            continue
        dev = noise[n].device
        if patch_type == "blend":
            new_noise = Blend(rng=rng, length=len(feature), size=noise[n].size, modulator=feature.to(dev))
        elif patch_type == "multiply":
            new_noise = Multiply(rng=rng, length=len(feature), size=noise[n].size, modulator=feature.to(dev))
        elif patch_type == "loop":
            n_loops = len(feature) / fps / 60 / tempo / 4 / loop_bars
            new_noise = Loop(rng=rng, length=len(feature), size=noise[n].size, n_loops=n_loops, device=dev)
        if merge_type == "average":
            noise[n] = Average(left=noise[n], right=new_noise)
        elif merge_type == "modulate":
            noise[n] = Modulate(left=noise[n], right=new_noise, modulator=(mod_feat_weight * features[mod_feat]).to(dev))
        else:
            noise[n] = new_noise
        noise[n] = ScaleBias(noise[n], scale=noise_std, bias=noise_mean)
    return noise


class NoiseProgram(Noise):
    """One layer's noise as the flat op list the module tree of noise.py:4-140 compiles to.  `Patch.forward` builds these
    directly (the trees it makes are chains: every sub-patch wraps the layer's previous value with ONE new leaf and a
    merge), which keeps the per-layer host cost at the generator draw plus a few list appends.  Same call surface as the
    modules: `forward(i, b) -> [b, H, W]`, `.size`, `.length`, `.device`, `.to(device)`."""

    def __init__(self, length, size, device):
        super().__init__(length, size)
        self._device = torch.device(device)
        self.steps = []  # (kind, noise tensor | None, mod tensor | None, M, p0, p1, to_acc)

    @property
    def device(self):
        return self._device

    def to(self, device):
        device = torch.device(device)
        if device != self._device:
            self.steps = [(k, n.to(device) if n is not None else None, m.to(device) if m is not None else None, M, p0, p1, ta)
                          for (k, n, m, M, p0, p1, ta) in self.steps]
            self._device = device
            self._compiled = None
        return self

    def leaf(self, kind, noise, mod, M=0, p0=0.0, to_acc=0):
        if to_acc:
            self.steps.clear()  # a leaf into `acc` replaces the layer's history (merge_type other than average / modulate)
        self.steps.append((kind, noise, mod, M, p0, 0.0, to_acc))
        self._compiled = None

    def merge(self, kind, mod=None, p0=0.0, p1=0.0):
        self.steps.append((kind, None, mod, 0, p0, p1, 0))
        self._compiled = None

    def forward(self, i, b):
        b = max(0, min(int(b), self.length - int(i)))
        # the op list does not depend on (i, b): its host array and device copy are made once per program, so a sweep of
        # forward(i, 32) over a track costs two launches per call and no host-to-device copy
        comp = getattr(self, "_compiled", None)
        if comp is None or comp[0] != len(self.steps) or comp[1] != self._device:
            prog = [_op(k, noise=n, mod=m, M=M, p0=p0, p1=p1, to_acc=ta) for (k, n, m, M, p0, p1, ta) in self.steps]
            comp = self._compiled = (len(self.steps), self._device, ops.noise_program(prog, self._device))
        return ops.noise_eval(comp[2], int(i), b, int(self.size[0]), int(self.size[1]), self._device)

    __call__ = forward


def noise_draw_shapes(sizes, noise_patches, features):
    """Shapes of every basis draw `base_noise_programs` + `noise_patch_programs` make, in draw order (patch.py:123-131,
    noise.py:89-140) — lets `Patch.forward` issue all of them as ONE batched Philox launch."""
    shapes = [(3, int(h), int(w)) for (h, w) in sizes]
    for sp in noise_patches:
        M = features[sp["seq_feat"]].shape[1]
        for n in NOISE_LAYER_RANGES[sp["merge_depth"]]:
            h, w = int(sizes[n][0]), int(sizes[n][1])
            shapes.append((2, M, h, w) if sp["patch_type"] == "blend" else (M, h, w) if sp["patch_type"] == "multiply"
                          else (3, h, w))
    return shapes


class LazyNoisePrograms:
    """The 17 per-layer noise programs of one Patch, built on first use.  `Patch.forward` draws every basis eagerly (one
    batched Philox launch: the generator ends in the reference's state) and hands the recipe — the sub-patch configs, the
    bound features and the drawn buffer — to this object; the per-layer op lists and the per-sub-patch modulator
    scalings (ssar/random/noise.py:89-140) materialise when a layer is first evaluated or inspected, like the lazy
    `forward(i, b)` of the reference's modules.  Building them costs ~0.5 ms of Python per Patch, which a batch pipeline
    that only consumes latents (or evaluates a few layers) should not pay up front."""

    def __init__(self, views_fn, length, sizes, n_loops, sigma, device, noise_patches, features, tempo, fps):
        self._args = (views_fn, length, sizes, n_loops, sigma, device, [dict(sp) for sp in noise_patches], features, tempo, fps)
        self._progs = None
        self._lock = __import__("threading").Lock()

    def programs(self):
        with self._lock:
            if self._progs is None:
                views_fn, length, sizes, n_loops, sigma, device, patches, feats, tempo, fps = self._args
                with torch.cuda.device(device):
                    draws = iter(views_fn())
                    progs = base_noise_programs(None, length, sizes, n_loops, sigma, device, draws=draws)
                    for sp in patches:
                        progs = noise_patch_programs(None, progs, feats, tempo, fps, draws=draws, **sp)
                self._progs, self._args = progs, None
            return self._progs


class LazyNoiseProgram(NoiseProgram):
    """Layer `index` of a `LazyNoisePrograms` recipe: a `NoiseProgram` whose `steps` appear on first access."""

    def __init__(self, owner, index, length, size, device):
        Noise.__init__(self, length, size)
        self._device = torch.device(device)
        self._owner, self._index, self._steps = owner, index, None

    @property
    def steps(self):
        if self._steps is None:
            self._steps = self._owner.programs()[self._index].steps
        return self._steps

    @steps.setter
    def steps(self, value):
        self._steps = value


def base_noise_programs(rng, length, sizes, n_loops, sigma, device, draws=None):
    """patch.py:123-131: one `Loop` per layer (same draw per layer: randn(3, H, W)).
    draws: iterator over pre-drawn bases (batched Philox launch, same values) instead of one torch.randn per layer."""
    progs = []
    for size in sizes:
        p = NoiseProgram(length, size, device)
        noise = next(draws) if draws is not None else torch.randn((3, size[0], size[1]), generator=rng, device=rng.device)
        if noise.device != p.device:
            noise = noise.to(p.device)
        p.leaf(LOOP, noise, plan.loop_phase(n_loops, int(length), str(device)), p0=sigma / 50, to_acc=1)
        progs.append(p)
    return progs


def noise_patch_programs(rng, progs, features, tempo, fps, patch_type, loop_bars, seq_feat, seq_feat_weight, mod_feat,
                         mod_feat_weight, merge_type, merge_depth, noise_mean, noise_std, draws=None):
    """ssar/random/noise.py:89-140 on `NoiseProgram`s: same generator draws (one basis per wrapped layer, in layer order),
    same arithmetic (the kernel evaluates the same ops the module tree would emit)."""
    layers = NOISE_LAYER_RANGES[merge_depth]
    if len(layers) == 0:
        return progs
    dev = progs[layers[0]].device
    feature = (seq_feat_weight * features[seq_feat]).to(dev)
    if feature.dtype != torch.float32 or not feature.is_contiguous():
        feature = feature.float().contiguous()
    length, M = feature.shape[0], feature.shape[1]
    mod_mean = None
    if merge_type == "modulate":
        mod_mean = (mod_feat_weight * features[mod_feat]).to(dev).float().mean(1).contiguous()
    if patch_type == "loop":
        phase = plan.loop_phase(length / fps / 60 / tempo / 4 / loop_bars, int(length), str(dev))
    merged = merge_type in ("average", "modulate")
    rdev = rng.device if rng is not None else None
    for n in layers:
        p = progs[n]
        h, w = p.size
        if draws is not None:
            noise = next(draws)
        elif patch_type == "blend":
            noise = torch.randn((2, M, h, w), generator=rng, device=rdev)
        elif patch_type == "multiply":
            noise = torch.randn((M, h, w), generator=rng, device=rdev)
        else:
            noise = torch.randn((3, h, w), generator=rng, device=rdev)
        if noise.device != dev:
            noise = noise.to(dev)
        to_acc = 0 if merged else 1
        if patch_type == "blend":
            p.leaf(BLEND, noise, feature, M=M, to_acc=to_acc)
        elif patch_type == "multiply":
            p.leaf(MULTIPLY, noise, feature, M=M, to_acc=to_acc)
        else:
            p.leaf(LOOP, noise, phase, p0=5 / 50, to_acc=to_acc)  # Loop's default sigma = 5 (noise.py:45)
        if merge_type == "average":
            p.merge(AVERAGE)
        elif merge_type == "modulate":
            p.merge(MODULATE, mod=mod_mean)
        p.merge(SCALEBIAS, p0=noise_std, p1=noise_mean)
    return progs
