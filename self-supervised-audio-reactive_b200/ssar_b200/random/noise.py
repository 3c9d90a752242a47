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
