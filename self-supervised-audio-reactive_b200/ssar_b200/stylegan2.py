"""Mirror of the noise-delivery part of maua/maua/GAN/wrappers/stylegan2.py: `make_noise_pyramid` (:193-210) and the
synthesizer's resize of a noise input whose size does not match the layer (:86-96).  The generator itself is out of
scope: layer sizes are arguments (default: the noise_const sizes of a 1024 px StyleGAN2 synthesis network)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _cabi as cabi
from ._cabi import SsarError, i32, ptr

# noise_const sizes of the 17 synthesis layers of a 1024 px StyleGAN2 (b4.conv1, b8.conv0, b8.conv1, ..., b1024.conv1)
LAYER_SIZES_1024 = [4] + [s for r in range(3, 11) for s in (2 ** r, 2 ** r)]


def resize_noise(noise: torch.Tensor, sizes, normalize: bool = False):
    """noise [T, 1, H, W] (CUDA) -> list of [T, 1, h, w], one per (h, w) in sizes, all from ONE launch
    (`ssar_noise_pyramid`): bicubic, align_corners=False; normalize: every frame / its own unbiased std."""
    if not noise.is_cuda:
        raise SsarError("ssar_b200 needs CUDA tensors (no CPU fallback)")
    if noise.dim() != 4 or noise.shape[1] != 1:
        raise SsarError(f"expected noise [T, 1, H, W], got {tuple(noise.shape)}")
    noise = noise.float().contiguous()
    T, _, H, W = noise.shape
    outs = []
    for c0 in range(0, len(sizes), 16):
        part = [(int(h), int(w)) for h, w in sizes[c0 : c0 + 16]]
        o = [torch.empty(T, 1, h, w, dtype=torch.float32, device=noise.device) for h, w in part]
        n = len(part)
        hs, ws = (C.c_int * n)(*[h for h, _ in part]), (C.c_int * n)(*[w for _, w in part])
        ps = (C.c_void_p * n)(*[t.data_ptr() for t in o])
        cabi.call("ssar_noise_pyramid", ptr(noise), i32(T), i32(H), i32(W), i32(n), hs, ws, ps, i32(1 if normalize else 0),
                  nbytes=cabi.nbytes_of(noise, *o))
        outs += o
    return outs


def make_noise_pyramid(noise: torch.Tensor, layer_limit: int = 8, layer_sizes=None):
    """stylegan2.py:193-210: {"noise{l}": [T, 1, h_l, w_l]} for the synthesis layers 1 .. layer_limit + 1 (the first,
    4 x 4 layer is skipped like `self.layer_names[1:]`), bicubic-resized and per-frame std-normalised.  The result stays
    on the GPU (the reference moves each level to the host: its generator call moves it back, :89)."""
    sizes = list(LAYER_SIZES_1024 if layer_sizes is None else layer_sizes)[1:]
    sizes = [(s, s) if isinstance(s, int) else tuple(s) for s in sizes[: layer_limit + 1]]
    return {f"noise{l}": t for l, t in enumerate(resize_noise(noise, sizes, normalize=True))}
