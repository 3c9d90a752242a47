"""Mirror of ssar/supervised/latent_augmenter.py (:14-20 spline_loop_latents twin, :23-105 LatentAugmenter).

The reference module cannot be imported even upstream (its line 8 asks supervised/data.py for a `normalize` that file does
not define; features/processing.py's is meant) and its constructor maps 16 384 random z through a StyleGAN2 checkpoint.
Here the mapped-latent bank `ws` [num, W, L] is an argument; the host-side draws (numpy's and Python's GLOBAL generators,
in the reference's order, so `np.random.seed` / `random.seed` reproduce a reference run) choose what to do, and every
latent update is one in-place mixing launch on the layer slice it touches:
    latent[:, a:b] = alpha(t) * latent[:, a:b] + beta(t) * sum_k S[t, k] * lats[k, a:b]."""
from __future__ import annotations

import random

import numpy as np
import torch

from .. import ops
from ..random.latent import spline_loop_latents as _spline_loop_latents


def spline_loop_latents(y, size):
    """latent_augmenter.py:14-20 (one loop)."""
    return _spline_loop_latents(y, size, 1)


class LatentAugmenter:
    FEAT_IDXS = {"chroma": (20, 32), "tonnetz": (32, 38), "onsets": (46, 47), "onsets_low": (47, 48), "onsets_mid": (48, 49),
                 "onsets_high": (49, 50), "volume": (51, 52), "volume_low": (52, 53), "volume_mid": (53, 54),
                 "volume_high": (54, 55), "volume_long": (55, 56), "volume_low_long": (56, 57),
                 "volume_mid_long": (57, 58), "volume_high_long": (58, 59)}

    def __init__(self, ws: torch.Tensor, n_patches: int) -> None:
        """ws: the bank of mapped latents [num, W, L] on the GPU (latent_augmenter.py:25-29 builds it from a checkpoint)."""
        if not ws.is_cuda:
            raise ops.SsarError("LatentAugmenter needs the latent bank on a CUDA device (no CPU fallback)")
        self.ws = ws.float().contiguous()
        self.n_patches = n_patches
        self.num, self.nw = ws.shape[0], ws.shape[1]
        self.feat_idxs = dict(self.FEAT_IDXS)  # the reference literal repeats keys; a dict keeps one of each (:31-56)
        self.feat_keys = list(self.feat_idxs.keys())
        self.single_dim = -12

    def __call__(self, features):
        residuals, offsets = zip(*(self.random_patch(f) for f in features))
        return torch.stack(residuals), torch.stack(offsets)

    def _rows(self, idx):
        return self.ws[torch.as_tensor(idx, device=self.ws.device, dtype=torch.long)]

    @torch.no_grad()
    def random_patch(self, feature: torch.Tensor):
        """feature [T, 59] (CUDA) -> (residual [T, W, L], offset [1, 1, L]) (latent_augmenter.py:68-105)."""
        feature = feature.float()
        T, L = feature.shape[0], self.ws.shape[2]
        n_knots = np.random.randint(3, 12)  # the size argument is evaluated before the call that draws the rows (:70-71)
        latent = spline_loop_latents(self._rows(np.random.randint(0, self.num, n_knots)), T)
        ones = torch.ones(T, 1, dtype=torch.float32, device=feature.device)

        def merge(S, lats, a, b, m):
            # latent[:, a:b] = (1 - m) * latent[:, a:b] + m * (S @ lats)[:, a:b]   (m None: overwrite)
            ops.mix_gemm(S, lats, out=latent, group_cols=self.nw * L, col_begin=a * L, col_end=b * L,
                         row_alpha=None if m is None else (1 - m).contiguous(), row_beta=None if m is None else m.contiguous())

        for _ in range(self.n_patches):
            start, stop = self.feat_idxs[random.choice(self.feat_keys)]
            if np.random.rand() > 0.5:
                lay_start = np.random.randint(0, self.nw - 6)
                lay_stop = np.random.randint(lay_start, self.nw)
            else:
                lay_start, lay_stop = 0, self.nw
            if stop - start == 1:
                lat = self._rows(np.random.randint(0, self.num, 1))
                m = ops.normalize(feature[:, start:stop].contiguous()).reshape(T)
                merge(ones, lat, lay_start, lay_stop, m)
            else:
                lats = self._rows(np.random.randint(0, self.num, stop - start))
                S = ops.normalize(feature[:, start:stop].contiguous())
                S = S / (S.sum(1, keepdim=True) + 1e-8)
                if np.random.rand() > 0.666:
                    a, b = self.feat_idxs[random.choice(self.feat_keys[self.single_dim:])]
                    merge(S, lats, lay_start, lay_stop, ops.normalize(feature[:, a:b].contiguous()).reshape(T))
                else:
                    merge(S, lats, lay_start, lay_stop, None)
        # offset = mean over frames and layers (:102); the subtraction is one more in-place mixing launch
        st = ops.feature_stats(latent.view(-1, L))
        offset = (st[1] / st[0]).float().reshape(1, 1, L)
        ops.mix_gemm(ones, offset.expand(1, self.nw, L).contiguous(), out=latent, group_cols=self.nw * L, row_alpha=ones.reshape(T),
                     row_beta=-ones.reshape(T))
        return latent, offset
