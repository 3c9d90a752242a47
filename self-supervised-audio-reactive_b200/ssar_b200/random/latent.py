"""Mirror of ssar/random/latent.py — latent-sequence construction as weight matrices + one mixing GEMM.

Every sequence here is linear in the palette, so smoothing and merging act on small [T, K] weight matrices and the
[T, 18, 512] tensor is touched once per call (`latent_patch`, in place like the reference) or once in total
(`LatentWeights`, used by `Patch.forward`)."""
from __future__ import annotations

import torch

from .. import _cabi as cabi
from .. import ops, plan

LAYER_SLICES = {"low": (0, 6), "mid": (6, 12), "high": (12, 18), "lowmid": (0, 12), "midhigh": (6, 18), "all": (0, 18)}


def spline_loop_latents(y, size, n_loops=1):
    """ssar/random/latent.py:7-13: y [K, W, L] -> [size, W, L]."""
    W = plan.spline_loop_weights_cpu(len(y), int(size), n_loops).to(y.device)
    return ops.mix_gemm(W, y)


def sequence_weights(rng, n_frames, n_palette, segmentations, features, tempo, fps, patch_type, segments, loop_bars,
                     seq_feat, seq_feat_weight, device):
    """The sub-patch's sequence (latent.py:35-54) as (selection [K] long on the rng device, weights [T, K] on `device`)
    with the Gaussian smoothing of :41 and :54 already applied to the weights (it commutes with the palette sum)."""
    feature = seq_feat_weight * features[seq_feat]
    segmentation = segmentations[(seq_feat, segments)]
    permutation = torch.randperm(n_palette, generator=rng, device=rng.device)
    if patch_type == "segmentation":
        selection = permutation[:segments]
        onehot = torch.zeros(n_frames, int(segments), dtype=torch.float32, device=device)
        onehot[torch.arange(n_frames, device=device), segmentation.to(device).long()] = 1.0
        S = ops.gauss1d(onehot, 5, "circular")
    elif patch_type == "feature":
        n_select = feature.shape[1]
        if n_select == 1:
            selection = permutation[:2]
            f = feature.to(device).float().reshape(n_frames, 1)
            S = torch.cat([f, 1 - f], dim=1)
        else:
            selection = permutation[:n_select]
            S = feature.to(device).float()
    elif patch_type == "loop":
        selection = permutation[:segments]
        n_loops = n_frames / fps / 60 / tempo / 4 / loop_bars
        S = plan.spline_loop_weights_cpu(len(selection), n_frames, n_loops).to(device)
    else:
        raise ValueError(patch_type)
    return selection, ops.gauss1d(S.contiguous(), 1, "circular")


def merge_coefficients(features, merge_type, mod_feat, mod_feat_weight, n_frames, device):
    """latent.py:68-78 as (alpha, beta): latents = alpha * latents + beta * sequence (floats or [T] tensors)."""
    if merge_type == "average":
        return 0.5, 0.5
    if merge_type == "modulate":
        m = (mod_feat_weight * features[mod_feat]).to(device).float().reshape(n_frames).contiguous()
        return (1 - m).contiguous(), m
    return 0.0, 1.0  # overwrite


def latent_patch(rng, latents, palette, segmentations, features, tempo, fps, patch_type, segments, loop_bars, seq_feat,
                 seq_feat_weight, mod_feat, mod_feat_weight, merge_type, merge_depth):
    """ssar/random/latent.py:16-80 — modifies `latents` [T, W, L] IN PLACE (like the reference) and returns it."""
    T, Wn, L = latents.shape
    dev = latents.device
    selection, S = sequence_weights(rng, T, len(palette), segmentations, features, tempo, fps, patch_type, segments,
                                    loop_bars, seq_feat, seq_feat_weight, dev)
    alpha, beta = merge_coefficients(features, merge_type, mod_feat, mod_feat_weight, T, dev)
    lo, hi = LAYER_SLICES[merge_depth]
    a_t = alpha if torch.is_tensor(alpha) else torch.full((T,), float(alpha), device=dev)
    b_t = beta if torch.is_tensor(beta) else torch.full((T,), float(beta), device=dev)
    ops.mix_gemm(S, palette[selection.to(palette.device)].to(dev), out=latents, col_begin=lo * L, col_end=min(hi, Wn) * L,
                 row_alpha=a_t, row_beta=b_t)
    return latents


class LatentWeights:
    """Accumulates a whole stack of sub-patches as weights [T, G, P] over the full palette (G layer groups of
    `group_layers` layers) and renders [T, W, L] with ONE mixing GEMM.  Each sub-patch is ONE kernel launch
    (`ssar_subpatch_merge`: feature scaling, the sigma = 1 smoothing, the merge coefficients and the scatter of the
    selected palette columns are fused), so a 20-sub-patch stack costs ~20 launches and no [T, 18, 512] traffic."""

    def __init__(self, n_frames, n_palette, n_layers=18, group_layers=6, device="cuda"):
        self.T, self.P, self.Wn, self.gl = n_frames, n_palette, n_layers, group_layers
        self.G = n_layers // group_layers
        self.device = torch.device(device)
        self.w = torch.empty(self.T, self.G, self.P, dtype=torch.float32, device=self.device)
        self._labels = {}
        # deferred mode (Patch.forward): sub-patches are queued as descriptors and applied by ONE ssar_latent_stack launch
        self.deferred = False
        self._queue, self._keep, self._fresh = [], [], True

    def _sel(self, selection):
        return selection.to(self.device, non_blocking=True)

    def _enqueue(self, selection, src, g_begin, g_end, src_mode=0, src_scale=1.0, sigma=0, mod=None, mod_scale=1.0,
                 alpha=0.5, beta=0.5, pre_sigma=None):
        d = cabi.SubpatchDesc()
        selection = self._sel(selection)
        d.selection, d.src, d.n_select = selection.data_ptr(), src.data_ptr(), int(selection.numel())
        d.src_ld = src.stride(0) if src_mode != 2 else 0
        d.src_scale, d.src_mode = float(src_scale), int(src_mode)
        d.g_begin, d.g_end = int(g_begin), int(g_end)
        d.radius = min(int(sigma * 4), 3 * self.T) if sigma else 0
        if mod is not None:
            d.mod, d.mod_ld, d.mod_scale = mod.data_ptr(), mod.stride(0), float(mod_scale)
        d.alpha, d.beta = float(alpha), float(beta)
        if src_mode == 2:
            d.pre_radius = min(int(pre_sigma * 4), 3 * self.T)
            pre = plan.gaussian_taps(float(pre_sigma), d.pre_radius, str(self.device))
            d.pre_taps = pre.data_ptr()
            self._keep.append(pre)
        self._keep += [selection, src, mod]
        self._queue.append(d)

    def flush(self):
        """Applies the queued sub-patches (deferred mode) with one launch per 24 of them."""
        if self._queue:
            taps = plan.gaussian_taps(1.0, min(4, 3 * self.T), str(self.device))
            ops.latent_stack(self.w, self._queue, taps, load_initial=not self._fresh)
            self._queue, self._keep, self._fresh = [], [], False

    def set_base(self, selection, S):
        """weights = S scattered to the selected palette columns in every layer group (overwrites)."""
        if self.deferred and len(selection) <= 32:
            self._enqueue(selection, S, 0, self.G, sigma=0, alpha=0.0, beta=1.0)
            return
        self._fresh = False
        ops.subpatch_merge(self.w, 0, self.G, self._sel(selection), S, sigma=0, alpha=0.0, beta=1.0)

    def merge(self, selection, S, alpha, beta, merge_depth):
        """Generic merge of an already smoothed sequence S [T, K] with float or [T] tensor coefficients."""
        lo, hi = LAYER_SLICES[merge_depth]
        self.flush()
        if torch.is_tensor(alpha) or torch.is_tensor(beta):
            D = torch.zeros(self.T, self.P, dtype=torch.float32, device=self.device)
            D.index_add_(1, self._sel(selection).long(), S)
            ops.weights_merge(self.w, D, lo // self.gl, hi // self.gl, alpha, beta)
        else:
            ops.subpatch_merge(self.w, lo // self.gl, hi // self.gl, self._sel(selection), S, sigma=0, alpha=float(alpha),
                               beta=float(beta))

    def labels(self, key, segmentation):
        """Segmentation labels as int64 on the compute device, copied once per key."""
        t = self._labels.get(key)
        if t is None:
            t = self._labels[key] = segmentation.to(self.device).long().contiguous()
        return t

    def apply_subpatch(self, rng, segmentations, features, tempo, fps, patch_type, segments, loop_bars, seq_feat,
                       seq_feat_weight, mod_feat, mod_feat_weight, merge_type, merge_depth, permutation=None):
        """latent.py:35-78 on the weights: same RNG draw (one randperm per sub-patch), same arithmetic per element.
        permutation: this sub-patch's torch.randperm(P) when the caller drew it already (batched Philox launch)."""
        T, P = self.T, self.P
        segmentation = segmentations[(seq_feat, segments)]  # looked up for every patch type, like latent.py:36
        if permutation is None:
            permutation = torch.randperm(P, generator=rng, device=rng.device)
        lerp, scale = False, 1.0
        use_stack = self.deferred and min(int(1 * 4), 3 * T) <= 8
        src_mode = None
        if patch_type == "segmentation":
            selection = permutation[:segments]
            if use_stack and int(segments) <= 32:
                src, src_mode = self.labels((seq_feat, segments), segmentation), 2
            else:
                src = ops.onehot_gauss(self.labels((seq_feat, segments), segmentation), int(segments), 5)
        elif patch_type == "feature":
            feature = features[seq_feat]
            scale = float(seq_feat_weight)
            if feature.shape[1] == 1:
                selection, src, lerp = permutation[:2], feature, True
            else:
                selection, src = permutation[: feature.shape[1]], feature
        elif patch_type == "loop":
            selection = permutation[:segments]
            n_loops = T / fps / 60 / tempo / 4 / loop_bars
            src = plan.spline_loop_weights(len(selection), T, n_loops, str(self.device))
        else:
            raise ValueError(patch_type)
        lo, hi = LAYER_SLICES[merge_depth]
        if merge_type == "modulate":
            mod, mod_scale, alpha, beta = features[mod_feat], float(mod_feat_weight), 0.0, 0.0
        else:
            mod, mod_scale = None, 1.0
            alpha, beta = (0.5, 0.5) if merge_type == "average" else (0.0, 1.0)
        if use_stack and len(selection) <= 32:
            self._enqueue(selection, src, lo // self.gl, hi // self.gl, src_mode=src_mode if src_mode is not None else
                          (1 if lerp else 0), src_scale=scale, sigma=1, mod=mod, mod_scale=mod_scale, alpha=alpha, beta=beta,
                          pre_sigma=5)
            return
        self.flush()
        ops.subpatch_merge(self.w, lo // self.gl, hi // self.gl, self._sel(selection), src, src_scale=scale, lerp=lerp,
                           sigma=1, mod=mod, mod_scale=mod_scale, alpha=alpha, beta=beta)

    def render(self, palette, out=None, palette_split=None):
        self.flush()
        return ops.mix_gemm(self.w, palette.to(self.device), out=out, basis_split=palette_split)
