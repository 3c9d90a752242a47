"""Mirror of the vendored maua "audioreactive" trajectory helpers the reference's patches are written against
(maua/maua/audiovisual/audioreactive/latent.py:13-103, signal.py:26-38, 108-157) — same names, arguments and shapes,
CUDA tensors in and out.  Every [T, W, L] sequence is produced by one kernel writing the trajectory once: weighted sums,
spline loops and quantised selections go through the mixing GEMM on their [T, K] weights (smoothing along time is
applied to the weights: it commutes with the palette sum), slerp loops through `ssar_slerp_loops`."""
from __future__ import annotations

import torch

from . import ops, plan


def normalize(x):
    """signal.py:26-38 (no epsilon): (x - min) / max(x - min) over the whole tensor."""
    x = x.contiguous().float()
    mn, mx = x.min(), x.max()  # two tiny reductions; the scaling below is the reference's expression
    return (x - mn) / (mx - mn)


def gaussian_filter(x, sigma, causal=None, mode="circular"):
    """signal.py:108-157: Gaussian along time with the right half of the kernel scaled by `causal`."""
    T = x.shape[0]
    if min(int(sigma * 4), 3 * T) > T:
        raise NotImplementedError("gaussian_filter: radius larger than the sequence (replicate-padded case) is not built")
    out = ops.gauss1d(x, sigma, mode, causal=causal)
    return out


def single_weighted(low_latent, high_latent, envelope):
    """latent.py:13-19: low * (1 - e) + high * e -> [T, W, L]."""
    e = envelope.float().reshape(-1, 1)
    return ops.mix_gemm(torch.cat([1 - e, e], dim=1).contiguous(), torch.stack([low_latent, high_latent]).float())


def multi_weighted(latents, envelopes):
    """latent.py:23-32: envelopes [T, A] are normalised IN PLACE (like the reference); latent a % len(latents) per
    envelope column."""
    envelopes /= envelopes.sum(dim=1, keepdim=True)
    A, n = envelopes.shape[1], len(latents)
    if A <= n:
        return ops.mix_gemm(envelopes.float().contiguous(), latents[:A].float())
    W = torch.zeros(envelopes.shape[0], n, dtype=torch.float32, device=envelopes.device)
    W.index_add_(1, torch.arange(A, device=envelopes.device) % n, envelopes.float())
    return ops.mix_gemm(W, latents.float())


def select_modulo(latents, envelope, smooth: float = 2):
    """latent.py:36-45: quartile-clamped envelope -> latent index per frame -> smoothing with a one-sided kernel
    (`causal=0`).  The gather + smoothing is a smoothed one-hot weight matrix times the latents."""
    env = envelope.float().reshape(1, -1, 1).contiguous()
    low = ops.col_quantile(env, 0.25)
    high = ops.col_quantile(env, 0.75)
    idx = normalize(ops.clamp_cols(env, low, high).reshape(-1))
    idx = (idx * (len(latents) - 1)).round().long()
    W = ops.onehot_gauss(idx.contiguous(), len(latents), smooth, causal=0)
    return ops.mix_gemm(W, latents.float())


def slerp_loops(y, size, n_loops):
    """latent.py:69-81."""
    return ops.slerp_loops(y, int(size), int(n_loops))


def spline_loops(y, size, n_loops):
    """latent.py:84-92."""
    W = plan.knot_loop_weights_cpu(len(y), int(size), int(n_loops)).to(y.device)
    return ops.mix_gemm(W, y.float())


def tempo_loops(latents, n_frames, fps, tempo, type="spline"):
    """latent.py:96-103."""
    n_loops = round(n_frames / fps * (tempo / 4 / 60))
    return spline_loops(latents, n_frames, n_loops) if type == "spline" else slerp_loops(latents, n_frames, n_loops)
