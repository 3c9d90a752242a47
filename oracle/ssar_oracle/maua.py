"""Oracle: the numpy/torch "audioreactive" trajectory helpers of the vendored maua package (torch CPU, fp32).
TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's CPU leg, never by the product.

Restates maua/maua/audiovisual/audioreactive/latent.py:13-103 (single_weighted, multi_weighted, select_modulo, slerp,
slerp_loops, spline_loops, tempo_loops) and signal.py:26-38, 108-157 (normalize, gaussian_filter with `causal`).
Pinned: tests/golden/maua.npz holds outputs of the reference's own two files imported in the build container
(tests/golden/make_golden.py:g_maua); spline_loops goes through torchcubicspline (un-vendored -> unpinned there).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .traj import natural_cubic_spline, spline_evaluate


def normalize(x):
    """signal.py:26-38: (x - min) / max(x - min) — no epsilon, unlike ssar's processing.normalize."""
    y = x - x.min()
    return y / y.max()


def gaussian_filter(x, sigma, causal=None, mode="circular"):
    """signal.py:108-157: radius min(int(4 sigma), 3 T); taps right of the centre scaled by `causal` (0 when `causal`
    is not a float, e.g. the integer 0 `select_modulo` passes) BEFORE the normalisation; circular pad; depthwise conv."""
    dim = x.dim()
    n_frames = x.shape[0]
    while x.dim() < 3:
        x = x[:, None]
    radius = min(int(sigma * 4), 3 * len(x))
    channels = x.shape[1]
    k = torch.arange(-radius, radius + 1, dtype=torch.float32)
    k = torch.exp(-0.5 / sigma ** 2 * k ** 2)
    if causal is not None:
        k[radius + 1:] *= 0 if not isinstance(causal, float) else causal
    k = k / k.sum()
    k = k.view(1, 1, -1).repeat(channels, 1, 1)
    if dim == 4:
        t, c, h, w = x.shape
        x = x.reshape(t, c, h * w)
    x = x.transpose(0, 2)
    if radius > n_frames:
        x = F.pad(x, (n_frames, n_frames), mode=mode)
        x = F.pad(x, (radius - n_frames, radius - n_frames), mode="replicate")
    else:
        x = F.pad(x, (radius, radius), mode=mode)
    x = F.conv1d(x, weight=k, groups=channels)
    x = x.transpose(0, 2)
    if dim == 4:
        x = x.reshape(t, c, h, w)
    if x.dim() > dim:
        x = x.squeeze()
    return x


def single_weighted(low_latent, high_latent, envelope):
    """latent.py:13-19."""
    e = envelope[:, None, None]
    return low_latent[None] * (1 - e) + high_latent[None] * e


def multi_weighted(latents, envelopes):
    """latent.py:23-32 (normalises `envelopes` IN PLACE like the reference)."""
    envelopes /= envelopes.sum(dim=1, keepdim=True)
    sel = latents[torch.arange(envelopes.shape[1]) % len(latents)]
    return torch.einsum("ta,awl->twl", envelopes, sel)


def select_modulo(latents, envelope, smooth=2):
    """latent.py:36-45."""
    low, high = torch.quantile(envelope, 0.25), torch.quantile(envelope, 0.75)
    idx = normalize(envelope.clamp(low, high))
    idx = (idx * (len(latents) - 1)).round().long()
    return gaussian_filter(latents[idx], smooth, causal=0)


def slerp(a, b, t):
    """latent.py:55-65: a, b [S, W, L], t [n] -> [n, S, W, L]."""
    a = a / a.norm(dim=-1, keepdim=True)
    b = b / b.norm(dim=-1, keepdim=True)
    d = (a * b).sum(dim=-1, keepdim=True)
    p = t * torch.acos(d)
    p = p.permute(2, 0, 1)[..., None]
    c = b - d * a
    c = c / c.norm(dim=-1, keepdim=True)
    out = a[None] * torch.cos(p) + c[None] * torch.sin(p)
    return out / out.norm(dim=-1, keepdim=True)


def slerp_loops(y, size, n_loops):
    """latent.py:69-81 — note the (t, segment) row order left by the reshape, reproduced as written."""
    y = torch.cat([y] * n_loops + [y[[0]]])
    t = torch.linspace(0, 1, round(size / len(y)))
    out = slerp(y[:-1], y[1:], t)
    out = out.reshape(-1, *out.shape[2:])
    out = F.interpolate(out.permute(1, 2, 0), size=size, mode="linear", align_corners=False)
    return out.permute(2, 0, 1)


def spline_loops(y, size, n_loops):
    """latent.py:84-92."""
    y = torch.cat([y] * n_loops + [y[[0]]])
    t_in = torch.linspace(0, 1, len(y))
    t_out = torch.linspace(0, 1, size)
    return spline_evaluate(t_in, natural_cubic_spline(t_in, y), t_out)


def tempo_loops(latents, n_frames, fps, tempo, type="spline"):
    """latent.py:96-103."""
    n_loops = round(n_frames / fps * (tempo / 4 / 60))
    return spline_loops(latents, n_frames, n_loops) if type == "spline" else slerp_loops(latents, n_frames, n_loops)


def make_noise_pyramid(noise, sizes):
    """maua/GAN/wrappers/stylegan2.py:193-210: noise [T, 1, H, W] -> one [T, 1, h, w] tensor per (h, w) in `sizes`
    (the noise_const shapes of synthesis layers 1..9), bicubic (align_corners=False), each FRAME divided by its own
    unbiased standard deviation."""
    out = []
    for h, w in sizes:
        n = F.interpolate(noise, (h, w), mode="bicubic", align_corners=False)
        out.append(n / n.std((1, 2, 3), keepdim=True))
    return out
