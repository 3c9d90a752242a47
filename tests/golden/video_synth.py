"""Deterministic synthetic video shared by the golden generator and the tests."""
import torch


def synth_video(T=70, C=3, H=48, W=40, seed=5):
    """Deterministic [T, C, H, W] video in [-1, 1]: drifting gradients + noise + a few constant frames / channels."""
    g = torch.Generator("cpu").manual_seed(8000 + seed)
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, H), torch.linspace(-1, 1, W), indexing="ij")
    t = torch.linspace(0, 6.28, T)[:, None, None, None]
    ph = torch.tensor([0.0, 2.1, 4.2])[None, :C, None, None]
    v = 0.6 * torch.sin(3 * xx[None, None] + t + ph) * torch.cos(2 * yy[None, None] - 0.5 * t) + 0.2 * torch.randn(T, C, H, W, generator=g)
    v = v.clamp(-1, 1)
    if T > 30:
        v[7, 1] = 0.25   # a constant channel: histc's degenerate range
        v[30] = v[29]    # a repeated frame: zero absdiff
    return v.contiguous()
