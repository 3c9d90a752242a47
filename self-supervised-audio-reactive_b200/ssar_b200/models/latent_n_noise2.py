"""Mirror of the fixed decoder of ssar/models/latent_n_noise2.py:134-178 (and its per-song twin, ssar/optimize.py:103-142).
The learned networks of that file (EnvelopeReactor, LearnedLatentNoiseDecoder) are out of scope."""
from __future__ import annotations

import torch

from .. import ops


class FixedLatentNoiseDecoder:
    """models/latent_n_noise2.py:134-178: envelopes [B, T, S*H + 2*n_noise] -> (latents [B, T, W, L], noise list)."""

    def __init__(self, latents, hidden_size=12, n_latent_split=3, n_noise=4):
        self.S = n_latent_split
        self.H = hidden_size
        assert (
            len(latents) == self.S * self.H
        ), f"Number of latent vectors supplied does not equal n_latent_split * hidden_size ({self.S * self.H})"
        self.latents = latents
        self.W = self.latents.shape[1] // self.S
        self.noise_sigma = 5

    def _weights(self, x):
        """Per-split normalised envelopes laid out as [rows, S, S*H] with zeros off the split's own block."""
        rows = x.shape[0]
        w = torch.zeros(rows, self.S, self.S * self.H, dtype=torch.float32, device=x.device)
        for i in range(self.S):
            env = x[:, i * self.H : (i + 1) * self.H]
            w[:, i, i * self.H : (i + 1) * self.H] = env / env.sum(dim=-1, keepdim=True)
        return w

    def forward(self, x, noise_draws=None):
        """`noise_draws`: optional list of pre-drawn standard-normal tensors [B, T, s, s] (s = 4, 8, ...); drawn with
        torch.randn on x's device when None (models/latent_n_noise2.py:172)."""
        B, T, _ = x.shape
        L = self.latents.shape[2]
        xf = x.reshape(B * T, -1).float()
        lat = ops.mix_gemm(self._weights(xf), self.latents.to(x.device).float(), group_cols=self.W * L)
        lat = lat.reshape(B, T, self.latents.shape[1], L)
        noise_envs = xf[:, self.S * self.H :]
        noise = []
        for i in range(noise_envs.shape[-1] // 2):
            mu = noise_envs[:, 2 * i].contiguous()
            sig = noise_envs[:, 2 * i + 1].contiguous()
            size = 2 ** (i + 2)
            z = noise_draws[i] if noise_draws is not None else torch.randn(B, T, size, size, device=x.device)
            zt = z.transpose(1, 0).contiguous()                      # [T, B, s, s]: smoothing runs along time
            n = ops.gauss1d(zt, self.noise_sigma, "circular").transpose(1, 0).contiguous()
            noise.append(ops.rows_affine(n.reshape(B * T, size * size), mu, sig).reshape(B, T, size, size))
        return lat, noise

    __call__ = forward
