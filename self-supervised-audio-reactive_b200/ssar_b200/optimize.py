"""Mirror of the per-song fixed decoder of ssar/optimize.py:103-142 (grouped soft-max envelopes -> latents, mu / sigma
envelopes -> noise maps).  The optimisation loop around it (HiPPO envelopes, losses) is out of scope."""
from __future__ import annotations

import torch

from . import ops


class FixedLatentNoiseDecoder:
    """optimize.py:103-142: x [T, S*G*H + 2*n_noise] -> (latents [T, W, L], noise list of [T, s, s], s = 4, 8, ...)."""

    def __init__(self, latents, n_latent_split=3, n_latent_groups=3, n_latent_per_group=3):
        self.S, self.G, self.H = n_latent_split, n_latent_groups, n_latent_per_group
        assert (
            len(latents) == self.S * self.G * self.H
        ), f"Number of latent vectors supplied does not equal n_latent_split * n_latent_per_split ({self.S * self.G * self.H})"
        self.latents = latents
        self.W = self.latents.shape[1] // self.S
        self.noise_sigma = 2

    def _weights(self, x):
        """[T, S, S*G*H]: per split the group-wise soft-max, renormalised to a total of 1 (optimize.py:121-124),
        zeros off the split's own block — a [T, 27] computation, done with torch glue ops."""
        T, GH = x.shape[0], self.G * self.H
        w = torch.zeros(T, self.S, self.S * GH, dtype=torch.float32, device=x.device)
        for i in range(self.S):
            env = x[:, i * GH : (i + 1) * GH].reshape(T, self.G, self.H).softmax(dim=2)
            env = env / (env.sum(dim=(1, 2), keepdim=True) + 1e-8)
            w[:, i, i * GH : (i + 1) * GH] = env.reshape(T, GH)
        return w

    def forward(self, x, noise_draws=None):
        """noise_draws: optional list of standard-normal tensors [T, s, s]; torch.randn on x's device when None."""
        x = x.float()
        T = x.shape[0]
        L = self.latents.shape[2]
        lat = ops.mix_gemm(self._weights(x), self.latents.to(x.device).float(), group_cols=self.W * L)
        lat = lat.reshape(T, self.latents.shape[1], L)
        noise_envs = x[:, self.S * self.G * self.H :]
        noise = []
        for i in range(noise_envs.shape[-1] // 2):
            mu, sig = noise_envs[:, 2 * i].contiguous(), noise_envs[:, 2 * i + 1].contiguous()
            size = 2 ** (i + 2)
            z = noise_draws[i] if noise_draws is not None else torch.randn(T, size, size, device=x.device)
            n = ops.gauss1d(z.contiguous(), self.noise_sigma, "circular")
            noise.append(ops.rows_affine(n.reshape(T, size * size), mu, sig).reshape(T, size, size))
        return lat, noise

    __call__ = forward
