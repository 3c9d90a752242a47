"""Mirror of the latent / noise builders of ssar/features/generate.py:155-230 (the synthetic audio-visual pairs with a
known chroma / onset correlation).  The reference functions also pick a random audio file and call the StyleGAN2
mapper / synthesizer (out of scope): here the audio features and the mapped latents `ws` are arguments, and each
function returns the (latents [T, W, L], noise [T, 1, s, s]) pair the reference hands to `get_video`.
As in `Patch.forward`, the time smoothing of a latent sequence is applied to its [T, K] weights (it commutes with the
palette sum) and the sequence is written once by the mixing GEMM."""
from __future__ import annotations

import torch

from .. import ops
from .processing import clamp_lower_percentile, clamp_peaks_percentile, gaussian_filter, normalize

FPS = 24  # generate.py:51


def audio2features(audio, sr, clamp=True, smooth=True):
    """generate.py:106-136: the generators' variant of audio2features.  audio [N] (or [B, N]) ALREADY at sr = 1024 * fps
    -> [T, 59] (or [B, T, 59]): the multi-column features and the flatness pre-smoothed with sigma 10, pulse = plp(audio)
    (not of the percussive part), clamp to each column's [q10, q97.5] (torch.quantile, linear), then sigma 2.  Same
    kernels and the same shared passes as supervised/data.py's audio2features (pipeline.feature_chain)."""
    from .. import pipeline

    squeeze = audio.dim() == 1
    x = audio[None] if squeeze else audio
    raw = pipeline.feature_chain(x.float(), sr, want_a2f=True, want_mir=False, pulse_of="audio")["a2f"]
    feats = torch.cat([ops.gauss1d(raw[:, :, :46].contiguous(), 10, "circular", batched=True), raw[:, :, 46:]], dim=2)
    feats = feats.contiguous()
    if clamp:
        feats = ops.clamp_cols(feats, ops.col_quantile(feats, 0.1), ops.col_quantile(feats, 0.975))
    if smooth:
        feats = ops.gauss1d(feats, 2, "circular", batched=True)
    return feats[0] if squeeze else feats


def _chroma_weights(chroma, fps):
    """generate.py:159-164: smooth, clamp peaks (97) / lower (5), L1-normalise per frame."""
    c = gaussian_filter(chroma, fps / 24)
    c = clamp_peaks_percentile(c, 97)
    c = clamp_lower_percentile(c, 5)
    return c / c.sum(1, keepdim=True)


def _onset_weights(ons, fps):
    """generate.py:182-187: smooth, clamp, min-max normalise -> [T]."""
    o = gaussian_filter(ons.reshape(-1), fps / 24)
    o = clamp_peaks_percentile(o, 97)
    o = clamp_lower_percentile(o, 5)
    return normalize(o).reshape(-1)


def _unit_std(noise, per_frame):
    # torch glue: one reduction over the noise tensor (generate.py:172, 196)
    return noise / (noise.std(dim=(1, 2, 3), keepdim=True) if per_frame else noise.std())


def high_chroma_correlation(chroma, ws, fps=FPS, noise_draw=None):
    """generate.py:155-174.  chroma: chromagram(audio, sr) [T, 12]; ws [12, W, L]; noise_draw: randn [T, 1, 32, 32]."""
    w = _chroma_weights(chroma, fps)
    latents = ops.mix_gemm(gaussian_filter(w, fps / 24).contiguous(), ws.float())
    z = noise_draw if noise_draw is not None else torch.randn((len(w), 1, 32, 32), device=chroma.device)
    return latents, _unit_std(gaussian_filter(z, fps), per_frame=False)


def high_onset_correlation(ons, ws, fps=FPS, noise_draw=None):
    """generate.py:177-199.  ons: onsets(audio, sr) [T, 1]; ws [2, W, L]; noise_draw: randn [T, 1, 64, 64]."""
    o = _onset_weights(ons, fps)
    w = torch.stack([o, 1 - o], dim=1)
    latents = ops.mix_gemm(gaussian_filter(w, fps / 24).contiguous(), ws.float())
    z = noise_draw if noise_draw is not None else torch.randn((len(o), 1, 64, 64), device=ons.device)
    noise = _unit_std(gaussian_filter(z, fps / 4), per_frame=True)
    return latents, noise + noise * o.reshape(-1, 1, 1, 1)


def high_both_correlation(chroma, ons, ws, fps=FPS, noise_draw=None):
    """generate.py:202-230.  ws [14, W, L]: 12 chroma latents, then the onset pair."""
    c, o = _chroma_weights(chroma, fps), _onset_weights(ons, fps)
    # (chroma_latents + onset_latents) / 2 as weights: the halving is exact in fp32
    w = torch.cat([c, o[:, None], (1 - o)[:, None]], dim=1) / 2
    latents = ops.mix_gemm(gaussian_filter(w, fps / 24).contiguous(), ws.float())
    z = noise_draw if noise_draw is not None else torch.randn((len(o), 1, 64, 64), device=ons.device)
    noise = _unit_std(gaussian_filter(z, fps / 4), per_frame=True)
    return latents, noise + noise * o.reshape(-1, 1, 1, 1)
