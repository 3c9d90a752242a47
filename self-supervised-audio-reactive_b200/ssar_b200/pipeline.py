"""Batched entry points for the BASELINE.json configurations (one call = all clips of a batch).

These compose the same kernels the per-function mirrors use, but with every clip of the batch in one launch
and with everything that can share an FFT fused into that FFT's epilogue.
"""
from __future__ import annotations

import torch

from . import ops, plan


def spectral_frontend(x: torch.Tensor, sr, n_fft: int = 2048, hop: int = 512, want_spec: bool = True):
    """BASELINE config 2: STFT power spectrogram + mel + onset strength + RMS for a batch of clips.

    x: [B, N] fp32 on the GPU.  Returns frame-major tensors:
      spec  [B, T, 1+n_fft/2]  == spectrogram(y, n_fft, hop, power=2)      (rosa/spectral.py:59-62)
      mel   [B, T, 128]        == melspectrogram(y, sr, n_fft, hop)        (rosa/spectral.py:65-70)
      onset [B, T]             == onset_strength(y, sr, hop, n_fft)        (rosa/beat.py:10-23)
      rms   [B, T]             == rms(y, sr, n_fft, hop)[:, 0]             (features/audio.py:31-37)
    with T = N // hop.  Two launches: the fused frame kernel, then the tiny dB/flux kernel (it needs the
    clip-global mel maximum, which the first kernel accumulates in its epilogue).
    """
    dev = str(x.device)
    T = x.shape[-1] // hop
    bank = plan.mel_bank(float(sr), n_fft, 128, 0.0, None, dev)
    bank_onset = plan.mel_bank(float(sr), n_fft, 128, 0.0, 11025.0, dev)
    out = ops.frames_fft(
        x, n_fft, hop, T, torch.hann_window,
        spec_power=2 if want_spec else None, want_rms=True,
        banks=[(bank, True, False), (bank_onset, True, True)],
    )
    onset = ops.onset_flux(out["bank1"], out["bank1_max"], 1 + n_fft // (2 * hop), "mean")
    return {"spec": out.get("spec"), "mel": out["bank0"], "onset": onset, "rms": out["rms"]}
