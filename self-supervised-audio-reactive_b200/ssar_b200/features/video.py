"""Mirror of the video-side feature twins of ssar/features/video.py that need only the tensor itself:
`redogram / greenogram / blueogram / rgb_hist` (:11-31), `visual_variance` (:58-59), `absdiff` (:62-73) — video [T, C, H, W]
fp32 in, [T, C'] out like the audio features.  `spectral_flux`, `onset_envelope` and `standardize` (the native quantile)
are in `processing.py`.  Not built: the HSV histograms (kornia's colour conversion), `video_spectrogram` (cv2.linearPolar)
and optical flow (cv2 Farneback on the CPU)."""
from __future__ import annotations

import torch

from .. import _cabi as cabi
from .._cabi import SsarError, i32, i64, ptr


def _video(video):
    if not video.is_cuda:
        raise SsarError("ssar_b200 needs CUDA tensors (no CPU fallback)")
    if video.dim() != 4:
        raise SsarError(f"expected video [T, C, H, W], got {tuple(video.shape)}")
    return video.contiguous().float()


def _hist(video, channel: int, bins: int, out=None, col0: int = 0):
    v = _video(video)
    T, Cn, H, W = v.shape
    if out is None:
        out = torch.empty(T, bins, dtype=torch.float32, device=v.device)
    cabi.call("ssar_video_hist", ptr(v), i32(T), i32(Cn), i32(H), i32(W), i32(channel), i32(bins), ptr(out), i64(out.shape[1]),
              i32(col0), nbytes=2 * T * H * W * 4)
    return out


def redogram(video, bins: int = 32):
    """video.py:11-13."""
    return _hist(video, 0, bins)


def greenogram(video, bins: int = 32):
    """video.py:17-19."""
    return _hist(video, 1, bins)


def blueogram(video, bins: int = 32):
    """video.py:23-25."""
    return _hist(video, 2, bins)


def rgb_hist(video, bins: int = 96):
    """video.py:29-30 -> [T, 3 * (bins // 3)]: the three histograms written side by side."""
    b = bins // 3
    out = torch.empty(video.shape[0], 3 * b, dtype=torch.float32, device=video.device)
    for c in range(3):
        _hist(video, c, b, out, c * b)
    return out


def visual_variance(video):
    """video.py:58-59 -> [T, 1]."""
    v = _video(video)
    out = torch.empty(v.shape[0], dtype=torch.float32, device=v.device)
    cabi.call("ssar_video_variance", ptr(v), i32(v.shape[0]), i64(v[0].numel()), ptr(out), nbytes=2 * cabi.nbytes_of(v))
    return out.unsqueeze(-1)


def absdiff(video, stride: int = 64):
    """video.py:62-73 -> [T, 1] (the reference's `stride` only bounds its temporaries; the values do not depend on it)."""
    v = _video(video)
    if v.shape[0] < 2:
        raise IndexError("list index out of range")  # the reference's y[-1] on an empty list
    out = torch.empty(v.shape[0], dtype=torch.float32, device=v.device)
    cabi.call("ssar_video_absdiff", ptr(v), i32(v.shape[0]), i64(v[0].numel()), ptr(out), nbytes=2 * cabi.nbytes_of(v))
    return out.unsqueeze(-1)
