"""Oracle: the video-side feature twins of ssar/features/video.py that need only the tensor (:11-31 colour histograms,
:58-59 visual_variance, :62-73 absdiff), torch CPU.  TEST INFRASTRUCTURE ONLY.  Pinned by tests/test_oracle_post.py against
fixtures made by running the reference's own function text (tests/golden/make_golden.py::g_video)."""
from __future__ import annotations

import torch


def channel_hist(video, channel, bins=32):
    """video.py:11-25: per frame torch.histc over the frame's own range, divided by the largest count."""
    hist = torch.stack([torch.histc(frame, bins=bins) for frame in video[:, channel]])
    return hist / hist.max(dim=1, keepdim=True).values


def rgb_hist(video, bins=96):
    """video.py:29-30."""
    return torch.cat([channel_hist(video, c, bins // 3) for c in range(3)], -1)


def visual_variance(video):
    """video.py:58-59."""
    return (video.std((1, 2, 3)) ** 2).unsqueeze(-1)


def absdiff(video):
    """video.py:62-73: sum |x[t + 1] - x[t]| per frame pair, the last value repeated."""
    d = torch.diff(video, dim=0).abs().flatten(1).sum(1)
    return torch.cat([d, d[-1:]]).unsqueeze(-1)
