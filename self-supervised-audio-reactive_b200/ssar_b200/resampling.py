"""`resample(waveform, orig_freq, new_freq)` — the torchaudio call the reference makes at
ssar/supervised/data.py:63 / ssar/random/sample.py:34 (sinc_interp_hann) and rosa/constantq.py:83 (kaiser_window)."""
from __future__ import annotations

from . import ops, plan


def resample(waveform, orig_freq, new_freq, resampling_method="sinc_interp_hann", post_scale=1.0, clip_n_in=None):
    """clip_n_in (int64 [B], device): ragged batch — each row holds that many samples followed by zeros."""
    if orig_freq == new_freq:
        return waveform
    squeeze = waveform.dim() == 1
    rp = plan.resample_plan(int(orig_freq), int(new_freq), resampling_method, str(waveform.device))
    y = ops.resample(waveform, rp, post_scale, clip_n_in=clip_n_in)
    return y[0] if squeeze else y
