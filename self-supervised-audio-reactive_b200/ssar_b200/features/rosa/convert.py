"""Mirror of ssar/features/rosa/convert.py (unit conversions + power_to_db)."""
from __future__ import annotations

import math

import numpy as np
import torch

from ... import ops, plan


def power_to_db(magnitude, ref_value=1.0, amin=1e-10, top_db=80.0):
    """rosa/convert.py:7-12.  The floor `max - top_db` uses the maximum of the WHOLE tensor (one clip)."""
    if float(ref_value) != 1.0 or float(amin) != 1e-10:
        raise NotImplementedError("power_to_db: only ref_value=1, amin=1e-10 (the reference's only use) is built")
    return ops.power_to_db(magnitude, n_clips=1, top_db=top_db).view(magnitude.shape)


def hz_to_mel(frequencies, htk=False, device="cpu"):
    assert not htk
    return plan.hz_to_mel_f32(torch.as_tensor(frequencies, dtype=torch.float32)).to(device)


def mel_to_hz(mels, htk=False):
    assert not htk
    return plan.mel_to_hz_f32(mels)


def note_to_hz(note):
    """Only "C1" is ever requested on the path (rosa/convert.py:129-130 via librosa): 440*2^((24-69)/12)."""
    if note != "C1":
        raise NotImplementedError(note)
    return torch.tensor(440.0 * 2.0 ** ((24 - 69) / 12.0)).float()


def hz_to_midi(frequencies):
    return 12 * (np.log2(frequencies) - np.log2(440.0)) + 69


def hz_to_octs(frequencies, tuning=0.0, bins_per_octave=12):
    A440 = 440.0 * 2.0 ** (tuning / bins_per_octave)
    return torch.log2(frequencies / (float(A440) / 16))
