"""Diagnostic (GPU box): per-stage rel-L2 of the CUDA path vs the CPU oracle along the audio2features chain."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT / "self-supervised-audio-reactive_b200", ROOT / "oracle"):
    sys.path.insert(0, str(p))
import torch

from ssar_b200 import ops, pipeline, plan
from ssar_b200.resampling import resample
from ssar_b200.synth import synth_clip
from ssar_oracle import dsp


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm())


x44 = synth_clip(44100 * 45 + 100, 44100, 8)
xo = dsp.resample(x44, 44100, 24576)
xg = resample(x44.cuda()[None], 44100, 24576)
print("resample", rel(xg[0], xo))
sr = 24576
ho, po = dsp.harmonic(xo), dsp.percussive(xo)
hg, pg = pipeline.hpss_pair(xo.cuda()[None])
print("harm", rel(hg[0], ho), "perc", rel(pg[0], po))
# onset envelopes on the ORACLE's perc (isolates the onset kernel) and on ours
for name, agg in (("mean", "mean"), ("median", "median")):
    eo = dsp.onset_strength(po, sr, aggregate=agg)
    m, med = pipeline.onset_pair(po.cuda()[None], sr, want_median=True)
    eg = m if agg == "mean" else med
    print("onset", name, "same input", rel(eg[0], eo), "chain", rel((pipeline.onset_pair(pg, sr, want_median=True)[0 if agg == 'mean' else 1])[0], eo))
eo = dsp.onset_strength(po, sr, aggregate="median")
pulse_o = dsp.plp(po, sr)
pulse_same = ops.plp_from_onset(eo.cuda()[None], sr)
print("plp from oracle envelope", rel(pulse_same[0], pulse_o))
eg = pipeline.onset_pair(po.cuda()[None], sr, want_median=True)[1]
print("plp from our envelope (oracle perc)", rel(ops.plp_from_onset(eg, sr)[0], pulse_o))
eg2 = pipeline.onset_pair(pg, sr, want_median=True)[1]
print("plp full chain", rel(ops.plp_from_onset(eg2, sr)[0], pulse_o))
# peak-bin agreement
G = dsp.stft(eo, 1024, 1)
Gg = torch.view_as_complex(ops.frames_fft(eg2, 1024, 1, len(eo) + 1, torch.hann_window, want_complex=True)["complex"])[0].T.cpu()
ko = G.abs()[43:129].argmax(0)
kg = Gg.abs()[43:129].argmax(0)
print("tempogram rel", rel(torch.view_as_real(Gg.contiguous()), torch.view_as_real(G.contiguous())), "peak-bin mismatches", int((ko != kg).sum()), "of", len(ko))
import numpy as np
g = np.load(ROOT / "tests/golden/a2f.npz")
gold = torch.from_numpy(g["features_raw"])[:, 50]
print("oracle(on this host) pulse vs golden(build container)", rel(pulse_o, gold))
from ssar_b200.supervised.data import audio2features
raw = audio2features(x44.cuda(), 44100, 24, clamp=False, smooth=False)
print("gpu a2f pulse vs oracle here", rel(raw[:, 50], pulse_o), "vs golden", rel(raw[:, 50], gold))
print("torch threads", torch.get_num_threads())
