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
po = dsp.percussive(xo)
pulse_o = dsp.plp(po, 24576)
eo = dsp.onset_strength(po, 24576, aggregate="median")
x = resample(x44.cuda()[None], 44100, 24576)
harm, perc = pipeline.hpss_pair(x)
print("perc", rel(perc[0], po))
on, on_med = pipeline.onset_pair(perc, 24576, want_median=True)
print("on_med", rel(on_med[0], eo))
p1 = ops.plp_from_onset(on_med, 24576, 1024)
print("pulse", rel(p1[0], pulse_o))
bs = pipeline.band_split(perc, 24576)
on2, on_med2 = pipeline.onset_pair(perc, 24576, want_median=True)
print("on_med after band_split", rel(on_med2[0], eo))
p2 = ops.plp_from_onset(on_med2, 24576, 1024)
print("pulse2", rel(p2[0], pulse_o))
ons = [pipeline.onset_pair(b, 24576)[0] for b in bs]
p3 = ops.plp_from_onset(on_med2, 24576, 1024)
print("pulse3", rel(p3[0], pulse_o))
for i in range(3):
    print("repeat", rel(ops.plp_from_onset(on_med2, 24576, 1024)[0], pulse_o))
