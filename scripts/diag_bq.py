import sys
sys.path.insert(0, "self-supervised-audio-reactive_b200"); sys.path.insert(0, "oracle")
import torch
from ssar_b200 import ops, plan
from ssar_b200.synth import synth_clip
from ssar_oracle import dsp
sr = 24576
xs = torch.stack([synth_clip(sr * 7 + 333, sr, i) for i in (2, 6, 9)]).cuda()
lo_c, hi_c = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
(y_lo, _), (y_hi, _) = ops.biquad_pair(xs, lo_c, hi_c)
for name, a, b in (("lo", y_lo, ops.biquad(xs, lo_c)), ("hi", y_hi, ops.biquad(xs, hi_c))):
    d = (a - b).abs()
    idx = d.flatten().argmax().item()
    print(name, "max abs diff", d.max().item(), "at", divmod(idx, a.shape[1]), "n diff", int((d > 0).sum()), "first diff idx", (d[0] > 0).nonzero()[:5].flatten().tolist())
ref_lo, ref_hi = dsp.low_pass(xs[0].cpu(), sr), dsp.high_pass(xs[0].cpu(), sr)
for name, a, b, r in (("lo", y_lo, ops.biquad(xs, lo_c), ref_lo), ("hi", y_hi, ops.biquad(xs, hi_c), ref_hi)):
    print(name, "pair vs oracle", ((a[0].cpu() - r).norm() / r.norm()).item(), "single vs oracle", ((b[0].cpu() - r).norm() / r.norm()).item())
