import sys
sys.path.insert(0, "self-supervised-audio-reactive_b200"); sys.path.insert(0, "oracle")
import torch
from ssar_b200 import pipeline
from ssar_b200.synth import synth_clip, synth_patch_inputs
from ssar_oracle import dsp, features, traj, post, cqt as ocqt

def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm())

for secs in (16, 45):
    x = synth_clip(44100 * secs, 44100, 5)
    x24 = dsp.resample(x, 44100, 24576)
    pf_ref = features.retrieve_music_features(x24, 24576)
    f, pf = pipeline.features_and_patch_inputs(x.cuda()[None], 44100, 24)
    print(secs, {k: f"{rel(pf[k][0], v):.2e}" for k, v in pf_ref.items()})
    raw_ref = {name: fn(x24, 24576) for name, fn in features.AFEATFNS.items()}
    ch = pipeline.feature_chain(pipeline.resample(x.cuda()[None], 44100, 24576) if hasattr(pipeline, "resample") else __import__("ssar_b200.resampling", fromlist=["resample"]).resample(x.cuda()[None], 44100, 24576), 24576, want_a2f=False, want_mir=True)["mir"][0]
    o = 0
    for (name, n) in pipeline.MIR_GROUPS:
        print("  raw", name, f"{rel(ch[:, o:o+n], raw_ref[name]):.2e}", "max abs", float((ch[:, o:o+n].cpu() - raw_ref[name]).abs().max()))
        o += n
    _, segs, palette = synth_patch_inputs(f.shape[1], 0)
    lat_ref, _ = traj.Patch(pf_ref, segs, 120.0, fps=24, seed=42).forward(palette, downscale_factor=4)
    lat_g, _ = traj.Patch({k: v[0].cpu() for k, v in pf.items()}, segs, 120.0, fps=24, seed=42).forward(palette, downscale_factor=4)
    print("  oracle Patch on GPU features vs on oracle features:", rel(lat_g, lat_ref))
