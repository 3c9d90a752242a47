"""GPU box: host wall time vs GPU time of the three parts of the full step (features, patch sampling, patch rendering)."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch
from ssar_b200 import pipeline
from ssar_b200.random.patch import Patch, precapture_sampler_graphs
from ssar_b200.synth import synth_batch_device, synth_patch_inputs

B = 32
x = synth_batch_device(B, 44100 * 240, 44100, "cuda")
_, segs, palette = synth_patch_inputs(5760, 0)
palette = palette.cuda()
precapture_sampler_graphs(torch.device("cuda"), segs)

def timed(fn, n=3):
    fn(); torch.cuda.synchronize()
    hs, gs = [], []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); t0 = time.perf_counter(); e0.record()
        r = fn()
        e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
        hs.append((t1 - t0) * 1e3); gs.append(e0.elapsed_time(e1))
    return min(hs), min(gs), r

for C in (8, 32):
    h, g, _ = timed(lambda: [pipeline.audio2features_batch(x[c:c + C], 44100, 24, deferred_checks=[]) for c in range(0, B, C)])
    print(f"a2f 32 clips in chunks of {C}: host {h:.1f} ms, gpu {g:.1f} ms")
feats = pipeline.audio2features_batch(x, 44100, 24)
h, g, pf = timed(lambda: pipeline.features_and_patch_inputs(x, 44100, 24)[1])
print(f"patch_features_batch: host {h:.2f} ms gpu {g:.2f} ms")
h, g, patches = timed(lambda: [Patch(None, segs, 120.0, seed=42 + b, device="cuda") for b in range(B)])
print(f"sampling 32 patches: host {h:.1f} ms, gpu {g:.1f} ms")
lat = torch.empty(B, 5760, 18, 512, device="cuda")
def render():
    out = []
    for b in range(B):
        patches[b].bind({k: v[b] for k, v in pf.items()})
        out.append(patches[b].forward(palette, downscale_factor=4, latents_out=lat[b])[1])
    return out
h, g, _ = timed(render)
print(f"rendering 32 patches: host {h:.1f} ms, gpu {g:.1f} ms")
import cProfile, pstats
pr = cProfile.Profile(); pr.enable(); render(); torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(16)
h, g, _ = timed(lambda: pipeline.features_and_latents(x, 44100, 24, palette, segs))
print(f"features_and_latents: host {h:.1f} ms, gpu {g:.1f} ms")
