"""GPU box: cProfile of one full step (32 clips) — where the host time of the step goes."""
import sys, time, cProfile, pstats
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch
from ssar_b200 import pipeline
from ssar_b200.synth import synth_batch_device, synth_patch_inputs

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
rng = sys.argv[2] if len(sys.argv) > 2 else "cuda"
x = synth_batch_device(B, 44100 * 240, 44100, "cuda")
_, segs, palette = synth_patch_inputs(5760, 0)
palette = palette.cuda()
def step():
    return pipeline.features_and_latents(x, 44100, 24, palette, segs, rng_device=rng)
for _ in range(2): step()
torch.cuda.synchronize()
t0 = time.perf_counter(); step(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"step: host {1e3*(t1-t0):.1f} ms, +drain {1e3*(t2-t1):.1f} ms")
t0 = time.perf_counter(); f = pipeline.audio2features_batch(x, 44100, 24); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"a2f: host {1e3*(t1-t0):.1f} ms, +drain {1e3*(t2-t1):.1f} ms")
pr = cProfile.Profile(); pr.enable(); step(); torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(35)
