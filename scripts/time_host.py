"""GPU box: wall-clock split of the host side of the full step (where the Python / launch overhead goes)."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch
from ssar_b200 import pipeline
from ssar_b200.random.patch import Patch
from ssar_b200.synth import synth_batch_device, synth_patch_inputs

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
x = synth_batch_device(B, 44100 * 240, 44100, "cuda")
_, segs, palette = synth_patch_inputs(5760, 0)
palette = palette.cuda()

def tm(fn, n=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        r = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3, r

ms, feats = tm(lambda: pipeline.audio2features_batch(x, 44100, 24))
print(f"a2f batch B={B}: {ms:.1f} ms  ({ms/B:.2f} ms/clip)")
pf = pipeline.features_and_patch_inputs(x, 44100, 24)[1]
for dev in ("cpu", "cuda"):
    ms, patch = tm(lambda: Patch(pf, segs, 120.0, seed=42, device=dev))
    print(f"Patch.__init__ rng={dev}: {ms:.2f} ms")
    ms, _ = tm(lambda: patch.forward(palette, downscale_factor=4))
    print(f"Patch.forward rng={dev}: {ms:.2f} ms")
ms, _ = tm(lambda: pipeline.features_and_patch_inputs(x, 44100, 24)[1])
print(f"patch_features_from_matrix: {ms:.2f} ms")
import cProfile, pstats
patch = Patch(pf, segs, 120.0, seed=42, device="cuda")
pr = cProfile.Profile(); pr.enable()
for _ in range(3): patch.forward(palette, downscale_factor=4)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
