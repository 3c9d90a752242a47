"""GPU box: host time of the per-clip Patch work of one step (sampling + bind + forward), without a profiler."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch
from ssar_b200 import ops
from ssar_b200.random.patch import Patch
from ssar_b200.synth import synth_patch_inputs

T = 5760
feats, segs, palette = synth_patch_inputs(T, 0)
dev = torch.device("cuda:0")
palette = palette.to(dev)
feats = {k: v.to(dev) for k, v in feats.items()}
lat = torch.empty(8, T, 18, 512, device=dev)
sp = ops.SplitBasis(palette)
for rep in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    patches = Patch.sample_batch([42 + i for i in range(32)], segs, 120.0, fps=24, device="cuda")
    t1 = time.perf_counter()
    for i, p in enumerate(patches):
        p.bind(feats)
        p.forward(palette, downscale_factor=4, latents_out=lat[i % 8], palette_split=sp)
    t2 = time.perf_counter()
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    print(f"rep {rep}: sample_batch(32) {1e3*(t1-t0):.1f} ms, 32 x forward host {1e3*(t2-t1):.1f} ms, drain {1e3*(t3-t2):.1f} ms")
