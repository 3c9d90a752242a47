"""GPU box: per-entry-point CUDA-event time of one audio2features_batch + Patch.forward pass."""
import sys
from collections import defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch

from ssar_b200 import _cabi as cabi
from ssar_b200 import pipeline
from ssar_b200.random.patch import Patch
from ssar_b200.synth import synth_batch_device, synth_patch_inputs

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
x = synth_batch_device(B, 44100 * 240, 44100, "cuda")
feats, segs, palette = synth_patch_inputs(5760, 0)
feats = {k: v.cuda() for k, v in feats.items()}
palette = palette.cuda()
patch = Patch(feats, segs, 120.0, seed=42, device="cpu")


class Log(dict):
    def get(self, k, d=None):
        return self.setdefault(k, [])


def step():
    f = pipeline.audio2features_batch(x, 44100, 24)
    lat, noise = patch.forward(palette, downscale_factor=4)
    return f, lat


for _ in range(2):
    step()
torch.cuda.synchronize()
cabi.event_log = Log()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
step()
e1.record()
torch.cuda.synchronize()
tot = e0.elapsed_time(e1)
rows = sorted(((sum(e[0].elapsed_time(e[1]) for e in v), len(v), k) for k, v in cabi.event_log.items()), reverse=True)
cabi.event_log = None
print(f"B={B} total {tot:.2f} ms  -> {B*240/(tot*1e-3):.0f} audio-s/s")
for ms, n, k in rows:
    print(f"{ms:9.3f} ms {n:4d}x {100*ms/tot:5.1f}%  {k}")
import time
t0 = time.perf_counter(); step(); torch.cuda.synchronize(); print("wall", (time.perf_counter() - t0) * 1e3, "ms")
print("mem GB", torch.cuda.max_memory_allocated() / 1e9)
