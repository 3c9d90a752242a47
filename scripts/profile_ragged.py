"""GPU box: host profile of the ragged pipeline (64 tracks of U[120, 360] s): where the per-chunk host time goes."""
import cProfile
import pstats
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

import bench  # noqa: E402
from ssar_b200 import pipeline  # noqa: E402
from ssar_b200.synth import synth_batch_device, synth_patch_inputs  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dur = bench.cfg5_durations(1000)[:n]
dev = torch.device("cuda:0")
clips = [synth_batch_device(1, int(round(d * 44100)), 44100, dev, first_index=i)[0] for i, d in enumerate(dur)]
_, segs, palette = synth_patch_inputs(360 * 24 + 2, 0)
palette, segs = palette.to(dev), {k: v.to(dev) for k, v in segs.items()}
stats = torch.zeros(3, 59, dtype=torch.float64, device=dev)


def step():
    pipeline.features_and_latents_ragged(clips, 44100, 24, palette, segs, on_clip=lambda i, f, lat: None, stats=stats)


for _ in range(2):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"{n} tracks, {sum(dur):.0f} audio-s: host {1e3 * (t1 - t0):.1f} ms, + drain {1e3 * (t2 - t1):.1f} ms -> {sum(dur) / (t2 - t0):.0f} audio-s/s")
pr = cProfile.Profile()
pr.enable()
step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
