"""GPU box: where the ragged pipeline's time goes (256 tracks of the cfg5 set): features only vs features + patches, chunk sizes."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200")); sys.path.insert(0, str(ROOT))
import torch
import bench
from ssar_b200 import pipeline
from ssar_b200.synth import synth_batch_device, synth_patch_inputs

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dur = bench.cfg5_durations(1000)[:n]
dev = torch.device("cuda:0")
clips = [synth_batch_device(1, int(round(d * 44100)), 44100, dev, first_index=i)[0] for i, d in enumerate(dur)]
_, segs, palette = synth_patch_inputs(360 * 24 + 2, 0)
palette, segs = palette.to(dev), {k: v.to(dev) for k, v in segs.items()}
total = sum(dur)

def feats_only(chunk):
    order = sorted(range(n), key=lambda i: -clips[i].shape[0])
    bucket = 30 * 44100
    for c0 in range(0, n, chunk):
        idxs = order[c0:c0 + chunk]
        cap = -(-clips[idxs[0]].shape[0] // bucket) * bucket
        g = pipeline.RaggedFeatureGraph.get(chunk, cap, 44100, 24, dev)
        g.run([clips[i] for i in idxs], deferred_checks=[])

for chunk in (8, 16):
    for name, fn in (("features only", lambda: feats_only(chunk)),
                     ("features + patches", lambda: pipeline.features_and_latents_ragged(clips, 44100, 24, palette, segs, chunk=chunk, on_clip=lambda i, f, l: None))):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter(); fn(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
        print(f"chunk {chunk:2d} {name:20s}: host {1e3*(t1-t0):7.1f} ms, total {1e3*(t2-t0):7.1f} ms -> {total/(t2-t0):8.0f} audio-s/s")
