"""GPU box: warm-up step(s) + ONE profiled full step (B clips) with a single launching thread — the command
`ncu --profile-from-start off` wraps for the launch list of the full path (every launch serialised, so per-kernel SHARES are comparable with bench.py's)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch
from ssar_b200 import ops, pipeline
from ssar_b200.synth import synth_batch_device, synth_patch_inputs

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
x = synth_batch_device(B, 44100 * 240, 44100, "cuda")
_, segs, palette = synth_patch_inputs(5760, 0)
palette = palette.cuda()
stats = torch.zeros(3, 59, dtype=torch.float64, device="cuda")
for i in range(n_steps):
    if i == n_steps - 1:
        # ncu --profile-from-start off: only the LAST step is listed (the first one also runs the one-time checks of
        # the batched Philox draws against torch's per-call draws: ~600 torch launches that never recur)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    feats, latents, _ = pipeline.features_and_latents(x, 44100, 24, palette, segs, rng_device="cuda", serial=True)
    stats.zero_()
    ops.feature_stats(feats, stats)
    torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok", feats.shape, latents.shape, float(stats[1].sum()))
