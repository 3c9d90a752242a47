"""Band filters of one chunk (8 clips x 240 s at 24576 Hz): the low / high / mid chain with signals and with RMS only,
the pair and the single filter.  The command ncu profiles."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch  # noqa: E402

from ssar_b200 import ops, plan  # noqa: E402
from ssar_b200.synth import synth_batch_device  # noqa: E402

sr, n = 24576, 24576 * 240
x = synth_batch_device(8, n, sr, torch.device("cuda"), first_index=0)
lo, hi = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
T = n // 1024
cases = {
    "chain signals": lambda: ops.biquad_chain(x, lo, hi, lo),
    "chain rms": lambda: ops.biquad_chain(x, lo, hi, lo, want_signal=(False,) * 3, want_rms=(True,) * 3, n_frames=T),
    "pair signals": lambda: ops.biquad_pair(x, lo, hi),
    "single": lambda: ops.biquad(x, lo),
}
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for name, fn in cases.items():
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / reps * 1e3:.1f} us")
