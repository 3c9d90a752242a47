"""One chunk (8 clips x 240 s at 24 576 Hz) through ops.cqt, a few times: the command ncu profiles for ssar_cqt_tc."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch  # noqa: E402

from ssar_b200 import ops  # noqa: E402
from ssar_b200.synth import synth_batch_device  # noqa: E402

sr = 24576
x = synth_batch_device(8, sr * 240, sr, torch.device("cuda:0"))
idx, _ = ops.estimate_tuning(x, sr, bins_per_octave=36)
tc = "fft" not in sys.argv
for _ in range(3):
    C = ops.cqt(x, sr, tuning_idx=idx, tensor_core=tc)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    C = ops.cqt(x, sr, tuning_idx=idx, tensor_core=tc)
e1.record()
torch.cuda.synchronize()
print("cqt", "tensor-core" if tc else "fft", "ms per 8-clip chunk (incl. 6 decimations):", e0.elapsed_time(e1) / 5, float(C.abs().mean()))
