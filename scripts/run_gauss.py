"""gauss1d sigma 80 (reflect) on the salience matrix of one chunk [8, 5760, 49]: the command ncu profiles."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch  # noqa: E402

from ssar_b200 import ops  # noqa: E402

x = torch.rand(8, 5760, 49, device="cuda")
for sigma, mode in ((80, "reflect"), (5, "reflect"), (2, "circular"), (10, "circular")):
    for _ in range(3):
        y = ops.gauss1d(x, sigma, mode, batched=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        y = ops.gauss1d(x, sigma, mode, batched=True)
    e1.record()
    torch.cuda.synchronize()
    print(f"gauss1d sigma {sigma} {mode}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us")
