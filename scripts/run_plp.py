"""plp_from_onset over one chunk (8 clips x 5760 onset frames): tempogram STFT (n_fft 1024, hop 1), band mask / peak pick,
inverse transform and the hop-1 overlap-add.  The command ncu profiles for overlap_add_dense_kernel."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch  # noqa: E402

from ssar_b200 import ops  # noqa: E402

g = torch.Generator().manual_seed(3)
on = torch.rand(8, 5760, generator=g).cuda()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for _ in range(2):
    ops.plp_from_onset(on, 24576, 1024)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    p = ops.plp_from_onset(on, 24576, 1024)
e1.record()
torch.cuda.synchronize()
print(f"plp_from_onset: {e0.elapsed_time(e1) / reps * 1e3:.1f} us per chunk")
