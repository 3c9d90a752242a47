"""One 4-minute track's latent render (weights [5760, 3, 180] x palette [180, 18, 512]) through ops.mix_gemm: the
command ncu profiles for ssar_mix_gemm_tc.  Prints the CUDA-event time per call with and without a pre-split palette."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch  # noqa: E402

from ssar_b200 import ops  # noqa: E402

g = torch.Generator("cpu").manual_seed(1)
W = torch.rand(5760, 3, 180, generator=g).cuda()
P = torch.randn(180, 18, 512, generator=g).cuda()
outs = [torch.empty(5760, 18, 512, device="cuda") for _ in range(8)]  # 8 x 212 MB: successive calls miss L2
sp = ops.SplitBasis(P)
for label, kw in (("per-call split", {}), ("pre-split palette", {"basis_split": sp})):
    for o in outs[:3]:
        ops.mix_gemm(W, P, out=o, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(4):
        for o in outs:
            ops.mix_gemm(W, P, out=o, **kw)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 32 * 1e3
    print(f"mix_gemm_tc {label}: {us:.1f} us per track, {212.336e6 / us / 1e3:.0f} GB/s of output")
