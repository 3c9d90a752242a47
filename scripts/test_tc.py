"""GPU box: tensor-core mix GEMM vs the CUDA-core fp32 kernel (numerics + time)."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch
from ssar_b200 import ops
g = torch.Generator("cpu").manual_seed(0)
for T, G, K in [(256, 3, 180), (5760, 3, 180), (5760, 3, 37), (1000, 1, 64), (333, 3, 192)]:
    W = torch.rand(T, G, K, generator=g).cuda() * (torch.rand(T, G, K, generator=g).cuda() < 0.5)
    P = torch.randn(K, 18, 512, generator=g).cuda()
    ops.USE_TENSOR_CORE_MIX = False
    ref = ops.mix_gemm(W, P)
    ops.USE_TENSOR_CORE_MIX = True
    out = ops.mix_gemm(W, P)
    torch.cuda.synchronize()
    ref64 = torch.einsum("tgk,kgn->tgn", W.double(), P.reshape(K, G, -1).double()).reshape(T, 18, 512)
    e = lambda a: float((a.double() - ref64).norm() / ref64.norm())
    print(T, G, K, "tc rel-L2 vs fp64:", e(out), " fp32 kernel vs fp64:", e(ref), " max abs diff:", float((out - ref).abs().max()))
    for flag in (False, True):
        ops.USE_TENSOR_CORE_MIX = flag
        for _ in range(3): ops.mix_gemm(W, P, out=out)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): ops.mix_gemm(W, P, out=out)
        e1.record(); torch.cuda.synchronize()
        print("   tc" if flag else "   fp32", f"{e0.elapsed_time(e1) / 10 * 1e3:.1f} us")
