"""GPU: `ssar_biquad_static` (the no-allocation variant of `ssar_biquad`: scan scratch in a static device array inside
the library) called through the C ABI, against `ssar_biquad` with caller-provided scratch on the same inputs — two calls
in a row with different inputs and sizes (the array keeps the previous call's contents), the size limit, and no device
allocation by the call.  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from ssar_b200.synth import synth_clip

pytestmark = pytest.mark.gpu


def _static(x, coeffs, clamp=True):
    from ssar_b200 import _cabi as cabi

    B, N = x.shape
    y = torch.empty_like(x)
    b0, b1, b2, a1, a2 = coeffs
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    free_before = torch.cuda.mem_get_info()[0]
    cabi.call("ssar_biquad_static", cabi.ptr(x), cabi.i64(N), cabi.i64(x.stride(0)), cabi.i32(B), cabi.f32(b0), cabi.f32(b1),
              cabi.f32(b2), cabi.f32(a1), cabi.f32(a2), cabi.i32(1 if clamp else 0), cabi.ptr(y), cabi.i64(N))
    torch.cuda.synchronize()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before  # no torch allocation
    assert torch.cuda.mem_get_info()[0] >= free_before                               # no cudaMalloc behind torch's back
    return y


def test_biquad_static_equals_biquad_twice_in_a_row():
    from ssar_b200 import ops, plan

    sr = 24576
    lo, hi = plan.biquad_coeffs("low", sr, 200), plan.biquad_coeffs("high", sr, 4000)
    xa = (torch.stack([synth_clip(sr * 21 + 313, sr, i) for i in (1, 4, 9)]) * 1.5).cuda()   # odd length, clamping
    xb = torch.stack([synth_clip(sr * 9 + 5, sr, i) for i in (2, 3)]).cuda()                  # smaller: stale tail in scratch
    ops.biquad(xa[:1, :4096], lo)  # (library + allocator warm before the allocation checks)
    for x, c in ((xa, lo), (xb, hi), (xa, hi)):
        got = _static(x, c)
        ref = ops.biquad(x, c)
        assert torch.equal(got, ref)  # same kernels, same arithmetic: tolerance 0
    assert float(got.abs().max()) <= 1.0


def test_biquad_static_rejects_sizes_beyond_the_array():
    from ssar_b200 import _cabi as cabi

    lib = cabi.lib()
    n_clips, n = 64, 24576 * 240  # 46 080 spans x 96 bytes = 4.4 MB of scratch: over the 2 MiB array
    assert lib.ssar_biquad_scratch_bytes(C.c_int(n_clips), C.c_longlong(n)) > (2 << 20)
    x = torch.zeros(1, 16, device="cuda")  # never touched: the size check comes first
    with pytest.raises(cabi.SsarError, match="static array holds"):
        cabi.call("ssar_biquad_static", cabi.ptr(x), cabi.i64(n), cabi.i64(n), cabi.i32(n_clips), cabi.f32(1), cabi.f32(0),
                  cabi.f32(0), cabi.f32(0), cabi.f32(0), cabi.i32(1), cabi.ptr(x), cabi.i64(n))
