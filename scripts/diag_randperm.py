"""GPU box: which assumption of ssar_philox_randperm differs from torch.randperm (offset increment, key bits, ...)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import ctypes as C
import torch
from ssar_b200 import _cabi as cabi, ops
dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 180
g = torch.Generator(dev); g.manual_seed(123)
ref = []
offs = [0]
for _ in range(64):
    ref.append(torch.randperm(n, generator=g, device=dev)); offs.append(g.get_offset())
ref = torch.stack(ref)
print("offsets:", offs[:6], "formula bits", ops.randperm_bits(n))
inc = offs[1]
# raw keys as torch would draw them: random_(INT_MIN, INT_MAX) on int32
g2 = torch.Generator(dev); g2.manual_seed(123)
keys = torch.empty(n, dtype=torch.int32, device=dev).random_(-2**31, 2**31 - 1, generator=g2)
print("random_ increment:", g2.get_offset(), "keys[:4]", keys[:4].tolist())
for bits in range(8, 33):
    out = torch.empty(64, n, dtype=torch.int64, device=dev)
    cabi.call("ssar_philox_randperm", C.c_longlong(123), C.c_longlong(0), C.c_longlong(inc), C.c_int(n), C.c_int(bits),
              C.c_int(64), cabi.ptr(out))
    rows = (out == ref).all(dim=1).sum().item()
    print(bits, "rows equal:", rows, "first row equal elems:", (out[0] == ref[0]).sum().item())
# does sorting torch's own keys by low bits reproduce row 0?
for bits in (18, 24, 31, 32):
    k = (keys.long() & ((1 << bits) - 1))
    order = torch.sort(k, stable=True).indices
    print("torch keys, bits", bits, "match row0:", (order == ref[0]).sum().item(), "/", n)
