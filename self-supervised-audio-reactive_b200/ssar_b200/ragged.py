"""Per-clip lengths of a RAGGED batch (clips of different durations padded to one capacity and processed by the same
launches — SURVEY 8d cfg 5: 1000 tracks of 120..360 s).  Every derived count the kernels need (samples at the working
rate, frame counts of the hop-1024 / hop-512 transforms, the octave signal lengths of the CQT chain) lives in two device
buffers with FIXED addresses: a CUDA graph captured with them serves any set of lengths after `update()`.

Convention shared by all ragged kernels: a clip's row holds its valid data followed by ZEROS up to the capacity; a kernel
treats the clip's own length as the end of the signal (reflection, wrap-around, window envelope, maxima, quantiles) and
writes zeros past it.  A batch then gives, for every clip, exactly the result of a per-clip call
(tests/test_gpu_ragged.py)."""
from __future__ import annotations

import torch

N_OCT = 7


class ClipLengths:
    I64_ROWS = 3 + N_OCT  # n_in, n, T (as int64), n_oct[0..6]
    I32_ROWS = 4          # T, T + 1, n // 512, T + 1 for the hop-1 tempogram (= T + 1)

    def __init__(self, n_clips: int, device):
        self.B, self.device = int(n_clips), torch.device(device)
        self._i64 = torch.zeros(self.I64_ROWS, self.B, dtype=torch.int64, device=self.device)
        self._i32 = torch.zeros(self.I32_ROWS, self.B, dtype=torch.int32, device=self.device)
        # pinned staging for update(): a ring, because the host may queue several chunks ahead of the GPU and an
        # asynchronous copy reads its pinned source when it EXECUTES (slot reuse waits on the slot's last copy)
        self._ring = [(torch.zeros(self.I64_ROWS, self.B, dtype=torch.int64).pin_memory(),
                       torch.zeros(self.I32_ROWS, self.B, dtype=torch.int32).pin_memory(), [None]) for _ in range(4)]
        self._slot = 0
        self.host_n_in = self.host_n = self.host_T = None

    def update(self, n_in, orig: int, new: int, hop: int = 1024):
        """n_in: samples of each clip at the input rate `orig`; the clips are resampled to `new` (= 1024 * fps)."""
        assert len(n_in) == self.B
        g = __import__("math").gcd(int(orig), int(new))
        o, nw = int(orig) // g, int(new) // g
        n_in = [int(v) for v in n_in]
        n = [-((-nw * v) // o) for v in n_in]          # ceil(new * n_in / orig): torchaudio's output length
        T = [v // hop for v in n]
        self.host_n_in, self.host_n, self.host_T = n_in, n, T
        h64, h32, ev = self._ring[self._slot]
        self._slot = (self._slot + 1) % len(self._ring)
        if ev[0] is not None:
            ev[0].synchronize()
        h64[0] = torch.tensor(n_in)
        h64[1] = torch.tensor(n)
        h64[2] = torch.tensor(T)
        cur = list(n)
        for q in range(N_OCT):
            h64[3 + q] = torch.tensor(cur)
            # every octave must offer at least T frames (rosa/constantq.py:106-110 trims to the shortest octave)
            assert all(c // (hop >> q) >= t for c, t in zip(cur, T)), "CQT octave shorter than the frame count"
            cur = [(c + 1) // 2 for c in cur]
        h32[0] = torch.tensor(T, dtype=torch.int32)
        h32[1] = torch.tensor([t + 1 for t in T], dtype=torch.int32)
        h32[2] = torch.tensor([v // (hop // 2) for v in n], dtype=torch.int32)
        h32[3] = h32[1]
        self._i64.copy_(h64, non_blocking=True)
        self._i32.copy_(h32, non_blocking=True)
        ev[0] = torch.cuda.Event()
        ev[0].record(torch.cuda.current_stream(self.device))
        return self

    # int64 views
    n_in = property(lambda s: s._i64[0])      # samples at the input rate
    n = property(lambda s: s._i64[1])         # samples at the working rate
    T64 = property(lambda s: s._i64[2])       # frames (hop) as int64: the "sample count" of a per-frame envelope
    # int32 views
    T = property(lambda s: s._i32[0])         # frames of the last-frame-dropping features
    T1 = property(lambda s: s._i32[1])        # frames of the complex STFT (1 + n // hop)
    P = property(lambda s: s._i32[2])         # frames of piptrack's hop / 2 transform

    def octave_n(self, q: int):
        return self._i64[3 + q]
