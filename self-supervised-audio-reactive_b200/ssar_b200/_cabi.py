"""ctypes binding of `libssar_b200.so` (the C ABI declared in `include/ssar_b200.h`).

There is deliberately NO fallback: if the library is missing or a tensor is not a CUDA tensor the call raises.
PyTorch is used for device memory, streams and `torch.distributed` only.
"""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

import torch

LIB_PATH = Path(__file__).resolve().parent / "libssar_b200.so"

_lib = None
launch_count = 0  # number of C-ABI kernel entry points invoked (bench.py reports it as gpu_launches)
event_log = None  # optional {entry point name: [(start_event, stop_event), ...]} filled by call()


class SsarError(RuntimeError):
    pass


class Filterbank(C.Structure):
    _fields_ = [
        ("n_bands", C.c_int), ("on_power", C.c_int),
        ("start", C.c_void_p), ("count", C.c_void_p), ("offset", C.c_void_p), ("weights", C.c_void_p),
        ("out", C.c_void_p), ("ld", C.c_longlong), ("max_key", C.c_void_p),
        ("lane_start4", C.c_void_p), ("lane_count4", C.c_void_p), ("lane_offset4", C.c_void_p),
        ("lane_weights", C.c_void_p),
    ]


class Projection(C.Structure):
    _fields_ = [
        ("n_rows", C.c_int),
        ("start", C.c_void_p), ("count", C.c_void_p), ("offset", C.c_void_p), ("weights", C.c_void_p),
        ("out", C.c_void_p), ("ld", C.c_longlong), ("scale", C.c_float),
        ("clip_slot", C.c_void_p), ("row_scale", C.c_void_p), ("row_scale_ld", C.c_longlong),
    ]


class PiptrackEpilogue(C.Structure):
    _fields_ = [
        ("count", C.c_void_p), ("pitch", C.c_void_p), ("mag", C.c_void_p), ("capacity", C.c_longlong),
        ("k_lo", C.c_int), ("k_hi", C.c_int), ("threshold", C.c_float), ("sr", C.c_float),
    ]


class FramesArgs(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("n_samples", C.c_longlong), ("clip_stride", C.c_longlong),
        ("n_clips", C.c_int), ("n_fft", C.c_int), ("hop", C.c_int), ("n_frames", C.c_int),
        ("window", C.c_void_p), ("twiddle", C.c_void_p),
        ("out_complex", C.c_void_p), ("ldc", C.c_longlong),
        ("out_spec", C.c_void_p), ("lds", C.c_longlong), ("spec_power", C.c_int),
        ("out_rms", C.c_void_p), ("out_flatness", C.c_void_p),
        ("fb", Filterbank * 2), ("proj", Projection), ("aux_frames", C.c_int), ("pip", PiptrackEpilogue),
        ("clip_samples", C.c_void_p), ("clip_frames", C.c_void_p),
        ("sample_offset", C.c_longlong), ("clip_total", C.c_longlong), ("frame_offset", C.c_int),
        ("sub_every", C.c_int),
    ]


class NoiseOp(C.Structure):
    _fields_ = [
        ("kind", C.c_int), ("M", C.c_int), ("to_acc", C.c_int), ("reserved", C.c_int),
        ("noise", C.c_void_p), ("mod", C.c_void_p), ("p0", C.c_float), ("p1", C.c_float),
    ]


class SubpatchDesc(C.Structure):
    """ssar_subpatch_desc (include/ssar_b200.h)."""
    _fields_ = [
        ("selection", C.c_void_p), ("src", C.c_void_p), ("mod", C.c_void_p), ("pre_taps", C.c_void_p),
        ("src_ld", C.c_longlong), ("mod_ld", C.c_longlong),
        ("src_scale", C.c_float), ("mod_scale", C.c_float), ("alpha", C.c_float), ("beta", C.c_float),
        ("g_begin", C.c_int), ("g_end", C.c_int), ("n_select", C.c_int), ("src_mode", C.c_int), ("radius", C.c_int),
        ("pre_radius", C.c_int),
    ]


def lib():
    """Loads the library once; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise SsarError(
                f"{LIB_PATH} not found: build the CUDA library first (python -m ssar_b200._build or "
                "__graft_entry__.build()). ssar_b200 has no CPU or PyTorch fallback."
            )
        _lib = C.CDLL(str(LIB_PATH))
        _lib.ssar_last_error.restype = C.c_char_p
        _lib.ssar_biquad_scratch_bytes.restype = C.c_longlong
        _lib.ssar_piptrack_capacity.restype = C.c_longlong
        _lib.ssar_mix_gemm_tc_scratch_bytes.restype = C.c_longlong
        if _lib.ssar_abi_version() != 1:
            raise SsarError("libssar_b200.so ABI version mismatch; rebuild")
    return _lib


def ptr(t):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise SsarError("ssar_b200 kernels need CUDA tensors (no CPU fallback)")
    return C.c_void_p(t.data_ptr())


_count_lock = threading.Lock()
_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr():
    """torch's current stream on the current device as a raw cudaStream_t (the fast private accessor when present:
    `torch.cuda.current_stream()` costs ~15 us per call, which matters at ~2 000 launches per step)."""
    if _raw_stream is not None:
        return C.c_void_p(_raw_stream(torch.cuda.current_device()))
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def call(name: str, *args, nbytes=None):
    """Invoke a C-ABI entry point on the current CUDA stream and raise on a non-zero status.
    nbytes: algorithmic bytes of this launch (compulsory inputs + outputs), recorded for the roofline when timing."""
    global launch_count
    fn = getattr(lib(), name)
    log = event_log.get(name) if event_log is not None else None
    if log is not None:  # bench.py: CUDA events on the launching stream around exactly this launch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = fn(*args, stream_ptr())
        e1.record()
        log.append((e0, e1, nbytes))
    else:
        rc = fn(*args, stream_ptr())
    with _count_lock:  # patches of different clips are built by several host threads
        launch_count += 1
    if rc != 0:
        raise SsarError(f"{name} failed ({rc}): {lib().ssar_last_error().decode()}")


def nbytes_of(*tensors):
    return sum(t.numel() * t.element_size() for t in tensors if t is not None)


def i32(v):
    return C.c_int(int(v))


def i64(v):
    return C.c_longlong(int(v))


def f32(v):
    return C.c_float(float(v))


def f64(v):
    return C.c_double(float(v))
