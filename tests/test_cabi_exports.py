"""CPU: the C-ABI library loads and exports every symbol include/ssar_b200.h declares (no compute calls)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "ssar_b200.h"


def declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(ssar_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_entry_points():
    syms = declared_symbols()
    assert "ssar_frames_fft" in syms and "ssar_last_error" in syms and len(syms) >= 8


def test_library_exports_all_declared_symbols():
    from ssar_b200 import _build, _cabi

    if not _cabi.LIB_PATH.exists():
        _build.build()
    lib = ctypes.CDLL(str(_cabi.LIB_PATH))
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, f"declared in the header but not exported: {missing}"
    assert lib.ssar_abi_version() == 1


def test_no_cpu_fallback():
    """The product path must fail loudly on CPU tensors instead of computing somewhere else."""
    import torch

    from ssar_b200._cabi import SsarError
    from ssar_b200.features.rosa import spectral

    with pytest.raises(SsarError):
        spectral.spectrogram(torch.zeros(4096))


def test_product_does_not_import_oracle():
    pkg = ROOT / "self-supervised-audio-reactive_b200" / "ssar_b200"
    for f in pkg.rglob("*.py"):
        assert "ssar_oracle" not in f.read_text(), f
