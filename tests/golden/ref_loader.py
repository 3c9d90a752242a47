"""Import the UNMODIFIED reference (`/root/reference/ssar/...`) in this container, on CPU.

Test infrastructure only: used by `make_golden.py` to generate the committed fixtures and by the
`not gpu` tests (when `/root/reference` exists) to cross-check `oracle/` against the real thing.
Nothing here is shipped, and nothing here is reachable from the product package.

The reference cannot be imported as-is: `ssar/__init__.py` does `from . import *`,
several third-party packages are absent, and `processing.py` needs the compiled
`efficient_quantile` extension.  So we build a throw-away package tree under /tmp that *symlinks*
the reference's source files (no copies enter the repo), add empty `__init__.py`s, compile
`efficient_quantile.cpp` from where it lies, and register import shims for the missing packages:

* ``librosa.core.convert.note_to_hz``  – only "C1" is ever asked for (`rosa/convert.py:129-130`).
* ``torchcubicspline`` – restated from the published upstream algorithm
  (patrick-kidger/torchcubicspline, `interpolate.py`, unpinned by the reference's `env.yaml`):
  natural cubic spline, coefficients returned as ``(t, a, b, two_c, three_d)``.
  PARITY UNPINNED at this boundary: the package is not installable here.
* empty stubs for ``ffcv.*``, ``npy_append_array``, ``torch_geometric.utils``, ``x_transformers``
  and the sibling model modules that `latent_n_noise2.py` imports but the path never calls.
"""
from __future__ import annotations

import importlib
import math
import os
import sys
import types
from pathlib import Path

import torch

REFERENCE_ROOT = Path(os.environ.get("SSAR_REFERENCE_ROOT", "/root/reference"))
TREE = Path(os.environ.get("SSAR_REF_TREE", "/tmp/ssar_ref_tree"))


def reference_available() -> bool:
    return (REFERENCE_ROOT / "ssar" / "features" / "audio.py").exists()


# ------------------------------------------------------------------------------------------------
# torchcubicspline shim (published algorithm; see module docstring)
# ------------------------------------------------------------------------------------------------
def _tridiagonal_solve(b, A_upper, A_diagonal, A_lower):
    # Thomas algorithm along the last axis; A_* are 1-D, b is (..., n)
    A_upper, _ = torch.broadcast_tensors(A_upper, b[..., :-1])
    A_lower, _ = torch.broadcast_tensors(A_lower, b[..., :-1])
    A_diagonal, b = torch.broadcast_tensors(A_diagonal, b)
    channels = b.size(-1)
    new_b = [None] * channels
    new_diag = [None] * channels
    outs = [None] * channels
    new_b[0] = b[..., 0]
    new_diag[0] = A_diagonal[..., 0]
    for i in range(1, channels):
        w = A_lower[..., i - 1] / new_diag[i - 1]
        new_diag[i] = A_diagonal[..., i] - w * A_upper[..., i - 1]
        new_b[i] = b[..., i] - w * new_b[i - 1]
    outs[channels - 1] = new_b[channels - 1] / new_diag[channels - 1]
    for i in range(channels - 2, -1, -1):
        outs[i] = (new_b[i] - A_upper[..., i] * outs[i + 1]) / new_diag[i]
    return torch.stack(outs, dim=-1)


def _natural_cubic_spline_coeffs(t, x):
    # x: (..., length, channels)
    path = x.transpose(-1, -2)  # (..., channels, length)
    length = path.size(-1)
    if length < 2:
        raise ValueError("Must have a time dimension of size at least 2.")
    if length == 2:
        a = path[..., :1]
        b = (path[..., 1:] - path[..., :1]) / (t[..., 1:] - t[..., :1])
        two_c = torch.zeros_like(a)
        three_d = torch.zeros_like(a)
    else:
        dt = t[1:] - t[:-1]
        rdt = dt.reciprocal()
        rdt2 = rdt**2
        three_dx = 3 * (path[..., 1:] - path[..., :-1])
        six_dx = 2 * three_dx
        dx_scaled = three_dx * rdt2
        diag = torch.empty(length, dtype=path.dtype, device=path.device)
        diag[:-1] = rdt
        diag[-1] = 0
        diag[1:] += rdt
        diag *= 2
        rhs = torch.empty_like(path)
        rhs[..., :-1] = dx_scaled
        rhs[..., -1] = 0
        rhs[..., 1:] += dx_scaled
        kd = _tridiagonal_solve(rhs, rdt, diag, rdt)
        a = path[..., :-1]
        b = kd[..., :-1]
        two_c = (six_dx * rdt - 4 * kd[..., :-1] - 2 * kd[..., 1:]) * rdt
        three_d = (-six_dx * rdt + 3 * (kd[..., :-1] + kd[..., 1:])) * rdt2
    return (t, a.transpose(-1, -2), b.transpose(-1, -2), two_c.transpose(-1, -2), three_d.transpose(-1, -2))


class _NaturalCubicSpline:
    def __init__(self, coeffs, **kwargs):
        self._t, self._a, self._b, self._two_c, self._three_d = coeffs

    def evaluate(self, t):
        maxlen = self._b.size(-2) - 1
        index = (torch.bucketize(t.detach(), self._t) - 1).clamp(0, maxlen)
        frac = (t - self._t[index]).unsqueeze(-1)
        inner = 0.5 * self._two_c[..., index, :] + self._three_d[..., index, :] * frac / 3
        inner = self._b[..., index, :] + inner * frac
        return self._a[..., index, :] + inner * frac


def _install_shims():
    def mod(name, **attrs):
        m = sys.modules.get(name)
        if m is None:
            m = types.ModuleType(name)
            sys.modules[name] = m
        for k, v in attrs.items():
            setattr(m, k, v)
        return m

    # librosa: only note_to_hz("C1")
    def note_to_hz(note):
        assert note == "C1", note
        return 440.0 * 2.0 ** ((24 - 69) / 12.0)

    conv = mod("librosa.core.convert", note_to_hz=note_to_hz)
    core = mod("librosa.core", convert=conv, note_to_hz=note_to_hz)
    # librosa.beat: tempo / beat tracking are OFF the path (random/mir.py:28-33; inputs of Patch) — fixed answers so that
    # the unmodified retrieve_music_information can run end to end for its feature part
    beat = mod("librosa.beat", tempo=lambda **k: 120.0,
               beat_track=lambda **k: (120.0, list(range(0, len(k["onset_envelope"]), 12))))
    mod("librosa", core=core, note_to_hz=note_to_hz, beat=beat)

    mod(
        "torchcubicspline",
        natural_cubic_spline_coeffs=_natural_cubic_spline_coeffs,
        NaturalCubicSpline=_NaturalCubicSpline,
    )

    class _Anything:
        def __init__(self, *a, **k):
            pass

        def __call__(self, *a, **k):
            return self

        def __getattr__(self, k):
            return _Anything()

    class _Stub(types.ModuleType):
        def __getattr__(self, k):
            if k.startswith("__"):
                raise AttributeError(k)
            return _Anything

    for name in [
        "ffcv", "ffcv.fields", "ffcv.fields.decoders", "ffcv.loader", "ffcv.transforms", "ffcv.writer",
        "npy_append_array", "torch_geometric", "torch_geometric.utils", "x_transformers", "joblib_stub",
        "sklearn_stub",
    ]:
        if name not in sys.modules:
            sys.modules[name] = _Stub(name)


# ------------------------------------------------------------------------------------------------
# throw-away package tree of symlinks
# ------------------------------------------------------------------------------------------------
_LINKS = [
    "ssar/features/audio.py",
    "ssar/features/processing.py",
    "ssar/features/rosa/spectral.py",
    "ssar/features/rosa/constantq.py",
    "ssar/features/rosa/convert.py",
    "ssar/features/rosa/beat.py",
    "ssar/features/rosa/pitch.py",
    "ssar/features/rosa/helpers.py",
    "ssar/random/latent.py",
    "ssar/random/noise.py",
    "ssar/random/patch.py",
    "ssar/supervised/data.py",
    "ssar/models/latent_n_noise2.py",
]


def _build_tree():
    pkgs = ["ssar", "ssar/features", "ssar/features/rosa", "ssar/random", "ssar/supervised", "ssar/models",
            "ssar/analysis", "ssar/analysis/efficient_quantile", "ssar/models/sashimi"]
    for p in pkgs:
        d = TREE / p
        d.mkdir(parents=True, exist_ok=True)
        init = d / "__init__.py"
        if not init.exists():
            init.write_text("")
    for rel in _LINKS:
        dst = TREE / rel
        if not dst.exists():
            dst.symlink_to(REFERENCE_ROOT / rel)
    # random/__init__ defines DEVICE (random/__init__.py:1-3) — keep CPU here
    (TREE / "ssar/random/__init__.py").write_text('import torch\nDEVICE = torch.device("cpu")\n')
    # random/mir.py imports librosa.beat + scipy + segment (off-path); only its three name lists are needed
    (TREE / "ssar/random/mir.py").write_text(
        'UNITFEATS = ["rms", "drop_strength", "onsets", "spectral_flatness"]\n'
        'ALLFEATS = ["chromagram", "tonnetz", "mfcc", "spectral_contrast"] + UNITFEATS\n'
    )
    # the REAL random/mir.py under another name (`mir_real`): its feature part (:14-22, :26, :44) is on the path.  Its
    # off-path imports are satisfied by stand-ins: rosa/segment.py (torch_geometric / sklearn / librosa) by a stub module
    # whose two functions return fixed label tensors, librosa.beat.tempo / beat_track by the shims below.
    dst = TREE / "ssar/random/mir_real.py"
    if not dst.exists():
        dst.symlink_to(REFERENCE_ROOT / "ssar/random/mir.py")
    (TREE / "ssar/features/rosa/segment.py").write_text(
        "import torch\n"
        "def laplacian_segmentation(feature, beats, ks=[2, 4, 6, 8, 12, 16]):\n"
        "    T = feature.shape[0]\n"
        "    return [torch.nn.functional.one_hot(torch.arange(T) % k, k).float() for k in ks]\n"
        "def laplacian_segmentation_rosa(audio, sr, n_frames, ks=[2, 4, 6, 8, 12, 16]):\n"
        "    return torch.stack([torch.arange(n_frames) % k for k in ks], dim=1)\n"
    )
    # sibling model modules imported by latent_n_noise2.py but never executed on the path
    for stub in ["ssar/models/convnext.py", "ssar/models/mlp.py", "ssar/models/audio2latent.py",
                 "ssar/models/sashimi/sashimi.py"]:
        f = TREE / stub
        if not f.exists():
            f.write_text(
                "class _S:\n    def __init__(self,*a,**k): pass\n"
                "ConvNeXt = MLP = Normalize = Sashimi = SaShiMi = LayerwiseConv = LayerwiseLinear = _S\n"
                "def __getattr__(name):\n    return _S\n"
            )


def _build_efficient_quantile():
    """Compile the reference's own C++ quantile from where it lies; outputs only under /tmp."""
    from torch.utils.cpp_extension import load

    build_dir = TREE / "_build_eq"
    build_dir.mkdir(parents=True, exist_ok=True)
    src = REFERENCE_ROOT / "ssar/analysis/efficient_quantile/efficient_quantile.cpp"
    m = load(name="efficient_quantile", sources=[str(src)], build_directory=str(build_dir), verbose=False)
    pkg = importlib.import_module("ssar.analysis.efficient_quantile")
    sub = types.ModuleType("ssar.analysis.efficient_quantile.efficient_quantile")
    sub._efficient_quantile = m._efficient_quantile
    sys.modules["ssar.analysis.efficient_quantile.efficient_quantile"] = sub

    def quantile(tensor, q):  # wrapper semantics of efficient_quantile/__init__.py:5-6
        return m._efficient_quantile(tensor.cpu().flatten(), torch.FloatTensor([q]), True, 3).squeeze().to(tensor.device)

    pkg.quantile = quantile
    pkg._efficient_quantile = m._efficient_quantile
    return m


_LOADED = None


def load_reference(build_quantile: bool = True):
    """Returns a namespace with the reference modules (imported unmodified through symlinks)."""
    global _LOADED
    if _LOADED is not None:
        return _LOADED
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    _install_shims()
    _build_tree()
    if str(TREE) not in sys.path:
        sys.path.insert(0, str(TREE))
    for k in [k for k in sys.modules if k == "ssar" or k.startswith("ssar.")]:
        del sys.modules[k]
    importlib.import_module("ssar")
    if build_quantile:
        _build_efficient_quantile()
    else:
        pkg = importlib.import_module("ssar.analysis.efficient_quantile")
        pkg.quantile = lambda t, q: torch.quantile(t.flatten(), q, interpolation="midpoint")
    ns = types.SimpleNamespace()
    ns.audio = importlib.import_module("ssar.features.audio")  # must come first (circular import otherwise)
    ns.processing = importlib.import_module("ssar.features.processing")
    ns.spectral = importlib.import_module("ssar.features.rosa.spectral")
    ns.constantq = importlib.import_module("ssar.features.rosa.constantq")
    ns.convert = importlib.import_module("ssar.features.rosa.convert")
    ns.beat = importlib.import_module("ssar.features.rosa.beat")
    ns.pitch = importlib.import_module("ssar.features.rosa.pitch")
    ns.helpers = importlib.import_module("ssar.features.rosa.helpers")
    ns.data = importlib.import_module("ssar.supervised.data")
    ns.latent = importlib.import_module("ssar.random.latent")
    ns.noise = importlib.import_module("ssar.random.noise")
    ns.patch = importlib.import_module("ssar.random.patch")
    ns.lnn2 = importlib.import_module("ssar.models.latent_n_noise2")
    import scipy.stats  # noqa: F401  (mir.py:29 uses scipy.stats through a bare `import scipy`)
    ns.mir = importlib.import_module("ssar.random.mir_real")
    ns.spline_coeffs = _natural_cubic_spline_coeffs
    ns.NaturalCubicSpline = _NaturalCubicSpline
    _LOADED = ns
    return ns


if __name__ == "__main__":
    ref = load_reference()
    x = torch.randn(24576 * 3)
    print(ref.audio.rms(x, 24576).shape, ref.audio.spectral_flatness(x, 24576).shape)
    print("C1", ref.convert.note_to_hz("C1"))


def load_maua_audioreactive():
    """The two trajectory files of the vendored maua package (`audioreactive/latent.py`, `audioreactive/signal.py`),
    imported unmodified through symlinks as the throw-away package `maua_ar` (the real package __init__ pulls librosa,
    madmom, openunmix...).  Shims: torchtyping.TensorType (annotations only), torchcubicspline (as above)."""
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    _install_shims()

    class _TT:
        def __class_getitem__(cls, item):
            return cls

    if "torchtyping" not in sys.modules:
        m = types.ModuleType("torchtyping")
        m.TensorType = _TT
        sys.modules["torchtyping"] = m
    d = TREE / "maua_ar"
    d.mkdir(parents=True, exist_ok=True)
    (d / "__init__.py").write_text("")
    for f in ("latent.py", "signal.py"):
        dst = d / f
        if not dst.exists():
            dst.symlink_to(REFERENCE_ROOT / "maua/maua/audiovisual/audioreactive" / f)
    if str(TREE) not in sys.path:
        sys.path.insert(0, str(TREE))
    ns = types.SimpleNamespace()
    ns.signal = importlib.import_module("maua_ar.signal")
    ns.latent = importlib.import_module("maua_ar.latent")
    return ns


def exec_reference_defs(relpath: str, names, namespace: dict) -> dict:
    """Runs the UNMODIFIED source text of the named top-level functions / classes of a reference file whose module
    cannot be imported (`ssar/features/generate.py` loads a GAN pickle and decord at import, `ssar/optimize.py` imports
    maua + matplotlib, `supervised/latent_augmenter.py` imports a name its sibling lacks, the maua StyleGAN2 wrapper
    imports the NVIDIA ops): the file is parsed with `ast`, only the requested definitions are compiled (with the
    reference's own file name and line numbers) and executed in `namespace`, which the caller fills with the reference's
    own imported functions (`load_reference()`) and stand-ins for the globals that are outside the path (the GAN
    mapper, the random audio picker, ...).  Nothing is copied: the source is read from `/root/reference` at run time."""
    import ast

    path = REFERENCE_ROOT / relpath
    tree = ast.parse(path.read_text(), filename=str(path))
    wanted = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in set(names)]
    missing = set(names) - {n.name for n in wanted}
    if missing:
        raise KeyError(f"{relpath}: no top-level definition named {sorted(missing)}")
    mod = ast.Module(body=wanted, type_ignores=[])
    exec(compile(mod, str(path), "exec"), namespace)
    return namespace


def exec_reference_methods(relpath: str, class_name: str, methods, namespace: dict) -> dict:
    """Like `exec_reference_defs` for methods of a class: the named methods are compiled as plain functions (taking
    `self`) — for classes whose constructor needs what is not here (a StyleGAN2 checkpoint)."""
    import ast

    path = REFERENCE_ROOT / relpath
    tree = ast.parse(path.read_text(), filename=str(path))
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == class_name)
    wanted = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in set(methods)]
    missing = set(methods) - {n.name for n in wanted}
    if missing:
        raise KeyError(f"{relpath}:{class_name}: no method named {sorted(missing)}")
    exec(compile(ast.Module(body=wanted, type_ignores=[]), str(path), "exec"), namespace)
    return namespace


class TorchOnCpu:
    """`torch` as seen by reference code that asks for device="cuda" in its draws (generate.py:166,171,...): the same
    module, with the device keyword of the factory functions dropped so the unmodified text runs on the CPU."""

    def __getattr__(self, k):
        v = getattr(torch, k)
        if k in ("randn", "rand", "zeros", "ones", "empty", "tensor", "arange", "linspace", "randint"):
            def f(*a, **kw):
                kw.pop("device", None)
                return v(*a, **kw)
            return f
        return v
