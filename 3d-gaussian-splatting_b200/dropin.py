"""Launcher glue that lets the reference's OWN code (src/train.py, src/eval.py, GaussianTrainer,
GaussianRasterizer) run unmodified on top of B200Backend.

The reference picks its backend by a string literal (``backend="gsplat"`` at src/training/trainer.py:706-708
and src/eval.py:130) resolved in ``GaussianRasterizer._initialize_backend``
(src/gaussians/gaussian_rasterizer.py:97-118), and imports ``gsplat`` / ``plyfile`` / ``matplotlib`` at
module top (src/backends/gsplat_backend.py:4, src/training/export.py:3, src/training/trainer.py:22).
``register()``:

* puts the reference tree and this repo on ``sys.path`` (both are PEP-420 namespace trees under ``src``),
* installs a ``gsplat`` module whose ``rasterization`` is this package's (so even the reference's own
  ``GSplatBackend`` would reach the sm_100a kernels), and stand-ins for ``plyfile`` / ``matplotlib`` /
  ``kornia`` ONLY when the real modules are absent from the image (probed first, never installed),
* patches the one registration point so ``backend in {"auto", "gsplat", "b200"}`` yields ``B200Backend``.

Nothing here touches the arithmetic; there is no CPU fallback (``B200Backend`` raises on CPU tensors).
"""
from __future__ import annotations

import contextlib
import importlib
import importlib.util
import os
import sys
import types
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)


def reference_root(explicit: Optional[str] = None) -> Optional[str]:
    """Where the unmodified reference tree lives: ``explicit`` / $B200_REFERENCE_ROOT, the build container's
    /root/reference, or the byte-identical copy staged by baseline/stage_reference.py (verified against the
    tracked sha256 manifest — a modified tree is refused)."""
    for cand in (explicit, os.environ.get("B200_REFERENCE_ROOT"), "/root/reference"):
        if cand and os.path.isfile(os.path.join(cand, "src", "training", "trainer.py")):
            return cand
    staged = os.path.join(_ROOT, "baseline", "_ref")
    if os.path.isfile(os.path.join(staged, "src", "training", "trainer.py")):
        spec = importlib.util.spec_from_file_location("_b200_stage_reference",
                                                      os.path.join(_ROOT, "baseline", "stage_reference.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        if not mod.verify(staged):
            raise RuntimeError(f"{staged} does not match baseline/REFERENCE_MANIFEST.json (modified reference?)")
        return staged
    return None


def _absent(name: str) -> bool:
    if name in sys.modules:
        return False
    try:
        return importlib.util.find_spec(name) is None
    except (ImportError, ValueError):
        return True


def _plyfile_shim() -> types.ModuleType:
    """``PlyElement.describe(structured, "vertex")`` + ``PlyData([el], text=False).write(path)`` as
    src/training/export.py:108-121 uses them: binary little-endian PLY, one property per dtype field."""
    import numpy as np

    m = types.ModuleType("plyfile")
    _names = {"f4": "float", "f8": "double", "i4": "int", "u1": "uchar", "i2": "short", "u2": "ushort", "u4": "uint",
              "i1": "char"}

    class PlyElement:
        def __init__(self, name, data):
            self.name, self.data = name, data

        @staticmethod
        def describe(data, name, **_):
            return PlyElement(name, np.asarray(data))

    class PlyData:
        def __init__(self, elements, text=False, byte_order="<", **_):
            self.elements, self.text = list(elements), text

        def write(self, stream):
            own = isinstance(stream, (str, os.PathLike))
            f = open(stream, "wb") if own else stream
            try:
                head = ["ply", "format binary_little_endian 1.0"]
                for el in self.elements:
                    head.append(f"element {el.name} {len(el.data)}")
                    for fname in el.data.dtype.names:
                        dt = el.data.dtype[fname]
                        head.append(f"property {_names[dt.str.lstrip('<|=')]} {fname}")
                head.append("end_header")
                f.write(("\n".join(head) + "\n").encode("ascii"))
                for el in self.elements:
                    f.write(el.data.astype(el.data.dtype.newbyteorder("<")).tobytes())
            finally:
                if own:
                    f.close()

    m.PlyElement, m.PlyData = PlyElement, PlyData
    m.__version__ = "b200-shim"
    return m


def _matplotlib_shim():
    """``plt.subplots`` / ``savefig`` / ``close`` as src/training/trainer.py:1003-1033 calls them (no figure is
    drawn: the box has no matplotlib; training_metrics.png is a convenience plot, not a result)."""

    class _Anything:
        def __getattr__(self, _):
            return lambda *a, **k: None

    mpl = types.ModuleType("matplotlib")
    plt = types.ModuleType("matplotlib.pyplot")

    def subplots(nrows=1, ncols=1, **_):
        n = nrows * ncols
        return _Anything(), (_Anything() if n == 1 else [_Anything() for _ in range(n)])

    plt.subplots = subplots
    for fn in ("tight_layout", "savefig", "close", "figure", "plot", "show"):
        setattr(plt, fn, lambda *a, **k: None)
    mpl.pyplot = plt
    mpl.use = lambda *a, **k: None
    mpl.__version__ = "b200-shim"
    return mpl, plt


def install_shims() -> dict:
    """Stand-ins for modules the reference imports at module top but the image lacks.  Returns what was done."""
    from .rasterization import rasterization

    done = {}
    if _absent("gsplat"):  # the arithmetic provider the reference calls (gsplat_backend.py:4,83,168)
        g = types.ModuleType("gsplat")
        g.rasterization = rasterization
        g.__version__ = "b200-shim"
        sys.modules["gsplat"] = g
        done["gsplat"] = "shim -> 3d-gaussian-splatting_b200.rasterization"
    else:
        done["gsplat"] = "present" if "gsplat" not in sys.modules or getattr(
            sys.modules["gsplat"], "__version__", "") != "b200-shim" else "shim (already installed)"
    if _absent("plyfile"):
        sys.modules["plyfile"] = _plyfile_shim()
        done["plyfile"] = "shim (binary little-endian writer)"
    if _absent("matplotlib"):
        mpl, plt = _matplotlib_shim()
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
        done["matplotlib"] = "shim (no-op plots)"
    return done


def register(reference: Optional[str] = None):
    """Make the reference's ``GaussianRasterizer`` hand out ``B200Backend``; returns the reference's
    ``src.gaussians.gaussian_rasterizer`` module.  Idempotent."""
    from .backend import B200Backend

    root = reference_root(reference)
    if root is None:
        raise RuntimeError("reference tree not found (set B200_REFERENCE_ROOT or run baseline/stage_reference.py)")
    for p in (root, _ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    install_shims()
    gr = importlib.import_module("src.gaussians.gaussian_rasterizer")

    def _initialize_backend(self, backend, K):
        if backend in ("auto", "gsplat", "b200"):
            return B200Backend(K)
        raise RuntimeError("No rasterization backend available!")

    gr.GaussianRasterizer._initialize_backend = _initialize_backend
    return gr


# ------------------------------------------------------------------------------------------------------------------
# Driving the reference's own trainer objects with this backend (used by bench.py's `reference_trainer` leg and by the
# drop-in tests; no oracle, no CPU path involved)
# ------------------------------------------------------------------------------------------------------------------
REFERENCE_PARAM_KEYS = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")


def reference_available() -> bool:
    try:
        return reference_root() is not None
    except RuntimeError:
        return False


def reference_modules():
    """(rasterizer module, trainer module, model module, config module) of the UNMODIFIED reference, with B200Backend
    registered at gaussian_rasterizer.py:97-118."""
    gr = register()
    tr = importlib.import_module("src.training.trainer")
    gm = importlib.import_module("src.gaussians.gaussian_model")
    cfg = importlib.import_module("src.gaussians.training_config")
    return gr, tr, gm, cfg


@contextlib.contextmanager
def backend_override(gr, factory):
    """Temporarily make GaussianRasterizer hand out ``factory(K)`` (None = leave the registered B200Backend)."""
    if factory is None:
        yield
        return
    saved = gr.GaussianRasterizer._initialize_backend
    gr.GaussianRasterizer._initialize_backend = lambda self, backend, K: factory(K)
    try:
        yield
    finally:
        gr.GaussianRasterizer._initialize_backend = saved


def make_reference_trainer(start, K, device: str, backend_factory=None, mixed_precision=None, **config_overrides):
    """The reference's GaussianTrainer / GaussianModel / Adam (``_setup_optimizer``, trainer.py:331-340) /
    GaussianRasterizer(..., backend="gsplat") (trainer.py:706-708), all unmodified, holding the raw parameters ``start``
    (dict over REFERENCE_PARAM_KEYS).  ``backend_factory``: K -> backend object, for a comparator arm (tests)."""
    import torch
    gr, tr, gm, cfg = reference_modules()
    config = cfg.TrainingConfig()
    config.wandb_mode = "disabled"
    if mixed_precision is not None:
        config.use_mixed_precision = bool(mixed_precision)
    for k, v in config_overrides.items():
        if not hasattr(config, k):
            raise AttributeError(f"TrainingConfig has no field {k}")
        setattr(config, k, v)
    trainer = tr.GaussianTrainer(config, device=device, wandb_run=None)
    n = start["xyz"].shape[0]
    model = gm.GaussianModel(n_gaussians=n, device=device)
    with torch.no_grad():
        for k in REFERENCE_PARAM_KEYS:
            getattr(model, k).data = start[k].to(device=device, dtype=torch.float32).clone()
    optimizer = trainer._setup_optimizer(model)
    with backend_override(gr, backend_factory):
        rasterizer = gr.GaussianRasterizer(K=K.to(device), device=device, enable_caching=True, backend="gsplat")
    return trainer, model, optimizer, rasterizer


def tensor_views(cams, images, K, W: int, H: int, device: str):
    """What ``_to_tensor_view`` (trainer.py:270-281) builds per frame."""
    import torch
    return [{"image": img.to(device=device, dtype=torch.float32), "pose": vm.to(device=device, dtype=torch.float32),
             "K": K.to(device=device, dtype=torch.float32), "width": W, "height": H} for vm, img in zip(cams, images)]


def main(argv=None) -> int:
    """``python -m src.backends.b200_launch train|eval <reference CLI arguments…>`` — runs the reference's
    unmodified ``src.train.main()`` / ``src.eval.main()`` with the B200 backend registered."""
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] not in ("train", "eval"):
        print("usage: b200_launch {train|eval} [reference CLI arguments]", file=sys.stderr)
        return 2
    register()
    mod = importlib.import_module("src.train" if argv[0] == "train" else "src.eval")
    sys.argv = [mod.__file__] + argv[1:]
    mod.main()
    return 0
