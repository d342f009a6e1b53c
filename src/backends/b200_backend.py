"""Reference-namespace shim: ``from src.backends.b200_backend import B200Backend``.

The reference tree has no ``__init__.py`` files, so ``src`` / ``src.backends`` are PEP-420 namespace
packages; with this repo's root AND the reference's root on ``sys.path`` this module sits next to
``src/backends/gsplat_backend.py`` and is selected by ``register()`` below, which patches the one
registration point the reference has (src/gaussians/gaussian_rasterizer.py:97-118) so that
``backend="gsplat"`` (hard-coded at src/training/trainer.py:706-708 and src/eval.py:130) resolves to the
B200 backend.  Do not add ``__init__.py`` under this repo's ``src/``.

``python -m src.backends.b200_launch train|eval …`` runs the reference's CLI entry points on top of it.
"""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

_pkg = importlib.import_module("3d-gaussian-splatting_b200")
B200Backend = _pkg.B200Backend
rasterization = _pkg.rasterization
register = _pkg.dropin.register  # shims for absent modules + the registration patch (dropin.py)
