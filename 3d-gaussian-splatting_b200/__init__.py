"""B200-native (sm_100a) Gaussian-splatting rasterizer — drop-in for the one hot path of
ShayManor/3d-gaussian-splatting (``backend.render_with_depth`` -> ``gsplat.rasterization``).

The directory name is not a Python identifier; load it with
``importlib.import_module("3d-gaussian-splatting_b200")`` (repo root on ``sys.path``) or through the
reference-namespace shim ``src/backends/b200_backend.py``.
"""
from . import _lib  # noqa: F401  (raises ImportError if the CUDA library is not built)
from .rasterization import rasterization, project_and_bin, RasterizeSplats, make_cam, set_gradient_arena  # noqa: F401
from .backend import B200Backend  # noqa: F401
from .stats import densify_stats  # noqa: F401
from .store import GaussianStore  # noqa: F401
from .loss import l1_ssim_loss  # noqa: F401
from .optim import StoreAdam  # noqa: F401
from .streams import ViewStreams  # noqa: F401
from . import dropin  # noqa: F401
from .distributed import FlatGradients, GradientArena, SymmetricBuffer, balance_views, shard_views, allreduce_gradients, allreduce_flat, ViewShardedStep  # noqa: F401

__all__ = [
    "rasterization", "project_and_bin", "RasterizeSplats", "make_cam", "B200Backend", "densify_stats", "GaussianStore", "l1_ssim_loss", "StoreAdam",
    "ViewStreams", "FlatGradients", "GradientArena", "SymmetricBuffer", "set_gradient_arena", "balance_views", "shard_views", "allreduce_gradients", "allreduce_flat", "ViewShardedStep",
]
