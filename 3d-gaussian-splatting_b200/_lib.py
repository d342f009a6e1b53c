"""ctypes binding of libb200splat.so (C ABI: include/b200splat.h).

There is no CPU fallback (BASELINE.json north_star): if the CUDA library is missing this module raises
at import, and every entry point raises ``RuntimeError`` on a non-zero return code, mirroring the
reference's convention of Python exceptions at the backend boundary
(src/gaussians/gaussian_rasterizer.py:115-118).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_LIB") or os.path.join(_HERE, "libb200splat.so")  # B200_LIB: a tuning build of the same ABI

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        f"or `make -C {os.path.join(_HERE, 'csrc')}` (there is no CPU fallback)"
    )

lib = C.CDLL(LIB_PATH)

_p = C.c_void_p
_i = C.c_int
_l = C.c_int64
_z = C.c_size_t

SIGNATURES = {
    "b200_version": (C.c_int, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_project_blocks": (_l, [_l]),
    "b200_cam_pack": (_i, [_p, _p, C.c_float, C.c_float, C.c_float, C.c_float, _i, _i, _p, _p]),
    "b200_project_sh_fwd": (_i, [_p, _p, _p, _p, _p, _i, _i, _l, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "b200_tile_count": (_i, [_p, _p, _l, _i, _i, _p, _p]),
    "b200_tile_prep_max_cells": (_z, []),
    "b200_tile_prep": (_i, [_p, _i, _i, _p, _p, _p, _p]),
    "b200_sort_pairs_k16_prepared": (_i, [_p, _p, _p, _p, _l, _i, _p, _p, _z, _p, _p]),
    "b200_sort_pairs_k32": (_i, [_p, _p, _p, _p, _l, _i, _p, _z, _p, _p]),
    "b200_sort_pairs_k32_static": (_i, [_p, _p, _p, _p, _l, _i, _p, _p]),
    "b200_sort_pairs_k16": (_i, [_p, _p, _p, _p, _l, _i, _p, _z, _p, _p]),
    "b200_bin_count_sorted": (_i, [_p, _p, _l, _p, _p]),
    "b200_bin_emit_sorted": (_i, [_p, _p, _p, _p, _p, _l, _i, _i, _l, _p, _p, _p]),
    "b200_tile_ranges_k16": (_i, [_p, _l, _i, _p, _p]),
    "b200_isect_ids": (_i, [_p, _p, _p, _l, _p, _p]),
    "b200_bin_coarse_supported": (_i, [_i, _i]),
    "b200_bin_coarse_scratch_bytes": (_z, [_l, _l]),
    "b200_bin_coarse": (_i, [_p, _p, _p, _p, _l, _i, _i, _p, _l, _p, _p, _p, _p, _p, _p, _z, _p]),
    "b200_isect_ids_from_offsets": (_i, [_p, _p, _i, _p, _l, _p, _p]),
    "b200_bin_scan": (_i, [_p, _l, _p, _p, _p]),
    "b200_bin_emit": (_i, [_p, _p, _p, _p, _l, _i, _i, _l, _p, _p, _p, _p]),
    "b200_sort_scratch_bytes": (_z, [_l]),
    "b200_sort_pairs": (_i, [_p, _p, _p, _p, _l, _i, _p, _z, _p, _p]),
    "b200_tile_ranges": (_i, [_p, _l, _i, _p, _p]),
    "b200_raster_fwd": (_i, [_p, _p, _p, _l, _i, _i, _i, _p, _i, _p, _p, _p, _p]),
    "b200_raster_bwd": (_i, [_p, _p, _p, _l, _i, _i, _i, _p, _i, _p, _p, _p, _p, _p, _p]),
    "b200_project_sh_bwd": (_i, [_p, _p, _p, _p, _p, _i, _i, _l, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p]),
    "b200_project_sh_fwd_raw": (_i, [_p, _p, _p, _p, _p, _i, _i, _l, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "b200_project_sh_bwd_raw": (_i, [_p, _p, _p, _p, _p, _i, _i, _l, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _i, _p]),
    "b200_l1_ssim_fwd_bwd": (_i, [_p, _p, _i, _i, C.c_float, C.c_float, _p, _p, _p]),
    "b200_adam_step": (_i, [_p, _p, _p, _p, _l, _p, _i, _i, C.c_float, C.c_float, C.c_float, _p]),
    "b200_adam_step_stats": (_i, [_p, _p, _p, _p, _l, _p, _i, _i, C.c_float, C.c_float, C.c_float, _p, _l, _p, _p, _p]),
    "b200_store_gather": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _l, _p]),
    "b200_densify_stats": (_i, [_p, _l, _p, _p, _p]),
    "b200_allreduce_multimem": (_i, [_p, _l, _i, _i, C.c_float, _i, _p]),
    "b200_multimem_slice": (None, [_l, _i, _i, _p, _p]),
    "b200_allreduce_adam_multimem": (_i, [_p, _p, _p, _p, _p, _l, _i, _i, C.c_float, _p, _i, _i, C.c_float, C.c_float, C.c_float,
                                          _l, _l, _i, _p]),
    "b200_share_slice_multimem": (_i, [_p, _p, _l, _i, _i, _p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here == header/library mismatch: fail loudly
    _fn.restype = _res
    _fn.argtypes = _args

CH_RGB, CH_D, CH_RGBD = 0, 1, 2
SPLAT_FLOATS = 12
TILE = 16
CAM_FLOATS = 24
CAM_DEV_FLOATS = 32


# Instrumentation used by bench.py: number of kernels of THIS library launched so far, and an optional
# per-stage CUDA-event recorder (object with .begin(name)/.end(name)) on the launching stream.
launch_count = 0
stage_timer = None


def check(rc: int, what: str, kernels: int = 1) -> None:
    global launch_count
    launch_count += kernels
    if rc != 0:
        msg = lib.b200_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")
