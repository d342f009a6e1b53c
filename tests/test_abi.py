"""CPU: the C-ABI library loads and exports every symbol include/b200splat.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200splat.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_stage_set():
    syms = _declared_symbols()
    for s in ("b200_project_sh_fwd", "b200_bin_scan", "b200_bin_emit", "b200_sort_pairs", "b200_tile_ranges",
              "b200_raster_fwd", "b200_raster_bwd", "b200_project_sh_bwd", "b200_densify_stats", "b200_version",
              "b200_last_error"):
        assert s in syms


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg._lib.LIB_PATH)
    for s in _declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/b200splat.h but not exported"
    assert set(pkg._lib.SIGNATURES) == set(_declared_symbols())


def test_version_and_scratch_query_need_no_gpu(pkg):
    lib = pkg._lib.lib
    assert lib.b200_version() >= 100
    assert lib.b200_sort_scratch_bytes(0) > 0
    assert lib.b200_sort_scratch_bytes(10_000_000) > lib.b200_sort_scratch_bytes(1_000_000)
    assert lib.b200_project_blocks(257) == 2


def test_bad_arguments_are_rejected_without_touching_the_gpu(pkg):
    lib = pkg._lib.lib
    rc = lib.b200_sort_pairs(None, None, None, None, 10, 45, None, 0, None, None)
    assert rc == -1 and b"null" in lib.b200_last_error()
    rc = lib.b200_raster_fwd(None, None, None, 0, 0, 0, 0, None, 0, None, None, None, None)
    assert rc == -1


def test_cpu_tensors_raise(pkg):
    import pytest
    import torch
    K = torch.eye(3)
    with pytest.raises(RuntimeError):
        pkg.B200Backend(K)
    with pytest.raises(RuntimeError):
        pkg.rasterization(torch.zeros(1, 3), torch.zeros(1, 4), torch.zeros(1, 3), torch.zeros(1),
                          torch.zeros(1, 16, 3), torch.eye(4)[None], torch.eye(3)[None], 16, 16, sh_degree=3)
