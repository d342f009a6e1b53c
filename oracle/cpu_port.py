"""ctypes wrapper of oracle/cpu_splat.c (TEST INFRASTRUCTURE ONLY — see that file's header).

Used by tests (checker at sizes the PyTorch oracle cannot reach) and by bench.py's ``cpu_baseline`` /
``--impl reference`` legs (the reference has no CPU implementation of this path to time:
src/backends/pytorch_backend.py:1-5).  Parity unpinned, like torch_oracle.py.  Never imported by the
product package.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess
from typing import Dict, Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_cpu", "libcpusplat.so")


def _load():
    if not os.path.exists(_SO):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    lib = C.CDLL(_SO)
    lib.cpu_num_threads.restype = C.c_int
    lib.cpu_cumsum.restype = C.c_int64
    lib.cpu_pairs_evaluated.restype = C.c_longlong
    lib.cpu_pairs_blended.restype = C.c_longlong
    return lib


_lib = _load()
_p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731


def num_threads() -> int:
    return int(_lib.cpu_num_threads())


def pairs_evaluated():
    """(pairs visited, pairs composited) of the last forward blend — P of SURVEY.md §8(d)."""
    return int(_lib.cpu_pairs_evaluated()), int(_lib.cpu_pairs_blended())


def set_threads(n: int) -> None:
    """Use ``n`` OpenMP threads from now on (overrides an inherited OMP_NUM_THREADS, e.g. torchrun's 1)."""
    _lib.cpu_set_threads(int(n))


def cam24(viewmat: torch.Tensor, K: torch.Tensor, near=0.01, far=1e10, eps2d=0.3, radius_clip=0.0) -> torch.Tensor:
    return torch.cat([viewmat.float().reshape(16), torch.tensor(
        [K[0, 0], K[1, 1], K[0, 2], K[1, 2], near, far, eps2d, radius_clip], dtype=torch.float32)]).contiguous()


def _f(t):
    return t.detach().to(torch.float32).contiguous()


def forward(act: Dict[str, torch.Tensor], viewmat, K, W: int, H: int, deg: int, bg: Optional[torch.Tensor] = None,
            mode: str = "RGB") -> Dict[str, torch.Tensor]:
    """Whole forward pipeline == gsplat.rasterization as gsplat_backend.py:168-182 calls it."""
    means, quats, scales = _f(act["means3D"]), _f(act["rotations"]), _f(act["scales"])
    opac, shs = _f(act["opacities"]).reshape(-1), _f(act["shs"])
    N = means.shape[0]
    cam = cam24(viewmat, K)
    radii = torch.empty(N, 2, dtype=torch.int32)
    means2d, depths = torch.empty(N, 2), torch.empty(N)
    conics, colors = torch.empty(N, 3), torch.empty(N, 3)
    tpg = torch.empty(N, dtype=torch.int32)
    _lib.cpu_project_sh_fwd(_p(means), _p(quats), _p(scales), _p(opac), _p(shs), C.c_int(shs.shape[1]), C.c_int(deg),
                            C.c_int64(N), _p(cam), C.c_int(W), C.c_int(H), _p(radii), _p(means2d), _p(depths),
                            _p(conics), _p(colors), _p(tpg))
    cum = torch.empty(N, dtype=torch.int64)
    M = int(_lib.cpu_cumsum(_p(tpg), C.c_int64(N), _p(cum)))
    ids = torch.empty(max(M, 1), dtype=torch.int64)
    flat = torch.empty(max(M, 1), dtype=torch.int32)
    _lib.cpu_bin_emit(_p(means2d), _p(radii), _p(depths), _p(cum), C.c_int64(N), C.c_int(W), C.c_int(H), _p(ids), _p(flat))
    tile_w, tile_h = (W + 15) // 16, (H + 15) // 16
    end_bit = 32 + int(math.floor(math.log2(tile_w * tile_h))) + 1
    tk, tv = torch.empty_like(ids), torch.empty_like(flat)
    _lib.cpu_sort_pairs(_p(ids), _p(flat), _p(tk), _p(tv), C.c_int64(M), C.c_int(end_bit))
    offsets = torch.empty(tile_h * tile_w, dtype=torch.int32)
    _lib.cpu_tile_ranges(_p(ids), C.c_int64(M), C.c_int(tile_h * tile_w), _p(offsets))
    if mode == "RGB":
        feat, CH = colors, 3
        bgv = _f(bg) if bg is not None else None
    elif mode in ("D", "ED"):
        feat, CH = depths[:, None].contiguous(), 1
        bgv = torch.zeros(1) if bg is not None else None
    else:
        feat, CH = torch.cat([colors, depths[:, None]], 1).contiguous(), 4
        bgv = torch.cat([_f(bg), torch.zeros(1)]) if bg is not None else None
    render = torch.empty(H, W, CH)
    alpha = torch.empty(H, W, 1)
    last = torch.empty(H, W, dtype=torch.int32)
    _lib.cpu_raster_fwd(_p(means2d), _p(conics), _p(feat), C.c_int(CH), _p(opac), _p(flat), _p(offsets), C.c_int64(M),
                        C.c_int(W), C.c_int(H), _p(bgv), _p(render), _p(alpha), _p(last))
    return dict(render=render, alpha=alpha, last_ids=last, radii=radii, means2d=means2d, depths=depths, conics=conics,
                colors=colors, tiles_per_gauss=tpg, cum_tiles=cum, isect_ids=ids[:M], flatten_ids=flat[:M],
                isect_offsets=offsets.view(tile_h, tile_w), M=M, feat=feat, CH=CH, bg=bgv, cam=cam,
                inputs=(means, quats, scales, opac, shs))


def backward(fw: Dict, W: int, H: int, deg: int, v_render: torch.Tensor, v_alpha: Optional[torch.Tensor] = None):
    """Gradients w.r.t. the five activated inputs for the raw (un-normalised) render of `forward`
    (mode RGB / D / RGB+D; the ED division is left to the caller)."""
    means, quats, scales, opac, shs = fw["inputs"]
    N, CH, M = means.shape[0], fw["CH"], fw["M"]
    v_render = _f(v_render)
    v_alpha = _f(v_alpha) if v_alpha is not None else None
    v_m2d, v_con = torch.zeros(N, 2), torch.zeros(N, 3)
    v_feat, v_op = torch.zeros(N, CH), torch.zeros(N)
    _lib.cpu_raster_bwd(_p(fw["means2d"]), _p(fw["conics"]), _p(fw["feat"]), C.c_int(CH), _p(opac),
                        _p(fw["flatten_ids"] if M else torch.zeros(1, dtype=torch.int32)), _p(fw["isect_offsets"]),
                        C.c_int64(M), C.c_int(W), C.c_int(H), _p(fw["bg"]), _p(fw["alpha"]), _p(fw["last_ids"]),
                        _p(v_render), _p(v_alpha), _p(v_m2d), _p(v_con), _p(v_feat), _p(v_op))
    if CH == 3:
        v_col, v_dep = v_feat, None
    elif CH == 1:
        v_col, v_dep = torch.zeros(N, 3), v_feat[:, 0].contiguous()
    else:
        v_col, v_dep = v_feat[:, :3].contiguous(), v_feat[:, 3].contiguous()
    v_means, v_quats = torch.empty(N, 3), torch.empty(N, 4)
    v_scales, v_shs = torch.empty(N, 3), torch.empty_like(shs)
    _lib.cpu_project_sh_bwd(_p(means), _p(quats), _p(scales), _p(shs), C.c_int(shs.shape[1]), C.c_int(deg), C.c_int64(N),
                            _p(fw["cam"]), C.c_int(W), C.c_int(H), _p(fw["radii"]), _p(v_m2d), _p(v_con), _p(v_col),
                            _p(v_dep), _p(v_means), _p(v_quats), _p(v_scales), _p(v_shs))
    return dict(means3D=v_means, rotations=v_quats, scales=v_scales, opacities=v_op, shs=v_shs,
                v_means2d=v_m2d, v_conics=v_con, v_colors=v_col)


def forward_backward(act, viewmat, K, W, H, deg, v_render, bg=None):
    fw = forward(act, viewmat, K, W, H, deg, bg=bg)
    grads = backward(fw, W, H, deg, v_render)
    return dict(M=fw["M"], render=fw["render"], alpha=fw["alpha"], grads=grads, fw=fw)


def densify_stats(accum: torch.Tensor, count: torch.Tensor, grad: torch.Tensor) -> None:
    _lib.cpu_densify_stats(_p(_f(grad)), C.c_int64(grad.shape[0]), _p(accum), _p(count))
