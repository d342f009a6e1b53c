"""Host side of the rasterizer: stage orchestration over the C ABI and the autograd bridge.

``rasterization(...)`` mirrors the call the reference makes into gsplat 1.5.3
(src/backends/gsplat_backend.py:83-97 and :168-182): same argument names, same return triple
``(renders[C,H,W,CH], alphas[C,H,W,1], meta)`` and the same ``meta`` keys the reference reads
(``means2d``, ``radii`` — gsplat_backend.py:204-208) plus gsplat's binning keys used by the parity
tests (``tiles_per_gauss, isect_ids, flatten_ids, isect_offsets``).

PyTorch is plumbing only here (device memory, current stream, autograd graph); every stage is a
hand-written sm_100a kernel reached through ``libb200splat.so``.  No CPU path exists.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
import os
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import lib, check

_MODES = {"RGB": _lib.CH_RGB, "D": _lib.CH_D, "ED": _lib.CH_D, "RGB+D": _lib.CH_RGBD, "RGB+ED": _lib.CH_RGBD}
_NCH = {_lib.CH_RGB: 3, _lib.CH_D: 1, _lib.CH_RGBD: 4}


class _stage:
    """Brackets one stage with CUDA events when bench.py installed a recorder in ``_lib.stage_timer``."""

    __slots__ = ("name",)

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if _lib.stage_timer is not None:
            _lib.stage_timer.begin(self.name)

    def __exit__(self, *exc):
        if _lib.stage_timer is not None:
            _lib.stage_timer.end(self.name)
        return False


class _stage_on:
    """``_stage(name)`` for a stage launched on ``stream`` by handle: only while a recorder is installed (``active``) is
    that stream made current around the block, so the recorder's events land on the stream the kernels run on."""

    __slots__ = ("stream", "active", "name", "prev")

    def __init__(self, stream, active: bool, name: str):
        self.stream, self.active, self.name = stream, active, name

    def __enter__(self):
        if self.active:
            self.prev = torch.cuda.current_stream(self.stream.device)
            torch.cuda.set_stream(self.stream)
            _lib.stage_timer.begin(self.name)

    def __exit__(self, *exc):
        if self.active:
            _lib.stage_timer.end(self.name)
            torch.cuda.set_stream(self.prev)
        return False


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


_NO_GUARD = contextlib.nullcontext()


def _on_device(dev: torch.device):
    """``torch.cuda.device(dev)`` that costs nothing when ``dev`` is already current (the usual case)."""
    return _NO_GUARD if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)


def _require_cuda(*tensors: torch.Tensor) -> torch.device:
    dev = tensors[0].device
    for t in tensors:
        if t.device.type != "cuda":
            raise RuntimeError(
                "b200 rasterizer: all inputs must be CUDA tensors (there is no CPU fallback); got "
                f"{t.device}"
            )
        if t.device != dev:
            raise RuntimeError("b200 rasterizer: inputs live on different devices")
    return dev


def _plain_f32(t: torch.Tensor, dev: torch.device) -> bool:
    """Usable by address as it is (the common case: skips three no-op torch calls per tensor on the host path)."""
    return t.dtype == torch.float32 and t.device == dev and t.is_contiguous()


def make_cam(viewmat: torch.Tensor, K: torch.Tensor, width: int, height: int, near_plane: float = 0.01,
             far_plane: float = 1e10, eps2d: float = 0.3, radius_clip: float = 0.0, stream: Optional[int] = None) -> torch.Tensor:
    """Packed DEVICE camera block (include/b200splat.h, b200_cam_pack).  The pose and the intrinsics are
    consumed on the device — like gsplat, which takes ``viewmats``/``Ks`` as CUDA tensors
    (gsplat_backend.py:174-175) — so rendering a view never synchronises on them.  CPU tensors are uploaded."""
    dev = torch.device("cuda", torch.cuda.current_device())
    if viewmat.device.type == "cuda":
        dev = viewmat.device
    elif K.device.type == "cuda":
        dev = K.device
    vm = viewmat if _plain_f32(viewmat, dev) else viewmat.detach().to(device=dev, dtype=torch.float32).contiguous()
    Kd = K if _plain_f32(K, dev) else K.detach().to(device=dev, dtype=torch.float32).contiguous()
    if vm.numel() != 16 or Kd.numel() != 9:
        raise ValueError("make_cam: viewmat must be [4,4] and K [3,3]")
    cam = torch.empty(_lib.CAM_DEV_FLOATS, dtype=torch.float32, device=dev)
    with _on_device(dev):
        check(lib.b200_cam_pack(_ptr(vm), _ptr(Kd), float(near_plane), float(far_plane), float(eps2d),
                                float(radius_clip), int(width), int(height), _ptr(cam), _stream() if stream is None else stream),
              "b200_cam_pack")
    return cam


_PINNED_M = {}


def _pinned_scalar(dev: torch.device) -> torch.Tensor:
    """One pinned int64 per device for the asynchronous read-back of the intersection count."""
    t = _PINNED_M.get(dev)
    if t is None:
        t = torch.empty(1, dtype=torch.int64).pin_memory()
        _PINNED_M[dev] = t
    return t


_SIDE_STREAMS = {}
# True: the tile-corner grid is built by b200_tile_count on the side stream (under the depth sort) instead of by four
# scattered REDs per Gaussian inside the projection kernel.  B200_GRID_IN_PROJECTION=1 restores the round-1 placement.
GRID_ON_SIDE_STREAM = os.environ.get("B200_GRID_IN_PROJECTION", "0") != "1"


def _side_stream(dev: torch.device) -> torch.cuda.Stream:
    """One helper stream per device for the single-CTA tile_prep kernel + the read-back of M: both depend only on the
    projection and run UNDER the depth sort of the launching stream (forked and joined with events)."""
    s = _SIDE_STREAMS.get(dev)
    if s is None:
        s = torch.cuda.Stream(device=dev, priority=-1 if FRONT_PRIORITY else 0)
        _SIDE_STREAMS[dev] = s
    return s


def tile_n_bits(tile_w: int, tile_h: int) -> int:
    """gsplat's key width: floor(log2(n_tiles)) + 1 — one more bit than the largest tile id needs when
    n_tiles is a power of two (65 536 tiles -> 17)."""
    return int(math.floor(math.log2(tile_w * tile_h))) + 1


def _tile_id_bits(n_tiles: int) -> int:
    """Bits that actually carry tile ids 0 .. n_tiles-1 (what the 16-bit split key has to sort)."""
    return max(1, (n_tiles - 1).bit_length())


def _sort_scratch(M: int, dev) -> torch.Tensor:
    return torch.empty(max(int(lib.b200_sort_scratch_bytes(M)), 256), dtype=torch.uint8, device=dev)


_COARSE_OK = {}


def _coarse_supported(width: int, height: int) -> bool:
    """Two-level binning available for this image size (tile grid within b200_tile_prep / b200_bin_coarse limits)."""
    ok = _COARSE_OK.get((width, height))
    if ok is None:
        tile_w, tile_h = (width + _lib.TILE - 1) // _lib.TILE, (height + _lib.TILE - 1) // _lib.TILE
        ok = bool(lib.b200_bin_coarse_supported(width, height)) and \
            (tile_w + 1) * (tile_h + 1) <= int(lib.b200_tile_prep_max_cells()) and tile_w * tile_h <= 65536
        _COARSE_OK[(width, height)] = ok
    return ok


def _carve(dev: torch.device, sizes):
    """ONE allocation for buffers that die with the call: returns (owner tensor, device addresses), 256-B aligned."""
    offs, total = [], 0
    for b in sizes:
        offs.append(total)
        total += (max(int(b), 0) + 255) & ~255
    owner = torch.empty(max(total, 256), dtype=torch.uint8, device=dev)
    base = owner.data_ptr()
    return owner, [base + o for o in offs]


# FRONT_PRIORITY (default on; B200_FRONT_PRIORITY=0 turns it off): everything of a view up to the forward blend — camera
# block, projection, tile_count / tile_prep (the read-back of M), depth sort, two-level binning — is LAUNCHED on
# high-priority companions of the caller's stream.  The block scheduler hands the SM slots that free up to the oldest
# pending kernel first, so with the views of a step in flight on alternating streams (streams.ViewStreams) these short
# latency-bound kernels used to wait behind every not-yet-dispatched CTA of the other view's blend and ran in its drain
# phase only (and the host, which needs M to size the intersection buffers, waited with them).  With priority their few CTAs
# are placed as blend CTAs retire: a view's front end runs UNDER the previous view's backward blend.  Measured at config A
# (profiles/overlap_trace.py, r02 notes): 4-view step 4.65 -> 4.35 ms.  Memory stays in the caller's stream pool (only the
# launches move), ordered by events both ways; on one stream it changes nothing but two event records.
FRONT_PRIORITY = int(os.environ.get("B200_FRONT_PRIORITY", "1"))
_FRONT_STREAMS = {}


def _front_stream(dev: torch.device, cur: torch.cuda.Stream) -> torch.cuda.Stream:
    key = (dev.index, cur.cuda_stream)
    s = _FRONT_STREAMS.get(key)
    if s is None:
        s = torch.cuda.Stream(device=dev, priority=-1)
        _FRONT_STREAMS[key] = s
    return s


def _enter_front(dev: torch.device) -> torch.cuda.Stream:
    """The high-priority companion of the current stream, ordered after everything queued on the current stream so far."""
    cur = torch.cuda.current_stream(dev)
    front = _front_stream(dev, cur)
    enter = torch.cuda.Event()
    enter.record(cur)
    front.wait_event(enter)
    return front


def _project_and_bin_coarse(means, quats, scales, opacities, colors, sh_coeffs: int, deg: int, cam, width: int, height: int,
                            tile_w: int, tile_h: int, N: int, raw: bool, want_grad_records: bool, coarse_gather: bool,
                            with_isect_ids: bool, dev: torch.device) -> Dict[str, torch.Tensor]:
    """The default path of project_and_bin (method "coarse"), written for a short HOST path: the tensors that outlive the call
    are allocated one by one, everything else lives in two carved workspaces addressed by raw pointers; the side-stream
    kernels are launched by stream handle (no stream context switch); M arrives in pinned host memory written by
    b200_tile_prep itself (unified addressing: the pinned scalar is device-visible), so there is no copy to enqueue."""
    n_tiles = tile_w * tile_h
    cur = torch.cuda.current_stream(dev)
    launch = _enter_front(dev) if FRONT_PRIORITY else cur
    st = launch.cuda_stream
    n1 = max(N, 1)
    i32, f32 = torch.int32, torch.float32
    radii = torch.empty(N, 2, dtype=i32, device=dev)
    splats = torch.empty(N, _lib.SPLAT_FLOATS, dtype=f32, device=dev)
    tiles_per_gauss = torch.empty(N, dtype=i32, device=dev)
    visibility = torch.empty(N, 2, dtype=torch.bool, device=dev)
    grad_records = torch.empty(N, _lib.SPLAT_FLOATS, dtype=f32, device=dev) if want_grad_records else None
    offsets = torch.empty(n_tiles, dtype=i32, device=dev)
    sort_bytes = max(int(lib.b200_sort_scratch_bytes(N)), 256)
    ws1, (p_dkeys_a, p_gids_a, p_dkeys_b, p_gids_b, p_rects, p_grid, p_digit, p_sort) = _carve(
        dev, (4 * n1, 4 * n1, 4 * n1, 4 * n1, 0 if coarse_gather else 8 * n1, 4 * (tile_h + 1) * (tile_w + 1), 4 * 512, sort_bytes))
    if coarse_gather:
        p_rects = None
    project_fwd = lib.b200_project_sh_fwd_raw if raw else lib.b200_project_sh_fwd  # raw: activations fused in
    # bench's per-stage events are recorded on the CURRENT stream: while a recorder is installed, each stage makes the stream
    # it launches on current (no allocation happens inside those blocks, so every buffer stays in the caller's pool)
    timed = _lib.stage_timer is not None
    with _stage_on(launch, timed, "project_sh_fwd"):
        check(project_fwd(_ptr(means), _ptr(quats), _ptr(scales), _ptr(opacities), _ptr(colors), sh_coeffs, deg, N, None,
                          _ptr(cam), width, height, _ptr(radii), _ptr(splats), _ptr(tiles_per_gauss), None, p_dkeys_a, p_gids_a,
                          None if GRID_ON_SIDE_STREAM else p_grid, _ptr(visibility), _ptr(grad_records), p_rects, st),
              "b200_project_sh_fwd", kernels=1 if N > 0 else 0)
    # tile_count + tile_prep (one CTA: 2-D prefix sum of the corner grid -> per-tile ranges, digit bases, M) depend only on the
    # projection: they go on a side stream and run under the depth sort.  The launching stream joins before it first needs
    # the ranges; the host collects M after the depth-sort launches are queued, so it never waits on an idle GPU.
    side = _side_stream(dev)
    fork = torch.cuda.Event()
    fork.record(launch)
    side.wait_event(fork)
    sst = side.cuda_stream
    m_host = _pinned_scalar(dev)
    if GRID_ON_SIDE_STREAM:  # the corner grid is counted here, not inside the projection kernel (b200_tile_count)
        with _stage_on(side, timed, "tile_count"):
            check(lib.b200_tile_count(_ptr(splats), _ptr(radii), N, width, height, p_grid, sst), "b200_tile_count",
                  kernels=1 if N > 0 else 0)
    with _stage_on(side, timed, "tile_prep"):
        check(lib.b200_tile_prep(p_grid, width, height, _ptr(offsets), m_host.data_ptr(), p_digit, sst), "b200_tile_prep")
    m_ready = torch.cuda.Event()
    m_ready.record(side)
    in_b = C.c_int(0)
    with _stage_on(launch, timed, "sort_depth"):
        check(lib.b200_sort_pairs_k32(p_dkeys_a, p_gids_a, p_dkeys_b, p_gids_b, N, 32, p_sort, sort_bytes, C.byref(in_b), st),
              "b200_sort_pairs_k32", kernels=6 if N > 0 else 0)
    p_perm = p_gids_b if in_b.value else p_gids_a
    launch.wait_event(m_ready)  # join: ranges / digit bases are used from here on
    m_ready.synchronize()  # the one host wait of the forward pass (sizes the intersection buffers)
    M = int(m_host.item())
    cap = max(M, 1)
    flat_ids = torch.empty(cap, dtype=i32, device=dev)
    sb = max(int(lib.b200_bin_coarse_scratch_bytes(N, M)), 256)
    ws2, (p_k0, p_k1, p_k2, p_k3, p_scr) = _carve(dev, (4 * cap, 4 * cap, 4 * cap, 4 * cap, sb))
    with _stage_on(launch, timed, "bin_coarse"):
        check(lib.b200_bin_coarse(_ptr(splats), _ptr(radii), p_rects, p_perm, N, width, height, _ptr(offsets), M,
                                  p_k0, p_k1, p_k2, p_k3, _ptr(flat_ids), p_scr, sb, st), "b200_bin_coarse",
              kernels=(3 + (2 if n_tiles > 16 * 256 else 1)) if (N > 0 and M > 0) else 0)
    ids = None
    if with_isect_ids:
        ids = torch.empty(cap, dtype=torch.int64, device=dev)
        check(lib.b200_isect_ids_from_offsets(_ptr(flat_ids), _ptr(offsets), n_tiles, _ptr(splats), M, _ptr(ids), st),
              "b200_isect_ids_from_offsets", kernels=1 if M > 0 else 0)
        ids = ids[:M]
    if launch is not cur:
        leave = torch.cuda.Event()
        leave.record(launch)
        cur.wait_event(leave)
    del ws1, ws2  # freed in the caller's stream order, after the join above
    return dict(radii=radii, splats=splats, tiles_per_gauss=tiles_per_gauss, tile_w=tile_w, tile_h=tile_h,
                visibility=visibility, grad_records=grad_records, M=M, flatten_ids=flat_ids[:M],
                isect_offsets=offsets.view(tile_h, tile_w), tile_keys=None, isect_ids=ids)


def project_and_bin(means, quats, scales, opacities, colors, sh_degree: Optional[int], cam, width: int,
                    height: int, keep_unsorted: bool = False, method: str = "auto",
                    with_isect_ids: bool = False, raw: bool = False, want_grad_records: bool = False) -> Dict[str, torch.Tensor]:
    """Stages 1-2 (projection+SH, tile binning, radix sort, tile ranges).  One host sync (reads M).

    method "keys64": literal gsplat formulation — emit (tile<<32|depth) int64 keys, sort M pairs.
    method "split" (default when tile_w*tile_h <= 65536): same result, depth digits sorted over the N
    Gaussians before duplication, 16-bit tile digits over the M intersections (include/b200splat.h); tile
    ranges and digit histograms come from the corner-difference grid (b200_tile_prep).
    method "split-hist": the split path with a histogram pass and tile_ranges over the sorted keys instead
    (what "split" falls back to when the tile grid does not fit b200_tile_prep).
    method "coarse" (the default whenever b200_bin_coarse_supported): depth sort over the Gaussians as in "split", then
    TWO-LEVEL binning — (Gaussian, 4x4-tile supertile) pairs are emitted in depth order, sorted on the supertile id, and
    expanded per supertile straight into the final per-tile lists (include/b200splat.h, b200_bin_coarse): the M
    intersections are written once (4 B each) and never sorted.  Same flatten_ids / isect_offsets, bit for bit.
    """
    dev = means.device
    N = means.shape[0]
    tile_w = (width + _lib.TILE - 1) // _lib.TILE
    tile_h = (height + _lib.TILE - 1) // _lib.TILE
    n_tiles = tile_w * tile_h
    coarse_ok = _coarse_supported(int(width), int(height))
    if method == "auto":
        method = "split" if (n_tiles <= 65536 and not keep_unsorted) else "keys64"
        if method == "split" and coarse_ok:
            method = "coarse"
    coarse_gather = method == "coarse-gather"  # "coarse" recomputing the rectangles from radii / means2d (test path)
    coarse = method in ("coarse", "coarse-gather")
    if sh_degree is None:
        sh_coeffs, deg = 1, -1
    else:
        sh_coeffs, deg = colors.shape[1], int(sh_degree)
    if coarse:
        if not coarse_ok:
            raise ValueError("binning='coarse' needs a tile grid within b200_bin_coarse_supported / b200_tile_prep limits")
        return _project_and_bin_coarse(means, quats, scales, opacities, colors, sh_coeffs, deg, cam, int(width), int(height),
                                       tile_w, tile_h, N, raw, want_grad_records, coarse_gather, with_isect_ids, dev)
    st = _stream()
    radii = torch.empty(N, 2, dtype=torch.int32, device=dev)
    splats = torch.empty(N, _lib.SPLAT_FLOATS, dtype=torch.float32, device=dev)
    tiles_per_gauss = torch.empty(N, dtype=torch.int32, device=dev)
    nblocks = int(lib.b200_project_blocks(N))
    block_sums = torch.empty(max(nblocks, 1), dtype=torch.int64, device=dev)
    block_offsets = torch.empty(max(nblocks, 1), dtype=torch.int64, device=dev)
    total = torch.empty(1, dtype=torch.int64, device=dev)  # written (not accumulated) by b200_bin_scan
    split = method in ("split", "split-hist")
    # tile ranges + radix digit bases from the corner-difference grid (b200_tile_prep) when the grid fits
    prep = method == "split" and (tile_w + 1) * (tile_h + 1) <= int(lib.b200_tile_prep_max_cells())
    tile_grid = torch.empty((tile_h + 1) * (tile_w + 1), dtype=torch.int32, device=dev) if prep else None
    if split:
        dkeys_a = torch.empty(max(N, 1), dtype=torch.int32, device=dev)
        gids_a = torch.empty(max(N, 1), dtype=torch.int32, device=dev)
    project_fwd = lib.b200_project_sh_fwd_raw if raw else lib.b200_project_sh_fwd  # raw: activations fused in
    # auxiliary outputs of the projection kernel: visibility_filter (= radii > 0) and, when a backward pass will follow,
    # the zero-initialised cotangent records it accumulates into — two eager torch launches per view otherwise
    visibility = torch.empty(N, 2, dtype=torch.bool, device=dev)
    grad_records = torch.empty(N, _lib.SPLAT_FLOATS, dtype=torch.float32, device=dev) if want_grad_records else None
    rects = None  # packed tile rectangles: the two-level path only (_project_and_bin_coarse)
    with _stage("project_sh_fwd"):
        check(project_fwd(_ptr(means), _ptr(quats), _ptr(scales), _ptr(opacities), _ptr(colors),
                                      sh_coeffs, deg, N, None, _ptr(cam), width, height, _ptr(radii), _ptr(splats),
                                      _ptr(tiles_per_gauss), None if split else _ptr(block_sums),  # split: recounted in depth order
                                      _ptr(dkeys_a) if split else None, _ptr(gids_a) if split else None,
                                      None if GRID_ON_SIDE_STREAM else _ptr(tile_grid), _ptr(visibility), _ptr(grad_records),
                                      _ptr(rects), st),
              "b200_project_sh_fwd", kernels=1 if N > 0 else 0)
    out = dict(radii=radii, splats=splats, tiles_per_gauss=tiles_per_gauss, tile_w=tile_w, tile_h=tile_h,
               visibility=visibility, grad_records=grad_records)
    offsets = torch.empty(n_tiles, dtype=torch.int32, device=dev)
    in_b = C.c_int(0)
    if not split:
        with _stage("bin_scan"):
            check(lib.b200_bin_scan(_ptr(block_sums), nblocks, _ptr(block_offsets), _ptr(total), st), "b200_bin_scan")
        M = int(total.item())  # the one device->host sync of the forward pass (sizes the key buffers)
        keys_a = torch.empty(max(M, 1), dtype=torch.int64, device=dev)
        vals_a = torch.empty(max(M, 1), dtype=torch.int32, device=dev)
        with _stage("bin_emit"):
            check(lib.b200_bin_emit(_ptr(splats), _ptr(radii), _ptr(tiles_per_gauss), _ptr(block_offsets), N, width,
                                    height, M, _ptr(keys_a), _ptr(vals_a), None, st), "b200_bin_emit",
                  kernels=1 if N > 0 else 0)
        if keep_unsorted:
            out["isect_ids_unsorted"] = keys_a[:M].clone()
            out["flatten_ids_unsorted"] = vals_a[:M].clone()
        keys_b = torch.empty_like(keys_a)
        vals_b = torch.empty_like(vals_a)
        scratch = _sort_scratch(M, dev)
        end_bit = 32 + tile_n_bits(tile_w, tile_h)
        with _stage("sort_pairs"):
            check(lib.b200_sort_pairs(_ptr(keys_a), _ptr(vals_a), _ptr(keys_b), _ptr(vals_b), M, end_bit,
                                      _ptr(scratch), scratch.numel(), C.byref(in_b), st), "b200_sort_pairs",
                  kernels=(2 + (end_bit + 7) // 8) if M > 0 else 0)
        keys_s, vals_s = (keys_b, vals_b) if in_b.value else (keys_a, vals_a)
        with _stage("tile_ranges"):
            check(lib.b200_tile_ranges(_ptr(keys_s), M, n_tiles, _ptr(offsets), st), "b200_tile_ranges",
                  kernels=1 if M > 0 else 0)
        out.update(M=M, isect_ids=keys_s[:M], flatten_ids=vals_s[:M], isect_offsets=offsets.view(tile_h, tile_w))
        return out

    # ---- split-key path ----
    if prep:
        digit_base = torch.empty(512, dtype=torch.int32, device=dev)
        total_prep = torch.empty(1, dtype=torch.int64, device=dev)
        # tile_prep (one CTA: 2-D prefix sum of the corner grid -> per-tile ranges, M) and the read-back of M depend only
        # on the projection: they go on a side stream and run under the depth sort.  The launching stream joins before it
        # first needs the ranges; the host collects M after the depth-sort launches are queued, so it never waits on an
        # idle GPU.
        cur, side = torch.cuda.current_stream(dev), _side_stream(dev)
        fork = torch.cuda.Event()
        fork.record(cur)
        side.wait_event(fork)
        m_host = _pinned_scalar(dev)
        m_ready = torch.cuda.Event()
        with torch.cuda.stream(side):
            if GRID_ON_SIDE_STREAM:  # the corner grid is counted here, not inside the projection kernel (b200_tile_count)
                with _stage("tile_count"):
                    check(lib.b200_tile_count(_ptr(splats), _ptr(radii), N, width, height, _ptr(tile_grid), side.cuda_stream),
                          "b200_tile_count", kernels=1 if N > 0 else 0)
            with _stage("tile_prep"):
                check(lib.b200_tile_prep(_ptr(tile_grid), width, height, _ptr(offsets), _ptr(total_prep), _ptr(digit_base),
                                         side.cuda_stream), "b200_tile_prep")
            m_host.copy_(total_prep, non_blocking=True)
            m_ready.record(side)
    dkeys_b = torch.empty_like(dkeys_a)
    gids_b = torch.empty_like(gids_a)
    scratch_n = _sort_scratch(N, dev)
    with _stage("sort_depth"):
        check(lib.b200_sort_pairs_k32(_ptr(dkeys_a), _ptr(gids_a), _ptr(dkeys_b), _ptr(gids_b), N, 32,
                                      _ptr(scratch_n), scratch_n.numel(), C.byref(in_b), st), "b200_sort_pairs_k32",
              kernels=6 if N > 0 else 0)
    perm = gids_b if in_b.value else gids_a
    if prep:
        torch.cuda.current_stream(dev).wait_event(m_ready)  # join: ranges / digit bases are used from here on
    with _stage("bin_scan"):
        check(lib.b200_bin_count_sorted(_ptr(tiles_per_gauss), _ptr(perm), N, _ptr(block_sums), st),
              "b200_bin_count_sorted", kernels=1 if N > 0 else 0)
        check(lib.b200_bin_scan(_ptr(block_sums), nblocks, _ptr(block_offsets), _ptr(total), st), "b200_bin_scan")
    if prep:
        m_ready.synchronize()  # the one host wait of the forward pass (sizes the intersection buffers)
        M = int(m_host.item())
    else:
        M = int(total.item())
    tkeys_a = torch.empty(max(M, 1), dtype=torch.int16, device=dev)
    vals_a = torch.empty(max(M, 1), dtype=torch.int32, device=dev)
    with _stage("bin_emit"):
        check(lib.b200_bin_emit_sorted(_ptr(splats), _ptr(radii), _ptr(tiles_per_gauss), _ptr(perm),
                                       _ptr(block_offsets), N, width, height, M, _ptr(tkeys_a), _ptr(vals_a), st),
              "b200_bin_emit_sorted", kernels=1 if N > 0 else 0)
    tkeys_b = torch.empty_like(tkeys_a)
    vals_b = torch.empty_like(vals_a)
    scratch_m = _sort_scratch(M, dev)
    tbits = _tile_id_bits(n_tiles)  # <= 16 on this path (n_tiles <= 65 536)
    with _stage("sort_pairs"):
        if prep:
            check(lib.b200_sort_pairs_k16_prepared(_ptr(tkeys_a), _ptr(vals_a), _ptr(tkeys_b), _ptr(vals_b), M, tbits,
                                                   _ptr(digit_base), _ptr(scratch_m), scratch_m.numel(),
                                                   C.byref(in_b), st), "b200_sort_pairs_k16_prepared",
                  kernels=((tbits + 7) // 8) if M > 0 else 0)
        else:
            check(lib.b200_sort_pairs_k16(_ptr(tkeys_a), _ptr(vals_a), _ptr(tkeys_b), _ptr(vals_b), M, tbits,
                                          _ptr(scratch_m), scratch_m.numel(), C.byref(in_b), st), "b200_sort_pairs_k16",
                  kernels=(2 + (tbits + 7) // 8) if M > 0 else 0)
    tkeys_s, vals_s = (tkeys_b, vals_b) if in_b.value else (tkeys_a, vals_a)
    if not prep:
        with _stage("tile_ranges"):
            check(lib.b200_tile_ranges_k16(_ptr(tkeys_s), M, n_tiles, _ptr(offsets), st), "b200_tile_ranges_k16",
                  kernels=1 if M > 0 else 0)
    out.update(M=M, flatten_ids=vals_s[:M], isect_offsets=offsets.view(tile_h, tile_w), tile_keys=tkeys_s[:M])
    if with_isect_ids:
        ids = torch.empty(max(M, 1), dtype=torch.int64, device=dev)
        check(lib.b200_isect_ids(_ptr(tkeys_s), _ptr(vals_s), _ptr(splats), M, _ptr(ids), st), "b200_isect_ids",
              kernels=1 if M > 0 else 0)
        out["isect_ids"] = ids[:M]
    else:
        out["isect_ids"] = None
    return out


# Optional provider of the gradient tensors the (non-accumulating) backward returns: distributed.GradientArena hands out, once
# per step, views of ONE flat buffer in symmetric memory, so that autograd's `.grad` tensors end up inside it without a copy
# and the multi-GPU all-reduce is a single multimem kernel (set with set_gradient_arena; None = torch.empty_like).
_GRAD_ARENA = None


def set_gradient_arena(arena) -> None:
    global _GRAD_ARENA
    _GRAD_ARENA = arena


_ACC_EVENTS = {}  # bare grad_into buffers only: (device, address) -> {"acc_event": event after the last in-place accumulation}


class _Aux:
    """Holder for the non-differentiable by-products of one RasterizeSplats.forward (filled in place).  A plain object on
    purpose: under autocast, custom_fwd's argument cast rebuilds dicts / lists, so a dict would be filled as a copy."""

    __slots__ = ("radii", "splats", "tiles_per_gauss", "isect_ids", "flatten_ids", "isect_offsets", "last_ids", "visibility")

    def __getitem__(self, key):
        return getattr(self, key)


class RasterizeSplats(torch.autograd.Function):
    """forward = stages 1-3, backward = stages 4a-4b (gsplat's autograd of rasterization())."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, means, quats, scales, opacities, colors, cam, width, height, sh_degree, chset, background,
                with_isect_ids=False, method="auto", raw=False, grad_into=None, aux=None):
        """Returns (render[H,W,CH], alpha[H,W,1]); the non-differentiable by-products (radii, splats, tiles_per_gauss,
        isect_ids, flatten_ids, isect_offsets, last_ids, visibility) are put into the holder ``aux`` (:class:`_Aux`) instead of being
        wrapped as autograd outputs (eight fewer output tensors to register per view on the host).
        raw=True: the five tensors are the RAW parameters (xyz, rotation, log-scaling, opacity logits,
        features) and the activations are fused into the kernels.  grad_into = (g_xyz, g_rotation, g_scaling,
        g_opacity, g_features): backward ADDS the raw-parameter gradients into these buffers in place (and
        returns None for the inputs) instead of handing fresh tensors to autograd."""
        _require_cuda(means, quats, scales, opacities, colors)
        means, quats, scales = means.contiguous(), quats.contiguous(), scales.contiguous()
        opacities, colors = opacities.contiguous(), colors.contiguous()
        dev = means.device
        if grad_into is not None and not raw:
            raise ValueError("grad_into needs raw=True")
        needs_grad = any(ctx.needs_input_grad[:5]) or grad_into is not None
        with _on_device(dev):
            b = project_and_bin(means, quats, scales, opacities, colors, sh_degree, cam, width, height,
                                method=method, with_isect_ids=with_isect_ids, raw=raw, want_grad_records=needs_grad)
            if b["isect_ids"] is None:
                b["isect_ids"] = torch.empty(0, dtype=torch.int64, device=dev)
            nch = _NCH[chset]
            render = torch.empty(height, width, nch, dtype=torch.float32, device=dev)
            alpha = torch.empty(height, width, 1, dtype=torch.float32, device=dev)
            last_ids = torch.empty(height, width, dtype=torch.int32, device=dev)
            bg, n_bg = background  # device tensor (or None) and the number of channels it covers
            with _stage("raster_fwd"):
                check(lib.b200_raster_fwd(_ptr(b["splats"]), _ptr(b["flatten_ids"]), _ptr(b["isect_offsets"]), b["M"],
                                          width, height, chset, _ptr(bg), n_bg, _ptr(render), _ptr(alpha),
                                          _ptr(last_ids), _stream()), "b200_raster_fwd")
        ctx.save_for_backward(means, quats, scales, opacities, colors, b["radii"], b["splats"], b["flatten_ids"],
                              b["isect_offsets"], alpha, last_ids)
        ctx.cfg = (cam, width, height, sh_degree, chset, bg, n_bg, b["M"], raw, grad_into)
        # zero-initialised cotangent records written by the projection kernel; consumed by the FIRST backward of this node
        # (a second one, retain_graph=True, allocates its own)
        ctx.grad_records = b["grad_records"]
        ctx.set_materialize_grads(False)  # unused outputs hand backward() None instead of freshly zero-filled tensors
        if aux is not None:
            aux.radii, aux.splats, aux.tiles_per_gauss, aux.isect_ids = b["radii"], b["splats"], b["tiles_per_gauss"], b["isect_ids"]
            aux.flatten_ids, aux.isect_offsets, aux.last_ids, aux.visibility = b["flatten_ids"], b["isect_offsets"], last_ids, b["visibility"]
        return render, alpha

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, v_render, v_alpha):
        means, quats, scales, opacities, colors, radii, splats, flatten_ids, offsets, alpha, last_ids = ctx.saved_tensors
        cam, width, height, sh_degree, chset, bg, n_bg, M, raw, grad_into = ctx.cfg
        dev = means.device
        N = means.shape[0]
        nch = _NCH[chset]
        if v_render is None:
            v_render = torch.zeros(height, width, nch, dtype=torch.float32, device=dev)
        v_render = v_render.to(torch.float32).contiguous()
        if v_alpha is not None:
            v_alpha = v_alpha.to(torch.float32).contiguous()
        with _on_device(dev):
            st = _stream()
            v_splats, ctx.grad_records = ctx.grad_records, None
            if v_splats is None:
                v_splats = torch.zeros(N, _lib.SPLAT_FLOATS, dtype=torch.float32, device=dev)
            with _stage("raster_bwd"):
                check(lib.b200_raster_bwd(_ptr(splats), _ptr(flatten_ids), _ptr(offsets), M, width, height, chset,
                                          _ptr(bg), n_bg, _ptr(alpha), _ptr(last_ids), _ptr(v_render), _ptr(v_alpha), _ptr(v_splats),
                                          st), "b200_raster_bwd")
            if sh_degree is None:
                sh_coeffs, deg = 1, -1
            else:
                sh_coeffs, deg = colors.shape[1], int(sh_degree)
            write_first = False
            if grad_into is not None:
                # optional sixth element: the store's lazy-zero state — the first backward after zero_grad(lazy=True)
                # writes the buffers (as the non-accumulating mode does), the following ones add to them
                state = grad_into[5] if len(grad_into) > 5 else None
                grad_into = tuple(grad_into[:5])
                if state is not None and state.get("pending_zero"):
                    write_first = True
                    state["pending_zero"] = False
                v_means, v_quats, v_scales, v_opac, v_colors = grad_into
                for gbuf, ref in zip(grad_into, (means, quats, scales, opacities, colors)):
                    if gbuf.numel() != ref.numel() or not gbuf.is_contiguous() or gbuf.dtype != torch.float32:
                        raise RuntimeError("grad_into buffers must be contiguous fp32 with the parameters' sizes")
            else:
                bufs = _GRAD_ARENA.take((means, quats, scales, opacities, colors)) if _GRAD_ARENA is not None else None
                if bufs is not None:  # first backward of the step: its outputs ARE the arena's slices (autograd adopts them as .grad)
                    v_means, v_quats, v_scales, v_opac, v_colors = bufs
                else:
                    v_means = torch.empty_like(means)
                    v_quats = torch.empty_like(quats)
                    v_scales = torch.empty_like(scales)
                    v_opac = torch.empty_like(opacities)
                    v_colors = torch.empty_like(colors)
            acc_holder = None
            if grad_into is not None:
                # the in-place "+=" of the accumulate mode is not atomic: views in flight on different streams
                # (streams.ViewStreams) take turns on the gradient buffer.  The event of the last accumulation lives WITH the
                # buffer's owner (the store's grad-state dict: it dies with the buffer, so a re-allocated store can never meet
                # a stale event); callers that pass bare buffers share one slot per (device, address)
                acc_holder = state if state is not None else _ACC_EVENTS.setdefault((dev.index, v_means.data_ptr()), {})
                prev = acc_holder.get("acc_event")
                if prev is not None:
                    torch.cuda.current_stream().wait_event(prev)
            with _stage("project_sh_bwd"):
                if raw:
                    check(lib.b200_project_sh_bwd_raw(_ptr(means), _ptr(quats), _ptr(scales), _ptr(opacities),
                                                      _ptr(colors), sh_coeffs, deg, N, None, _ptr(cam), width, height,
                                                      _ptr(radii), _ptr(v_splats), _ptr(v_means), _ptr(v_quats),
                                                      _ptr(v_scales), _ptr(v_opac), _ptr(v_colors),
                                                      1 if (grad_into is not None and not write_first) else 0, st),
                          "b200_project_sh_bwd_raw")
                else:
                    check(lib.b200_project_sh_bwd(_ptr(means), _ptr(quats), _ptr(scales), _ptr(opacities), _ptr(colors),
                                                  sh_coeffs, deg, N, None, _ptr(cam), width, height, _ptr(radii),
                                                  _ptr(v_splats), _ptr(v_means), _ptr(v_quats), _ptr(v_scales), _ptr(v_opac),
                                                  _ptr(v_colors), st), "b200_project_sh_bwd")
        if grad_into is not None:
            ev = acc_holder.get("acc_event") or torch.cuda.Event()
            with _on_device(dev):
                ev.record()
            acc_holder["acc_event"] = ev
            return (None,) * 16
        return v_means, v_quats, v_scales, v_opac, v_colors, None, None, None, None, None, None, None, None, None, None, None


def _render_one(means, quats, scales, opacities, colors, viewmat, K, width, height, sh_degree, render_mode,
                background, near_plane, far_plane, eps2d, radius_clip, with_isect_ids=False, method="auto",
                raw=False, grad_into=None):
    chset = _MODES[render_mode]
    nch = _NCH[chset]
    # the background stays on the device (no host read-back); the depth channel composites over 0
    if background is None or chset == _lib.CH_D:
        bg = (None, 0)
    else:
        bgt = background if (background.dim() == 1 and _plain_f32(background, means.device)) else \
            background.detach().reshape(-1).to(device=means.device, dtype=torch.float32).contiguous()
        if bgt.numel() < 3:
            raise ValueError("backgrounds must hold 3 colour channels")
        bg = (bgt, 3)
    # (FRONT_PRIORITY: the camera block is packed on the high-priority companion stream too — a one-CTA kernel on the
    # caller's stream would wait behind every CTA of the other view's blend, and the projection behind it)
    front = _enter_front(means.device).cuda_stream if (FRONT_PRIORITY and method in ("auto", "coarse", "coarse-gather")
                                                       and _coarse_supported(int(width), int(height))) else None
    cam = make_cam(viewmat, K, int(width), int(height), near_plane, far_plane, eps2d, radius_clip, stream=front)
    aux = _Aux()
    render, alpha = RasterizeSplats.apply(means, quats, scales, opacities, colors, cam, int(width), int(height), sh_degree,
                                          chset, bg, with_isect_ids, method, raw, grad_into, aux)
    if render_mode in ("ED", "RGB+ED"):
        render = torch.cat([render[..., :-1], render[..., -1:] / alpha.clamp(min=1e-10)], dim=-1)
    return render, alpha, aux


def rasterization(means: torch.Tensor, quats: torch.Tensor, scales: torch.Tensor, opacities: torch.Tensor,
                  colors: torch.Tensor, viewmats: torch.Tensor, Ks: torch.Tensor, width: int, height: int,
                  near_plane: float = 0.01, far_plane: float = 1e10, radius_clip: float = 0.0,
                  eps2d: float = 0.3, sh_degree: Optional[int] = None, packed: bool = False,
                  tile_size: int = 16, backgrounds: Optional[torch.Tensor] = None, render_mode: str = "RGB",
                  meta_isect_ids: bool = False, binning: str = "auto", raw_params: bool = False,
                  grad_into=None, **unsupported) -> Tuple[torch.Tensor, torch.Tensor, Dict]:
    """Drop-in for ``gsplat.rasterization`` with the argument set the reference uses.

    means[N,3] quats[N,4](wxyz) scales[N,3] opacities[N] colors=[N,K,3] SH coefficients (``sh_degree``
    given) or [N,3] colours; viewmats[C,4,4] world->camera, Ks[C,3,3] pixel intrinsics.  Returns
    ``(renders[C,H,W,CH], alphas[C,H,W,1], meta)``.

    Extensions (not gsplat arguments): ``meta_isect_ids=True`` materialises the sorted int64
    ``meta["isect_ids"]`` (the fast path never needs them: the reference reads only ``means2d`` and
    ``radii``, gsplat_backend.py:204-208); ``binning`` selects "split" / "keys64" / "auto" (project_and_bin);
    ``raw_params=True`` takes the RAW parameters of the reference's GaussianModel (scales = log-scaling,
    opacities = logits, quats un-normalised, colors = features_dc ++ features_rest) and applies the getters'
    activations inside the kernels; ``grad_into`` (raw only) accumulates the gradients into caller buffers
    (see :class:`RasterizeSplats`, :mod:`store`).
    """
    if render_mode not in _MODES:
        raise ValueError(f"render_mode must be one of {sorted(_MODES)}, got {render_mode!r}")
    if packed:
        raise NotImplementedError("packed=True is not used by the reference (gsplat_backend.py:178)")
    if tile_size != _lib.TILE:
        raise NotImplementedError("tile_size is fixed at 16 (gsplat default, the only value the reference uses)")
    for k, v in unsupported.items():
        if v not in (None, False, "classic", "pinhole"):
            raise NotImplementedError(f"rasterization(): argument {k}={v!r} is outside the reference's call")
    _require_cuda(means, quats, scales, opacities, colors, viewmats, Ks)
    if opacities.dim() == 2:
        opacities = opacities.squeeze(-1)
    if viewmats.dim() == 2:
        viewmats, Ks = viewmats[None], Ks[None]
    Ccam = viewmats.shape[0]
    renders, alphas, metas = [], [], []
    for c in range(Ccam):
        bg = None if backgrounds is None else (backgrounds[c] if backgrounds.dim() == 2 else backgrounds)
        r, a, m = _render_one(means, quats, scales, opacities, colors, viewmats[c], Ks[c], width, height, sh_degree,
                              render_mode, bg, near_plane, far_plane, eps2d, radius_clip, meta_isect_ids, binning,
                              raw_params, grad_into)
        renders.append(r)
        alphas.append(a)
        metas.append(m)
    tile_w = (width + _lib.TILE - 1) // _lib.TILE
    tile_h = (height + _lib.TILE - 1) // _lib.TILE
    def _stk(key):  # C == 1 (the reference's only case) stays a zero-copy view
        return metas[0][key][None] if Ccam == 1 else torch.stack([m[key] for m in metas])

    sp = _stk("splats")
    meta = {
        "radii": _stk("radii"),
        "means2d": sp[..., 0:2],
        "conics": sp[..., 2:5],
        "depths": sp[..., 9],
        "colors": sp[..., 6:9],
        "opacities": opacities[None].expand(Ccam, -1),
        "tiles_per_gauss": _stk("tiles_per_gauss"),
        "isect_ids": metas[0]["isect_ids"] if Ccam == 1 else [m["isect_ids"] for m in metas],
        "flatten_ids": metas[0]["flatten_ids"] if Ccam == 1 else [m["flatten_ids"] for m in metas],
        "isect_offsets": _stk("isect_offsets"),
        "last_ids": _stk("last_ids"),
        "visibility_filter": _stk("visibility"),  # radii > 0, written by the projection kernel
        "tile_width": tile_w, "tile_height": tile_h, "tile_size": _lib.TILE,
        "width": width, "height": height, "n_cameras": Ccam,
    }
    if Ccam == 1:
        return renders[0][None], alphas[0][None], meta
    return torch.stack(renders), torch.stack(alphas), meta
