"""CPU oracle (TEST INFRASTRUCTURE ONLY) — pure-PyTorch restatement of the rasterizer the reference calls.

Status: **parity unpinned**.  The arithmetic the reference runs on this path lives in the third-party
wheel ``gsplat==1.5.3`` (reference ``requirements.txt:9``; called at
``src/backends/gsplat_backend.py:83-97`` and ``:168-182``).  That wheel is not vendored under
``/root/reference``, is not installed, and cannot be fetched (no network); the reference's own tests
never touch the rasterizer.  This file therefore restates gsplat 1.5.x's *published*
algorithm (EWA splatting / 3DGS as documented by gsplat; SURVEY.md Appendix A is the spec) and is
self-validated by finite differences and closed-form cases in ``tests/``.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module.  The product path (``3d-gaussian-splatting_b200``) never does.

Every stage is a separate function returning every intermediate so that the CUDA path can be compared
stage by stage.  All functions are dtype-generic (fp32 for parity, fp64 for gradient self-checks) and
differentiable by autograd where the math is differentiable.

Conventions: quaternions (w,x,y,z); viewmat = world->camera, OpenCV axes; pixel
centre (j+0.5, i+0.5); images HWC; tile_size 16; C = 1 camera; packed=False.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch

ALPHA_THRESHOLD = 1.0 / 255.0
MAX_ALPHA = 0.999
T_STOP = 1e-4
TILE = 16

SH_C0 = 0.2820947917738781
SH_C1 = 0.48860251190292


# --------------------------------------------------------------------------------------------------
# A.1  projection  (gsplat call site: gsplat_backend.py:168-182, argument ``means/quats/scales``)
# --------------------------------------------------------------------------------------------------
def quat_to_rotmat(q: torch.Tensor) -> torch.Tensor:
    """Normalises (gsplat normalises internally although gaussian_model.py:98-100 already did)."""
    q = q / q.norm(dim=-1, keepdim=True)
    w, x, y, z = q.unbind(-1)
    R = torch.stack(
        [
            1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y),
        ],
        dim=-1,
    )
    return R.reshape(q.shape[:-1] + (3, 3))


def project(
    means: torch.Tensor,      # [N,3]
    quats: torch.Tensor,      # [N,4]
    scales: torch.Tensor,     # [N,3]  (post-exp)
    opacities: torch.Tensor,  # [N]    (post-sigmoid)
    viewmat: torch.Tensor,    # [4,4]
    K: torch.Tensor,          # [3,3]
    W: int,
    H: int,
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    eps2d: float = 0.3,
    radius_clip: float = 0.0,
) -> Dict[str, torch.Tensor]:
    """SURVEY.md A.1.  Returns radii[N,2] int32, means2d[N,2], depths[N], conics[N,3] (0 where culled)."""
    dt = means.dtype
    R = viewmat[:3, :3].to(dt)
    t = viewmat[:3, 3].to(dt)
    fx, fy, cx, cy = K[0, 0].to(dt), K[1, 1].to(dt), K[0, 2].to(dt), K[1, 2].to(dt)

    pc = means @ R.T + t
    x, y, z = pc.unbind(-1)
    valid = (z >= near_plane) & (z <= far_plane)
    zs = torch.where(valid, z, torch.ones_like(z))  # keep culled rows finite

    Rq = quat_to_rotmat(quats)
    M = Rq * scales[:, None, :]
    Sigma = M @ M.transpose(1, 2)
    Sc = R @ Sigma @ R.T

    tanx = 0.5 * W / fx
    tany = 0.5 * H / fy
    lim_xp = (W - cx) / fx + 0.3 * tanx
    lim_xn = cx / fx + 0.3 * tanx
    lim_yp = (H - cy) / fy + 0.3 * tany
    lim_yn = cy / fy + 0.3 * tany
    rz = 1.0 / zs
    rz2 = rz * rz
    tx = zs * torch.minimum(lim_xp, torch.maximum(-lim_xn, x * rz))
    ty = zs * torch.minimum(lim_yp, torch.maximum(-lim_yn, y * rz))
    zero = torch.zeros_like(rz)
    J = torch.stack(
        [fx * rz, zero, -fx * tx * rz2, zero, fy * rz, -fy * ty * rz2], dim=-1
    ).reshape(-1, 2, 3)
    S2 = J @ Sc @ J.transpose(1, 2)
    mx = fx * x * rz + cx
    my = fy * y * rz + cy

    a = S2[:, 0, 0] + eps2d
    b = S2[:, 0, 1]
    c = S2[:, 1, 1] + eps2d
    det = a * c - b * b
    valid = valid & (det > 0)
    dets = torch.where(valid, det, torch.ones_like(det))
    conics = torch.stack([c / dets, -b / dets, a / dets], dim=-1)

    valid = valid & (opacities >= ALPHA_THRESHOLD)
    o_safe = torch.where(valid, opacities, torch.ones_like(opacities)).detach()
    extend = torch.minimum(
        torch.full_like(o_safe, 3.33), torch.sqrt(2.0 * torch.log(o_safe / ALPHA_THRESHOLD))
    )
    rx = torch.ceil(extend * torch.sqrt(a.detach().clamp_min(0)))
    ry = torch.ceil(extend * torch.sqrt(c.detach().clamp_min(0)))
    valid = valid & ~((rx <= radius_clip) & (ry <= radius_clip))
    mxd, myd = mx.detach(), my.detach()
    valid = valid & ~((mxd + rx <= 0) | (mxd - rx >= W) | (myd + ry <= 0) | (myd - ry >= H))

    vm = valid[:, None]
    radii = torch.where(vm, torch.stack([rx, ry], -1), torch.zeros_like(means[:, :2])).to(torch.int32)
    means2d = torch.where(vm, torch.stack([mx, my], -1), torch.zeros_like(means[:, :2]))
    depths = torch.where(valid, z, torch.zeros_like(z))
    conics = torch.where(vm, conics, torch.zeros_like(conics))
    return {"radii": radii, "means2d": means2d, "depths": depths, "conics": conics}


# --------------------------------------------------------------------------------------------------
# A.2  spherical harmonics -> RGB
# --------------------------------------------------------------------------------------------------
def sh_basis(dirs: torch.Tensor, degree: int) -> torch.Tensor:
    """Real SH basis values [N, 16] for unit directions; columns >= (degree+1)^2 are zero."""
    x, y, z = dirs.unbind(-1)
    N = dirs.shape[0]
    out = [torch.full_like(x, SH_C0)]
    if degree >= 1:
        out += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if degree >= 2:
        z2 = z * z
        fT0B = -1.092548430592079 * z
        fC1 = x * x - y * y
        fS1 = 2 * x * y
        out += [
            0.5462742152960395 * fS1,
            fT0B * y,
            0.9461746957575601 * z2 - 0.3153915652525201,
            fT0B * x,
            0.5462742152960395 * fC1,
        ]
    if degree >= 3:
        fT0C = -2.285228997322329 * z2 + 0.4570457994644658
        fT1B = 1.445305721320277 * z
        fC2 = x * fC1 - y * fS1
        fS2 = x * fS1 + y * fC1
        out += [
            -0.5900435899266435 * fS2,
            fT1B * fS1,
            fT0C * y,
            z * (1.865881662950577 * z2 - 1.119528997770346),
            fT0C * x,
            fT1B * fC1,
            -0.5900435899266435 * fC2,
        ]
    B = torch.stack(out, dim=-1)
    if B.shape[1] < 16:
        B = torch.cat([B, B.new_zeros(N, 16 - B.shape[1])], dim=1)
    return B


def sh_colors(
    means: torch.Tensor, shs: torch.Tensor, viewmat: torch.Tensor, degree: int, visible: torch.Tensor
) -> torch.Tensor:
    """colours [N,3] = max(SH(dir)+0.5, 0) for visible Gaussians, 0 elsewhere."""
    dt = means.dtype
    R = viewmat[:3, :3].to(dt)
    t = viewmat[:3, 3].to(dt)
    campos = -(R.T @ t)
    u = means - campos
    d = u / u.norm(dim=-1, keepdim=True)
    B = sh_basis(d, degree)[:, : shs.shape[1]]
    rgb = (B[:, :, None] * shs).sum(dim=1)
    col = torch.clamp_min(rgb + 0.5, 0.0)
    return torch.where(visible[:, None], col, torch.zeros_like(col))


# --------------------------------------------------------------------------------------------------
# A.3  tile binning, sort, per-tile ranges  (integer, bit-exact target)
# --------------------------------------------------------------------------------------------------
def tile_bounds(means2d: torch.Tensor, radii: torch.Tensor, tile_w: int, tile_h: int):
    m = means2d.detach().to(torch.float32)
    r = radii.to(torch.float32)
    tx = m[:, 0] / TILE
    ty = m[:, 1] / TILE
    rx = r[:, 0] / TILE
    ry = r[:, 1] / TILE
    # (uint32) cast of a negative float saturates to 0 on the device; clamp before the cast here.
    minx = torch.floor(tx - rx).clamp(0, tile_w).to(torch.int64)
    miny = torch.floor(ty - ry).clamp(0, tile_h).to(torch.int64)
    maxx = torch.ceil(tx + rx).clamp(0, tile_w).to(torch.int64)
    maxy = torch.ceil(ty + ry).clamp(0, tile_h).to(torch.int64)
    vis = (radii > 0).all(dim=-1)
    zero = torch.zeros_like(minx)
    return (
        torch.where(vis, minx, zero), torch.where(vis, miny, zero),
        torch.where(vis, maxx, zero), torch.where(vis, maxy, zero),
    )


def tile_n_bits(tile_w: int, tile_h: int) -> int:
    return int(math.floor(math.log2(tile_w * tile_h))) + 1


def bin_tiles(
    means2d: torch.Tensor, radii: torch.Tensor, depths: torch.Tensor, W: int, H: int
) -> Dict[str, torch.Tensor]:
    """Emits unsorted (isect_id int64, flatten_id int32) pairs in gsplat's emission order."""
    tile_w = (W + TILE - 1) // TILE
    tile_h = (H + TILE - 1) // TILE
    minx, miny, maxx, maxy = tile_bounds(means2d, radii, tile_w, tile_h)
    bw = maxx - minx
    tiles_per_gauss = ((maxy - miny) * bw).to(torch.int32)
    cum = torch.cumsum(tiles_per_gauss.to(torch.int64), dim=0)
    M = int(cum[-1]) if cum.numel() else 0
    g = torch.repeat_interleave(torch.arange(means2d.shape[0]), tiles_per_gauss.to(torch.int64))
    start = cum - tiles_per_gauss.to(torch.int64)
    k = torch.arange(M) - start[g]
    bwg = bw[g].clamp_min(1)
    ti = miny[g] + k // bwg
    tj = minx[g] + k % bwg
    tile_id = ti * tile_w + tj
    depth_bits = depths.detach().to(torch.float32).contiguous().view(torch.int32).to(torch.int64)
    isect_ids = (tile_id << 32) | (depth_bits[g] & 0xFFFFFFFF)
    return {
        "tiles_per_gauss": tiles_per_gauss,
        "cum_tiles": cum,
        "isect_ids": isect_ids,
        "flatten_ids": g.to(torch.int32),
        "tile_w": tile_w,
        "tile_h": tile_h,
    }


def sort_and_ranges(isect_ids: torch.Tensor, flatten_ids: torch.Tensor, tile_w: int, tile_h: int):
    """Stable ascending sort by key; offsets[t] = first sorted index whose tile id >= t."""
    order = torch.sort(isect_ids, stable=True).indices
    ids_s = isect_ids[order]
    flat_s = flatten_ids[order]
    n_tiles = tile_w * tile_h
    tile_of = ids_s >> 32
    offsets = torch.searchsorted(tile_of, torch.arange(n_tiles, dtype=torch.int64)).to(torch.int32)
    return ids_s, flat_s, offsets.reshape(tile_h, tile_w)


# --------------------------------------------------------------------------------------------------
# A.4  forward blend
# --------------------------------------------------------------------------------------------------
def rasterize(
    means2d: torch.Tensor,    # [N,2]
    conics: torch.Tensor,     # [N,3]
    colors: torch.Tensor,     # [N,CH]
    opacities: torch.Tensor,  # [N]
    W: int,
    H: int,
    isect_offsets: torch.Tensor,  # [tile_h, tile_w] int32
    flatten_ids: torch.Tensor,    # [M] int32
    background: Optional[torch.Tensor],  # [CH] or None
    margin_eps: float = 0.0,
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
    """SURVEY.md A.4, one tile at a time, the sequential per-pixel loop restated with cumprod.

    Returns render[H,W,CH], alpha[H,W,1], last_ids[H,W] int32 and (if margin_eps>0) a bool map of
    pixels for which some skip/stop decision was within a relative ``margin_eps`` of its threshold;
    those pixels may legitimately differ between two fp32 implementations by one blend term.
    """
    dt = means2d.dtype
    CH = colors.shape[1]
    tile_h, tile_w = isect_offsets.shape
    M = flatten_ids.shape[0]
    offs = isect_offsets.reshape(-1).to(torch.int64).tolist() + [M]
    render = torch.zeros(H, W, CH, dtype=dt)
    alpha_img = torch.zeros(H, W, 1, dtype=dt)
    last_ids = torch.zeros(H, W, dtype=torch.int32)
    marginal = torch.zeros(H, W, dtype=torch.bool) if margin_eps > 0 else None
    bg = background.to(dt) if background is not None else None
    out_tiles = []
    for t in range(tile_h * tile_w):
        s, e = offs[t], offs[t + 1]
        ti, tj = divmod(t, tile_w)
        y0, x0 = ti * TILE, tj * TILE
        y1, x1 = min(y0 + TILE, H), min(x0 + TILE, W)
        ph, pw = y1 - y0, x1 - x0
        if e <= s:
            if bg is not None:
                out_tiles.append((y0, y1, x0, x1, bg.expand(ph, pw, CH), torch.zeros(ph, pw, 1, dtype=dt)))
            continue
        ids = flatten_ids[s:e].to(torch.int64)
        py = (torch.arange(y0, y1, dtype=dt) + 0.5)[:, None].expand(ph, pw).reshape(-1)
        px = (torch.arange(x0, x1, dtype=dt) + 0.5)[None, :].expand(ph, pw).reshape(-1)
        m = means2d[ids]
        cn = conics[ids]
        o = opacities[ids]
        dx = m[:, 0:1] - px[None, :]
        dy = m[:, 1:2] - py[None, :]
        sigma = 0.5 * (cn[:, 0:1] * dx * dx + cn[:, 2:3] * dy * dy) + cn[:, 1:2] * dx * dy
        a_raw = o[:, None] * torch.exp(-sigma)
        a = torch.clamp_max(a_raw, MAX_ALPHA)
        valid = (sigma >= 0) & (a >= ALPHA_THRESHOLD)
        a = torch.where(valid, a, torch.zeros_like(a))
        Tn = torch.cumprod(1.0 - a, dim=0)
        stop = valid & (Tn <= T_STOP)
        done = torch.cumsum(stop.to(torch.int32), dim=0) > 0
        a_eff = torch.where(done, torch.zeros_like(a), a)
        T_incl = torch.cumprod(1.0 - a_eff, dim=0)
        T_before = torch.cat([torch.ones_like(T_incl[:1]), T_incl[:-1]], dim=0)
        w = a_eff * T_before
        T_final = T_incl[-1]
        col = w.transpose(0, 1) @ colors[ids]
        if bg is not None:
            col = col + T_final[:, None] * bg[None, :]
        comp = a_eff > 0
        G = e - s
        idxs = torch.arange(G)[:, None].expand(G, ph * pw)
        lastk = torch.where(comp, idxs, torch.full_like(idxs, -1)).max(dim=0).values
        last = torch.where(lastk >= 0, lastk + s, torch.zeros_like(lastk)).to(torch.int32)
        out_tiles.append((y0, y1, x0, x1, col.reshape(ph, pw, CH), (1.0 - T_final).reshape(ph, pw, 1)))
        last_ids[y0:y1, x0:x1] = last.reshape(ph, pw)
        if marginal is not None:
            with torch.no_grad():
                first_stop = stop & (torch.cumsum(stop.to(torch.int32), dim=0) == 1)
                live = (~done) | first_stop  # decisions the sequential loop actually evaluates
                near_a = ((a_raw - ALPHA_THRESHOLD).abs() <= margin_eps * ALPHA_THRESHOLD) & (sigma >= 0)
                near_t = ((Tn - T_STOP).abs() <= margin_eps * T_STOP) & valid
                near_s = sigma.abs() <= margin_eps * 1e-3
                mg = ((near_a | near_t | near_s) & live).any(dim=0)
                marginal[y0:y1, x0:x1] = mg.reshape(ph, pw)
    # assemble without in-place writes on a leaf so autograd stays simple
    if out_tiles:
        render_rows = [[None] * tile_w for _ in range(tile_h)]
        alpha_rows = [[None] * tile_w for _ in range(tile_h)]
        for (y0, y1, x0, x1, c, al) in out_tiles:
            render_rows[y0 // TILE][x0 // TILE] = c
            alpha_rows[y0 // TILE][x0 // TILE] = al
        for ti in range(tile_h):
            for tj in range(tile_w):
                if render_rows[ti][tj] is None:
                    ph = min(TILE, H - ti * TILE)
                    pw = min(TILE, W - tj * TILE)
                    render_rows[ti][tj] = torch.zeros(ph, pw, CH, dtype=dt)
                    alpha_rows[ti][tj] = torch.zeros(ph, pw, 1, dtype=dt)
        render = torch.cat([torch.cat(r, dim=1) for r in render_rows], dim=0)
        alpha_img = torch.cat([torch.cat(r, dim=1) for r in alpha_rows], dim=0)
    return render, alpha_img, last_ids, marginal


# --------------------------------------------------------------------------------------------------
# Whole pipeline == gsplat.rasterization(...) as the reference calls it (gsplat_backend.py:168-182)
# --------------------------------------------------------------------------------------------------
def rasterization(
    means: torch.Tensor,
    quats: torch.Tensor,
    scales: torch.Tensor,
    opacities: torch.Tensor,
    colors: torch.Tensor,          # shs [N,K,3] when sh_degree is not None, else [N,3]
    viewmat: torch.Tensor,         # [4,4]
    K: torch.Tensor,               # [3,3]
    width: int,
    height: int,
    sh_degree: Optional[int] = 3,
    background: Optional[torch.Tensor] = None,
    render_mode: str = "RGB",
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    eps2d: float = 0.3,
    radius_clip: float = 0.0,
    margin_eps: float = 0.0,
):
    """Returns (render[H,W,CH], alpha[H,W,1], meta) with every intermediate in ``meta``."""
    assert render_mode in ("RGB", "D", "ED", "RGB+D", "RGB+ED")
    if opacities.dim() == 2:
        opacities = opacities.squeeze(-1)
    proj = project(means, quats, scales, opacities, viewmat, K, width, height,
                   near_plane, far_plane, eps2d, radius_clip)
    visible = (proj["radii"] > 0).all(dim=-1)
    if sh_degree is not None:
        rgb = sh_colors(means, colors, viewmat, sh_degree, visible)
    else:
        rgb = colors
    depths = proj["depths"]
    bg = background
    if render_mode in ("D", "ED"):
        feat = depths[:, None]
        bg = torch.zeros(1, dtype=means.dtype) if background is not None else None
    elif render_mode in ("RGB+D", "RGB+ED"):
        feat = torch.cat([rgb, depths[:, None]], dim=-1)
        bg = torch.cat([background.to(means.dtype), torch.zeros(1, dtype=means.dtype)]) if background is not None else None
    else:
        feat = rgb
    b = bin_tiles(proj["means2d"], proj["radii"], depths, width, height)
    ids_s, flat_s, offsets = sort_and_ranges(b["isect_ids"], b["flatten_ids"], b["tile_w"], b["tile_h"])
    render, alpha, last_ids, marginal = rasterize(
        proj["means2d"], proj["conics"], feat, opacities, width, height, offsets, flat_s, bg, margin_eps
    )
    if render_mode in ("ED", "RGB+ED"):
        render = torch.cat(
            [render[..., :-1], render[..., -1:] / alpha.clamp(min=1e-10)], dim=-1
        )
    meta = dict(proj)
    meta.update(
        colors=rgb, opacities=opacities, tiles_per_gauss=b["tiles_per_gauss"], cum_tiles=b["cum_tiles"],
        isect_ids_unsorted=b["isect_ids"], flatten_ids_unsorted=b["flatten_ids"],
        isect_ids=ids_s, flatten_ids=flat_s, isect_offsets=offsets, last_ids=last_ids,
        tile_width=b["tile_w"], tile_height=b["tile_h"], width=width, height=height, tile_size=TILE,
        n_cameras=1, marginal=marginal,
    )
    return render, alpha, meta


# --------------------------------------------------------------------------------------------------
# A.7  stage 5: densification statistics (reference trainer.py:786-791)
# --------------------------------------------------------------------------------------------------
def densify_stats(accum: torch.Tensor, count: torch.Tensor, xyz_grad: torch.Tensor) -> None:
    """accum[N,3] += ||grad||_2 (broadcast to 3 columns); count[N,1] += 1 — in place, like the reference."""
    accum += torch.norm(xyz_grad, dim=-1, keepdim=True)
    count += 1


# --------------------------------------------------------------------------------------------------
# Synthetic scene generator — deterministic on CPU so oracle and GPU see the same bits
# --------------------------------------------------------------------------------------------------
def synthetic_scene(N: int, W: int, H: int, seed: int = 0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    fx = fy = (24.0 / 36.0) * max(W, H)
    cx, cy = W / 2.0, H / 2.0
    z = 2.0 + 8.0 * torch.rand(N, generator=g)
    u = torch.rand(N, generator=g)
    v = torch.rand(N, generator=g)
    x = z * (u - 0.5) * 1.1 * W / fx
    y = z * (v - 0.5) * 1.1 * H / fy
    xyz = torch.stack([x, y, z], dim=-1)
    log_scale = (math.log(0.012) + 0.6 * torch.randn(N, 3, generator=g)).clamp(math.log(1e-4), math.log(0.3))
    rot = torch.randn(N, 4, generator=g)
    opacity_logit = 2.0 * torch.randn(N, 1, generator=g)
    features_dc = (torch.rand(N, 1, 3, generator=g) * 2 - 1) * 1.77
    features_rest = 0.1 * torch.randn(N, 15, 3, generator=g)
    Kmat = torch.tensor([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=torch.float32)
    out = dict(
        xyz=xyz, scaling=log_scale, rotation=rot, opacity=opacity_logit,
        features_dc=features_dc, features_rest=features_rest,
        K=Kmat, viewmat=torch.eye(4),
    )
    return {k: (t.to(dtype) if t.is_floating_point() else t) for k, t in out.items()}


def activate(scene: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Activations exactly as the reference's getters (gaussian_model.py:90-108)."""
    return {
        "means3D": scene["xyz"],
        "scales": torch.exp(scene["scaling"]),
        "rotations": torch.nn.functional.normalize(scene["rotation"], dim=-1),
        "opacities": torch.sigmoid(scene["opacity"]),
        "shs": torch.cat([scene["features_dc"], scene["features_rest"]], dim=1),
    }


def circle_cameras(n: int, radius: float = 1.0, target=(0.0, 0.0, 6.0)) -> torch.Tensor:
    """Config-B cameras: n w2c matrices on a circle in the xz-plane looking at target."""
    mats = []
    tgt = torch.tensor(target, dtype=torch.float64)
    for i in range(n):
        th = 2 * math.pi * i / n
        # camera sits on a circle around a point `radius` in front of the target so that it keeps
        # looking roughly down +z (the scene only exists in front of the origin camera)
        pos = torch.tensor([radius * math.sin(th), 0.0, radius * (1 - math.cos(th)) * 0.5], dtype=torch.float64)
        fwd = tgt - pos
        fwd = fwd / fwd.norm()
        up = torch.tensor([0.0, -1.0, 0.0], dtype=torch.float64)
        right = torch.linalg.cross(fwd, up)
        right = right / right.norm()
        down = torch.linalg.cross(fwd, right)
        Rc = torch.stack([right, down, fwd], dim=0)  # rows = camera axes in world coords
        m = torch.eye(4, dtype=torch.float64)
        m[:3, :3] = Rc
        m[:3, 3] = -Rc @ pos
        mats.append(m)
    return torch.stack(mats).to(torch.float32)
