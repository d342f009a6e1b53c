"""Full-size GPU parity at BASELINE.json's own sizes — configs[1] = A (1 M Gaussians SH3, 1920x1080; also with a posed
camera and asymmetric intrinsics), configs[2] = B (3 M, three of the 64 circle cameras), configs[3] = C (6 M, 3840x2160)
— and the committed golden fixture.

At this size the PyTorch oracle is too slow, so the checker is the C/OpenMP port of the same spec
(oracle/cpu_splat.c, validated against the PyTorch oracle and the golden fixture in tests/test_oracle_cpu.py),
plus size-independent properties of the pipeline: bit-exact binning given the GPU's own projection outputs,
sortedness, range consistency, background linearity, cotangent linearity, invariance under a permutation of
the Gaussians.  Tolerances: integers bit-exact; render max-abs <= 1e-4; gradients relative error <= 1e-3.
"""
import ctypes as C
import math
import os

import pytest
import torch

from parity_util import apply_camera, grad_rows, image_close

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "splat_small.pt")
N_A, W_A, H_A, DEG = 1_000_000, 1920, 1080, 3


@pytest.fixture(scope="module")
def scene_a(oracle):
    sc = oracle.synthetic_scene(N_A, W_A, H_A, seed=0)
    return sc, oracle.activate(sc)


def _render(pkg, act, sc, W, H, deg, bg=None, requires_grad=False, perm=None):
    leaves = {}
    for k in ("means3D", "rotations", "scales", "opacities", "shs"):
        t = act[k]
        if perm is not None:
            t = t[perm]
        t = t.to(DEV).clone()
        if k == "opacities":
            t = t.reshape(-1)
        leaves[k] = t.requires_grad_(requires_grad)
    r, al, meta = pkg.rasterization(
        leaves["means3D"], leaves["rotations"], leaves["scales"], leaves["opacities"], leaves["shs"],
        sc["viewmat"].to(DEV)[None], sc["K"].to(DEV)[None], W, H, sh_degree=deg,
        backgrounds=None if bg is None else bg.to(DEV)[None])
    return r[0], al[0], meta, leaves


# BASELINE.json configs at their full sizes: A = configs[1] (1 M, 1080p), B = configs[2] (3 M, the 64 circle cameras —
# three of them here), C = configs[3] (6 M, 3840x2160: 32 400 tiles -> 15-bit tile key, 288 MB of records > L2)
FULL = {
    "A": dict(N=1_000_000, W=1920, H=1080, seed=0, cam="identity"),
    "A-asymK": dict(N=1_000_000, W=1920, H=1080, seed=0, cam="asymK"),
    "B-cam0": dict(N=3_000_000, W=1920, H=1080, seed=0, cam=("circle", 0)),
    "B-cam21": dict(N=3_000_000, W=1920, H=1080, seed=0, cam=("circle", 21)),
    "B-cam43": dict(N=3_000_000, W=1920, H=1080, seed=0, cam=("circle", 43)),
    "C": dict(N=6_000_000, W=3840, H=2160, seed=0, cam="identity"),
}
_scene_cache = {}


def _full_scene(oracle, name):
    cfg = FULL[name]
    key = (cfg["N"], cfg["W"], cfg["H"], cfg["seed"])
    if key not in _scene_cache:
        _scene_cache.clear()  # one full-size scene resident at a time
        sc = oracle.synthetic_scene(cfg["N"], cfg["W"], cfg["H"], seed=cfg["seed"])
        _scene_cache[key] = (sc, oracle.activate(sc))
    sc, act = _scene_cache[key]
    if isinstance(cfg["cam"], tuple):
        sc = dict(sc)
        sc["viewmat"] = oracle.circle_cameras(64)[cfg["cam"][1]]
    else:
        sc = apply_camera(sc, cfg["cam"], cfg["W"], cfg["H"])
    return sc, act, cfg


@pytest.mark.parametrize("name", list(FULL))
def test_full_size_binning_bit_exact_against_c_port(pkg, oracle, name):
    """tile keys / sort order / per-tile ranges at full size, from the GPU's own means2d / radii / depths."""
    from oracle import cpu_port
    cpu_port.set_threads(os.cpu_count() or 1)
    sc, act, cfg = _full_scene(oracle, name)
    N, W, H = cfg["N"], cfg["W"], cfg["H"]
    with torch.no_grad():
        _, _, meta, _ = _render(pkg, act, sc, W, H, DEG)
    lib, p = cpu_port._lib, cpu_port._p
    radii = meta["radii"][0].cpu().contiguous()
    means2d = meta["means2d"][0].cpu().contiguous()
    depths = meta["depths"][0].cpu().contiguous()
    tpg = meta["tiles_per_gauss"][0].cpu().contiguous()
    cum = torch.empty(N, dtype=torch.int64)
    M = int(lib.cpu_cumsum(p(tpg), C.c_int64(N), p(cum)))
    assert M == meta["flatten_ids"].numel() == int(tpg.sum())
    ids, flat = torch.empty(M, dtype=torch.int64), torch.empty(M, dtype=torch.int32)
    lib.cpu_bin_emit(p(means2d), p(radii), p(depths), p(cum), C.c_int64(N), C.c_int(W), C.c_int(H), p(ids), p(flat))
    tile_w, tile_h = (W + 15) // 16, (H + 15) // 16
    end_bit = 32 + int(math.floor(math.log2(tile_w * tile_h))) + 1
    tk, tv = torch.empty_like(ids), torch.empty_like(flat)
    lib.cpu_sort_pairs(p(ids), p(flat), p(tk), p(tv), C.c_int64(M), C.c_int(end_bit))
    offs = torch.empty(tile_h * tile_w, dtype=torch.int32)
    lib.cpu_tile_ranges(p(ids), C.c_int64(M), C.c_int(tile_h * tile_w), p(offs))
    print(f"{name}: N={N} M={M} tiles={tile_w}x{tile_h} visible={(radii > 0).all(-1).sum().item()}")
    assert torch.equal(meta["flatten_ids"].cpu(), flat)
    assert torch.equal(meta["isect_offsets"][0].cpu().reshape(-1), offs)


@pytest.mark.parametrize("name", list(FULL))
def test_full_size_render_and_gradients_match_c_port(pkg, oracle, name):
    """Image and parameter gradients at full size against the C port (fp32 like the GPU, so BOTH sides round):
    image max-abs <= 1e-4 except on <= 0.01 % threshold-tied pixels (measured 5e-6 .. 1e-5 of the pixels), themselves
    capped at 0.01 (measured <= 3.3e-3: one faint Gaussian's contribution); gradients per Gaussian (parity_util)."""
    from oracle import cpu_port
    cpu_port.set_threads(os.cpu_count() or 1)
    sc, act, cfg = _full_scene(oracle, name)
    W, H = cfg["W"], cfg["H"]
    v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1))
    ref = cpu_port.forward_backward(act, sc["viewmat"], sc["K"], W, H, DEG, v)
    r, al, meta, leaves = _render(pkg, act, sc, W, H, DEG, requires_grad=True)
    r.backward(v.to(DEV))
    # Gaussians whose integer radius differs by fp32 rounding of the projection change M; bounded at 0.01 %
    mism = (meta["radii"][0].cpu() != ref["fw"]["radii"]).any(-1).float().mean().item()
    print(f"{name}: M gpu {meta['flatten_ids'].numel()} / cpu {ref['M']}; radii mismatch fraction {mism:.2e}")
    assert mism <= 1e-4, mism
    image_close(r, ref["render"], None, max_marginal=1e-4, cap=0.01, what=f"{name}/render")
    image_close(al, ref["alpha"], None, max_marginal=1e-4, cap=0.01, what=f"{name}/alpha")
    for k, g in leaves.items():
        grad_rows(g.grad, ref["grads"][k].reshape(g.shape), f"{k}/{name}")


def test_config_a_structure_properties(pkg, scene_a):
    sc, act = scene_a
    with torch.no_grad():
        r, al, meta, _ = _render(pkg, act, sc, W_A, H_A, DEG)
    flat, offs = meta["flatten_ids"].long(), meta["isect_offsets"][0].reshape(-1).long()
    M = flat.numel()
    assert int(meta["tiles_per_gauss"].sum()) == M
    assert int(offs[0]) == 0 and bool((offs[1:] >= offs[:-1]).all()) and int(offs[-1]) <= M
    # tile id of every sorted position (from the ranges) and depth order inside each tile
    tile_of = torch.searchsorted(offs, torch.arange(M, device=DEV), right=True) - 1
    d = meta["depths"][0][flat]
    same_tile = tile_of[1:] == tile_of[:-1]
    assert bool((d[1:][same_tile] >= d[:-1][same_tile]).all())
    # ties inside a tile keep ascending Gaussian index (stable sort of emission order)
    tie = same_tile & (d[1:] == d[:-1])
    assert bool((flat[1:][tie] > flat[:-1][tie]).all())
    # every listed Gaussian is visible and its rectangle contains the tile
    assert bool((meta["radii"][0][flat] > 0).all())
    assert float(al.min()) >= 0.0 and float(al.max()) <= 1.0 + 1e-6
    assert bool(torch.isfinite(r).all())


def test_config_a_background_and_cotangent_linearity(pkg, scene_a):
    sc, act = scene_a
    bg = torch.tensor([0.25, 0.5, 0.75])
    with torch.no_grad():
        r0, a0, _, _ = _render(pkg, act, sc, W_A, H_A, DEG)
        r1, a1, _, _ = _render(pkg, act, sc, W_A, H_A, DEG, bg=bg)
    assert torch.equal(a0, a1)
    assert (r1 - r0 - (1.0 - a0) * bg.to(DEV)).abs().max().item() <= 1e-6
    # the backward pass is linear in the cotangent: g(v1 + 2 v2) == g(v1) + 2 g(v2)
    gen = torch.Generator().manual_seed(7)
    v1, v2 = torch.randn(H_A, W_A, 3, generator=gen).to(DEV), torch.randn(H_A, W_A, 3, generator=gen).to(DEV)
    grads = []
    for v in (v1, v2, v1 + 2.0 * v2):
        r, _, _, leaves = _render(pkg, act, sc, W_A, H_A, DEG, requires_grad=True)
        r.backward(v)
        grads.append({k: t.grad for k, t in leaves.items()})
    for k in grads[0]:
        want = grads[0][k] + 2.0 * grads[1][k]
        rel = ((grads[2][k] - want).norm() / want.norm()).item()
        assert rel <= 1e-4, f"{k}: {rel}"


def test_config_a_invariant_under_gaussian_permutation(pkg, scene_a):
    sc, act = scene_a
    perm = torch.randperm(N_A, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        r0, a0, m0, _ = _render(pkg, act, sc, W_A, H_A, DEG)
        r1, a1, m1, _ = _render(pkg, act, sc, W_A, H_A, DEG, perm=perm)
    assert m0["flatten_ids"].numel() == m1["flatten_ids"].numel()
    assert torch.equal(m0["isect_offsets"], m1["isect_offsets"])
    # same Gaussians per tile in the same depth order (ties may swap), hence the same image up to fp32 reassociation
    assert (r0 - r1).abs().max().item() <= 1e-4
    assert torch.equal(m0["radii"][0][perm.to(DEV)], m1["radii"][0])


def test_golden_fixture_on_gpu(pkg):
    """The committed fixture (tests/golden/make_golden.py: fp32 oracle forward, fp64 oracle autograd)."""
    gold = torch.load(GOLDEN)
    W, H, deg = gold["W"], gold["H"], gold["deg"]
    sc = gold["scene"]
    r, al, meta, leaves = _render(pkg, gold["act"], sc, W, H, deg, bg=gold["bg"], requires_grad=True)
    r.backward(gold["v_render"].to(DEV))
    if torch.equal(meta["radii"][0].cpu(), gold["radii"]):
        assert torch.equal(meta["flatten_ids"].cpu(), gold["flatten_ids"])
        assert torch.equal(meta["isect_offsets"][0].cpu(), gold["isect_offsets"])
    image_close(r, gold["render"], None, max_marginal=0.02, what="golden/render")
    for k, gref in gold["grads"].items():
        grad_rows(leaves[k].grad.reshape(gref.shape), gref, f"{k}/golden")


def test_more_than_65536_tiles_takes_the_64bit_key_path(pkg, oracle):
    """4352x4096 = 69 632 tiles: tile ids no longer fit the 16-bit split key, the backend must fall back to packed
    64-bit (tile, depth) keys on its own and still match the C port (image, ranges, gradients)."""
    from oracle import cpu_port
    cpu_port.set_threads(os.cpu_count() or 1)
    N, W, H = 6000, 4352, 4096
    assert ((W + 15) // 16) * ((H + 15) // 16) > 65536
    sc = oracle.synthetic_scene(N, W, H, seed=12)
    act = oracle.activate(sc)
    v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(2))
    ref = cpu_port.forward_backward(act, sc["viewmat"], sc["K"], W, H, DEG, v)
    r, al, meta, leaves = _render(pkg, act, sc, W, H, DEG, requires_grad=True)
    r.backward(v.to(DEV))
    if torch.equal(meta["radii"][0].cpu(), ref["fw"]["radii"]):  # same projection bits -> binning must be identical
        assert torch.equal(meta["flatten_ids"].cpu(), ref["fw"]["flatten_ids"])
        assert torch.equal(meta["isect_offsets"][0].cpu(), ref["fw"]["isect_offsets"])
    image_close(r, ref["render"], None, max_marginal=1e-4, cap=0.01, what="69632-tiles/render")
    for k, g in leaves.items():
        grad_rows(g.grad, ref["grads"][k].reshape(g.shape), f"{k}/69632-tiles")


def test_ten_million_gaussians_structure(pkg, oracle):
    """The reference's population cap (training_config.py:22, max_gaussians = 1e7): sizes, 32-bit indices and the
    sort scratch hold at 10 M Gaussians / ~65 M intersections; checked through size-independent properties."""
    N, W, H = 10_000_000, 1920, 1080
    sc = oracle.synthetic_scene(N, W, H, seed=13)
    act = oracle.activate(sc)
    with torch.no_grad():
        r, al, meta, _ = _render(pkg, act, sc, W, H, DEG, bg=torch.tensor([0.1, 0.2, 0.3]))
    flat, offs = meta["flatten_ids"], meta["isect_offsets"][0].reshape(-1).long()
    tpg = meta["tiles_per_gauss"][0]
    M = flat.numel()
    assert M == int(tpg.sum().item()) and M > 40_000_000
    assert int(flat.min()) >= 0 and int(flat.max()) < N
    assert bool((offs[1:] >= offs[:-1]).all()) and int(offs[0]) == 0 and int(offs[-1]) <= M
    # every Gaussian appears exactly tiles_per_gauss times in the sorted list
    assert torch.equal(torch.bincount(flat.long(), minlength=N), tpg.long())
    # depth order inside a few tiles spread over the image
    depths = meta["depths"][0]
    ends = torch.cat([offs[1:], torch.tensor([M], device=offs.device)])
    for t in (0, 1234, 4050, 8099):
        d = depths[flat[offs[t]: ends[t]].long()]
        assert bool((d[1:] >= d[:-1]).all()), t
    assert torch.isfinite(r).all() and float(al.min()) >= 0.0 and float(al.max()) <= 1.0 + 1e-6
