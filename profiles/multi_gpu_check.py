"""On-hardware multi-GPU correctness check.  Launch:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        profiles/multi_gpu_check.py [--n 1000000] [--views-per-rank 2]

Asserts, with the REAL rasterizer over NCCL:
  1. the N-rank reduced gradient (views sharded round-robin, `allreduce_flat` of the store's flat buffer; and the drop-in
     path's `allreduce_gradients` over the five attribute tensors) equals the 1-rank gradient of the mean loss over the same
     N*V views — per-Gaussian criterion of tests/parity_util.py (>= 99.9 % rows within 1e-3, whole tensor <= 1e-3; fp32
     atomics make the two summation orders differ, so not bit-exact);
  2. after one identical StoreAdam step (with the fused densification statistic) `store.flat`, both Adam moments and the
     statistic buffers are BIT-identical on every rank.
Prints one JSON line on rank 0.
"""
import argparse
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def rows_ok(got, ref, rtol=1e-3, floor=1e-2):
    n = ref.shape[0]
    g, r = got.reshape(n, -1).double(), ref.reshape(n, -1).double()
    scale = r.abs().amax().clamp_min(1e-300)
    tol = rtol * torch.maximum(r.abs().amax(1), floor * scale)
    err = (g - r).abs().amax(1)
    return float((err <= tol).double().mean()), float((g - r).norm() / r.norm().clamp_min(1e-300)), float(err.amax() / scale)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", dest="n", type=int, default=1_000_000)  # not "--n": torchrun's own parser would claim it as an abbreviation
    ap.add_argument("--views-per-rank", type=int, default=2)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pkg = importlib.import_module("3d-gaussian-splatting_b200")
    syn = importlib.import_module("3d-gaussian-splatting_b200.synthetic")
    N, W, H, deg = a.n, a.width, a.height, 3
    sc = syn.synthetic_scene(N, W, H, seed=0)
    cams = syn.circle_cameras(64)
    V = a.views_per_rank * world
    K = sc["K"].to(dev)
    bg = torch.zeros(3, device=dev)

    def cotangent(v):
        return torch.randn(H, W, 3, generator=torch.Generator().manual_seed(100 + v)).to(dev)

    def make_store():
        return pkg.GaussianStore.from_tensors(*(sc[k].to(dev) for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")))

    def render_views(store, views):
        store.zero_grad()
        for v in views:
            vp = {"world_view_transform": cams[v].to(dev), "image_width": W, "image_height": H}
            res = store.render(K, vp, bg_color=bg, sh_degree=deg)
            res["render"].backward(cotangent(v))

    out = {"world": world, "N": N, "views": V}
    # ---- store path: sharded + one flat all-reduce vs all views on this rank
    st = make_store()
    render_views(st, pkg.shard_views(V, rank, world))
    pkg.allreduce_flat(st.flat_grad, 1.0 / V)
    ref = make_store()
    render_views(ref, list(range(V)))
    ref.flat_grad.mul_(1.0 / V)
    worst = [1.0, 0.0, 0.0]
    for name in ("xyz", "scaling", "rotation", "opacity", "features"):
        f, l2, mx = rows_ok(getattr(st, name).grad, getattr(ref, name).grad)
        out[f"store/{name}"] = {"rows_ok": f, "l2_rel": l2, "max_err_over_scale": mx}
        worst = [min(worst[0], f), max(worst[1], l2), max(worst[2], mx)]
    # ---- drop-in path: five attribute tensors reduced back to back
    leaves = {k: v.to(dev).clone().requires_grad_(True) for k, v in syn.activate(sc).items()}
    be = pkg.B200Backend(K)
    for v in pkg.shard_views(V, rank, world):
        vp = {"world_view_transform": cams[v].to(dev), "image_width": W, "image_height": H}
        be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)["render"].backward(cotangent(v))
    pkg.allreduce_gradients(list(leaves.values()), V)
    leaves1 = {k: v.to(dev).clone().requires_grad_(True) for k, v in syn.activate(sc).items()}
    for v in range(V):
        vp = {"world_view_transform": cams[v].to(dev), "image_width": W, "image_height": H}
        be.render_with_depth(leaves1, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)["render"].backward(cotangent(v))
    for k in leaves:
        f, l2, mx = rows_ok(leaves[k].grad, leaves1[k].grad / V)
        out[f"dropin/{k}"] = {"rows_ok": f, "l2_rel": l2, "max_err_over_scale": mx}
        worst = [min(worst[0], f), max(worst[1], l2), max(worst[2], mx)]
    # ---- replicas after an identical optimiser step
    lrs = {"xyz": 1.6e-4, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
    opt = pkg.StoreAdam(st, lrs)
    opt.step(densify_stats=True)
    identical = True
    for t in (st.flat, opt.exp_avg, opt.exp_avg_sq, st.xyz_gradient_accum.reshape(-1), st.xyz_gradient_count.reshape(-1)):
        bits = t.view(torch.int32)
        hi, lo = bits.clone(), bits.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        identical = identical and bool(torch.equal(hi, lo))
    out["replicas_bit_identical_after_step"] = identical
    flags = torch.tensor([worst[0], -worst[1], -worst[2], float(identical)], device=dev, dtype=torch.float64)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    out["all_ranks"] = {"min_rows_ok": float(flags[0]), "max_l2_rel": -float(flags[1]), "max_err_over_scale": -float(flags[2]),
                        "bit_identical": bool(flags[3] == 1.0)}
    ok = flags[0] >= 0.999 and -flags[1] <= 1e-3 and flags[3] == 1.0
    out["ok"] = bool(ok)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
