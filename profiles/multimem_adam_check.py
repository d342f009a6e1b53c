"""torchrun-launched check + timing of the fused multi-GPU optimiser step (csrc/collective.cu):
  A: NCCL all-reduce of the flat gradient, then StoreAdam.step(densify_stats=True) on every rank   (round-1 path)
  B: store.make_symmetric(); StoreAdam(sharded=True).step_allreduce(...)  — reduce-scatter + Adam on the slice +
     parameter all-gather in ONE multimem kernel, sharded moments
Asserts after 3 steps: parameters of B == A (same arithmetic; the in-switch sum may reassociate, so a tight allclose plus
the bit-equality count is reported), replicas of B bit-identical, statistic buffers equal, B's moments (after
gather_moments) == A's; then a densify_and_prune on both and one more step.  Prints one JSON line with the timings."""
import importlib, json, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
pkg = importlib.import_module("3d-gaussian-splatting_b200")
syn = importlib.import_module("3d-gaussian-splatting_b200.synthetic")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
sc = syn.synthetic_scene(N, 1920, 1080, seed=0)
mk = lambda: pkg.GaussianStore.from_tensors(*(sc[k].to(dev) for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")))  # noqa: E731
lrs = {"xyz": 1.6e-4, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
out = {"world": world, "N": N}
A = mk()
oA = pkg.StoreAdam(A, lrs, scale_ceiling=0.5)
try:
    B = mk().make_symmetric()
    oB = pkg.StoreAdam(B, lrs, scale_ceiling=0.5, sharded=True)
except Exception as e:
    out["unavailable"] = f"{type(e).__name__}: {e}"[:300]
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()
    sys.exit(0)
gen = torch.Generator(device=dev).manual_seed(100 + rank)

def grads(scale=1.0):
    g = torch.randn(A.flat.numel(), generator=gen, device=dev) * scale
    for st in (A, B):
        st.zero_grad()
        st.flat_grad.copy_(g)

def same_everywhere(t):
    bits = t.contiguous().view(torch.int32)
    hi, lo = bits.clone(), bits.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX); dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    return bool(torch.equal(hi, lo))

ok = True
for it in range(3):
    grads(10.0 ** (it - 2))
    oA.step_allreduce(1.0 / world, densify_stats=True)
    oB.step_allreduce(1.0 / world, densify_stats=True)
torch.cuda.synchronize()
n = A.flat.numel()
out["params_max_abs_diff"] = float((A.flat - B.flat[:n]).abs().max())
out["params_bit_equal_fraction"] = float((A.flat.view(torch.int32) == B.flat[:n].view(torch.int32)).float().mean())
out["replicas_bit_identical"] = same_everywhere(B.flat) and same_everywhere(B.xyz_gradient_accum) and same_everywhere(B.xyz_gradient_count)
out["stats_equal"] = bool(torch.allclose(A.xyz_gradient_accum, B.xyz_gradient_accum, rtol=1e-6, atol=1e-7) and torch.equal(A.xyz_gradient_count, B.xyz_gradient_count))
oB.gather_moments()
out["moments_max_abs_diff"] = float(max((oA.exp_avg - oB.exp_avg[:n]).abs().max(), (oA.exp_avg_sq - oB.exp_avg_sq[:n]).abs().max()))
# (the in-switch N-way sum and NCCL's reduction associate differently for N > 2: gradients differ in the last bit, so a
# small fraction of the parameters differs by 1-2 ulp; the replicas of either path are bit-identical among themselves)
ok = ok and out["params_max_abs_diff"] <= 1e-5 and out["params_bit_equal_fraction"] >= 0.97 and out["replicas_bit_identical"] \
    and out["stats_equal"] and out["moments_max_abs_diff"] <= 1e-6
# population change on both, then one more step
for st, op in ((A, oA), (B, oB)):
    torch.manual_seed(7)
    st.xyz_gradient_accum.fill_(1e-3); st.xyz_gradient_count.fill_(1.0)
    stats = st.densify_and_prune(2e-4, 0.05, 0.2, None, optimizer=op, max_gaussians=int(1.05 * N))
out["n_after_densify"] = len(B)
grads(0.1)
oA.step_allreduce(1.0 / world); oB.step_allreduce(1.0 / world)
torch.cuda.synchronize()
n = A.flat.numel()
out["after_densify_params_max_abs_diff"] = float((A.flat - B.flat[:n]).abs().max())
ok = ok and len(A) == len(B) and out["after_densify_params_max_abs_diff"] <= 1e-5 and same_everywhere(B.flat)

def timed(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / it], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t), 4)
out["nccl_allreduce_plus_adam_ms"] = timed(lambda: oA.step_allreduce(1.0 / world, densify_stats=True))
out["fused_multimem_step_ms"] = timed(lambda: oB.step_allreduce(1.0 / world, densify_stats=True))
out["multimem_allreduce_only_ms"] = timed(lambda: B.allreduce_grads(1.0))
out["nccl_allreduce_only_ms"] = timed(lambda: A.allreduce_grads(1.0))
out["ok"] = bool(ok)
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
sys.exit(0 if ok else 1)
