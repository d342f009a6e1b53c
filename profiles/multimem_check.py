"""torchrun-launched: this library's multimem all-reduce (csrc/collective.cu) against NCCL on the 236 MB flat gradient of
config A — same sums (fp32 reassociation only), bit-identical replicas, and the time of both.  Prints one JSON line."""
import importlib, json, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
pkg = importlib.import_module("3d-gaussian-splatting_b200")
n = 59_000_000
out = {"world": world, "floats": n}
try:
    sb = pkg.SymmetricBuffer(n, dev)
except Exception as e:
    out["unavailable"] = f"{type(e).__name__}: {e}"[:300]
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()
    sys.exit(0)
g = torch.Generator(device=dev).manual_seed(rank)
src = torch.randn(n, generator=g, device=dev)
ref = src.clone()
dist.all_reduce(ref)
ref.mul_(1.0 / world)
sb.tensor.copy_(src)
sb.allreduce_(1.0 / world)
torch.cuda.synchronize()
err = ((sb.tensor - ref).abs().max() / ref.abs().max()).item()
bits = sb.tensor.view(torch.int32)
hi, lo = bits.clone(), bits.clone()
dist.all_reduce(hi, op=dist.ReduceOp.MAX); dist.all_reduce(lo, op=dist.ReduceOp.MIN)
out["max_err_over_max"] = err
out["replicas_bit_identical"] = bool(torch.equal(hi, lo))
def timed(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / it], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
buf = src.clone()
out["nccl_allreduce_ms"] = timed(lambda: dist.all_reduce(buf))
out["two_barriers_ms"] = timed(lambda: (sb.handle.barrier(channel=0), sb.handle.barrier(channel=1)))
for unroll in (4, 8, 16):
    for ctas in (32, 64, 148, 296):
        sb.max_ctas = ctas | (unroll << 16)
        out[f"multimem_ms_u{unroll}_c{ctas}"] = round(timed(lambda: sb.allreduce_(1.0)), 4)
sb.max_ctas = 0
out["ok"] = bool(err <= 1e-5 and out["replicas_bit_identical"])
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
