"""All-reduce timing of the 59 N-float gradient buffer (N = 1 M) — diagnostic for the scaling leg."""
import os, torch, torch.distributed as dist
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
for n in (59_000_000, 48_000_000, 4_000_000):
    x = torch.ones(n, device="cuda")
    for _ in range(5): dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    if rank == 0:
        print(f"{os.environ.get('TAG','default')}: all_reduce {n*4/1e6:.0f} MB x{world}: {ms:.3f} ms  algbw {n*4/ms/1e6:.0f} GB/s busbw {n*4/ms/1e6*2*(world-1)/world:.0f} GB/s", flush=True)
dist.destroy_process_group()
