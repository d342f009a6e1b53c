"""GPU timeline of config-A steps with the views on alternating streams (streams.ViewStreams), torch.profiler / CUPTI:
per-kernel durations when views overlap, union-busy time and the launch sequence with stream ids.  Diagnostic only."""
import importlib, os, sys, torch
from torch.profiler import profile, ProfilerActivity
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("3d-gaussian-splatting_b200")
orc = importlib.import_module("3d-gaussian-splatting_b200.synthetic")
N, W, H, deg = bench.CONFIGS["A"]
nstreams = int(sys.argv[1]) if len(sys.argv) > 1 else 2
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
dev = torch.device("cuda", 0)
leaves = {k: act[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
params = list(leaves.values())
be = pkg.B200Backend(sc["K"].to(dev))
v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1)).to(dev)
cams = [c.to(dev) for c in orc.circle_cameras(64)]
bg = torch.zeros(3, device=dev)
vs = pkg.ViewStreams(dev, nstreams, params)
def step(s):
    for p in params: p.grad = None
    vs.begin()
    for i in range(4):
        with vs.view(i):
            vp = {"world_view_transform": cams[(4 * s + i) % 64], "image_width": W, "image_height": H}
            res = be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
            res["render"].backward(v)
    vs.join()
for s in range(4): step(s)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for s in range(4, 7): step(s)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0, t1 = ev[0].time_range.start, max(e.time_range.end for e in ev)
# union of busy intervals
busy = 0.0; cur_s, cur_e = None, None
for e in ev:
    s_, e_ = e.time_range.start, e.time_range.end
    if cur_e is None or s_ > cur_e:
        if cur_e is not None: busy += cur_e - cur_s
        cur_s, cur_e = s_, e_
    else:
        cur_e = max(cur_e, e_)
busy += cur_e - cur_s
tot = sum(e.time_range.end - e.time_range.start for e in ev)
print(f"streams {nstreams} FRONT_PRIORITY {os.environ.get('B200_FRONT_PRIORITY', '0')}: span {(t1-t0)/1e3:.3f} ms for 12 views, union busy {busy/1e3:.3f} ms, sum of kernel durations {tot/1e3:.3f} ms")
agg = {}
for e in ev:
    a = agg.setdefault(e.name[:58], [0, 0.0]); a[0] += 1; a[1] += e.time_range.end - e.time_range.start
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:16]:
    print(f"{t/n:9.1f} us x{n:3d}  {k}")
if os.environ.get("SEQ"):
    def sid(e):
        for attr in ("cuda_stream", "stream", "device_resource_id"):
            if hasattr(e, attr): return getattr(e, attr)
        return -1
    k0 = len(ev) // 3
    for e in ev[k0:k0 + int(os.environ["SEQ"])]:
        print(f"  +{(e.time_range.start - t0)/1e3:8.3f} .. {(e.time_range.end - t0)/1e3:8.3f} ms  s{sid(e)}  {e.name[:60]}")
