"""GPU timeline of a few config-A views through the public backend call (torch.profiler / CUPTI): busy time,
idle gaps and which kernel each gap follows.  Diagnostic only — never a source of bench numbers."""
import importlib, os, sys, json, torch
from torch.profiler import profile, ProfilerActivity
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("3d-gaussian-splatting_b200")
from oracle import torch_oracle as orc
cfg = sys.argv[1] if len(sys.argv) > 1 else "A"
N, W, H, deg = bench.CONFIGS[cfg]
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
dev = "cuda:0"
leaves = {k: act[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
be = pkg.B200Backend(sc["K"].to(dev))
v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1)).to(dev)
vp = {"world_view_transform": sc["viewmat"].to(dev), "image_width": W, "image_height": H}
bg = torch.zeros(3, device=dev)
mode = sys.argv[2] if len(sys.argv) > 2 else "backend"
store = pkg.GaussianStore.from_tensors(*[sc[k].to(dev) for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")])
Kd = sc["K"].to(dev)
def step():
    if mode == "store":
        store.zero_grad()
        for _ in range(4):
            res = store.render(Kd, vp, bg_color=bg, render_mode="RGB", sh_degree=deg)
            res["render"].backward(v)
        return
    for p in leaves.values(): p.grad = None
    for _ in range(4):
        res = be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
        res["render"].backward(v)
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2): step()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0, t1 = ev[0].time_range.start, max(e.time_range.end for e in ev)
busy = sum(e.time_range.end - e.time_range.start for e in ev)
print(f"span {(t1-t0)/1e3:.3f} ms, busy {busy/1e3:.3f} ms, idle {(t1-t0-busy)/1e3:.3f} ms, kernels {len(ev)}  (8 views)")
agg = {}
for e in ev:
    a = agg.setdefault(e.name[:60], [0, 0.0]); a[0] += 1; a[1] += e.time_range.end - e.time_range.start
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:14]:
    print(f"{t/n:9.1f} us x{n:3d}  {k}")
gaps = {}
for a, b in zip(ev, ev[1:]):
    g = b.time_range.start - a.time_range.end
    k = (a.name[:40], b.name[:40])
    s = gaps.setdefault(k, [0, 0.0]); s[0] += 1; s[1] += max(g, 0)
for k, (n, g) in sorted(gaps.items(), key=lambda x: -x[1][1])[:8]:
    print(f"{g/n:8.1f} us x{n:3d}  after {k[0]:40s} before {k[1]}")
if os.environ.get("SEQ"):  # launch order of the first ~2 views with start offsets, durations and the gap in front
    prev = ev[0].time_range.start
    for e in ev[:int(os.environ["SEQ"])]:
        print(f"  +{(e.time_range.start - t0)/1e3:8.3f} ms  gap {max(e.time_range.start - prev, 0):7.1f} us  dur {e.time_range.end - e.time_range.start:7.1f} us  {e.name[:70]}")
        prev = e.time_range.end
