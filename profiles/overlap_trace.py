"""Unprofiled timeline of two-stream steps from CUDA events (the package's per-stage hooks): where, relative to the
step's start, each stage of each view began and ended on the GPU, and when the HOST issued it.  Diagnostic only."""
import importlib, os, sys, time, torch
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
class Rec:
    def __init__(self): self.on = False; self.open = {}; self.rows = []; self.view = 0
    def begin(self, name):
        if not self.on: return
        e = torch.cuda.Event(enable_timing=True); e.record(); self.open[name] = (e, time.perf_counter())
    def end(self, name):
        if not self.on: return
        e = torch.cuda.Event(enable_timing=True); e.record()
        b, t = self.open.pop(name); self.rows.append((self.view, name, b, e, t))
rec = Rec(); pkg._lib.stage_timer = rec
def step(s):
    for p in params: p.grad = None
    vs.begin()
    for i in range(4):
        with vs.view(i):
            rec.view = 4 * s + i
            vp = {"world_view_transform": cams[(4 * s + i) % 64], "image_width": W, "image_height": H}
            res = be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
            res["render"].backward(v)
    vs.join()
for s in range(5): step(s)
torch.cuda.synchronize()
base = torch.cuda.Event(enable_timing=True); base.record(); t_base = time.perf_counter(); rec.on = True
for s in range(5, 8): step(s)
end = torch.cuda.Event(enable_timing=True); end.record(); torch.cuda.synchronize(); rec.on = False
print(f"streams {nstreams} FRONT_PRIORITY {os.environ.get('B200_FRONT_PRIORITY', '0')}: 3 steps (12 views) in {base.elapsed_time(end):.3f} ms (with stage events)")
print("view stage            host_issue_ms  gpu_begin_ms  gpu_end_ms  dur_ms")
for view, name, b, e, t in rec.rows:
    if view < 22 or view > 27: continue
    print(f"{view:4d} {name:16s} {1e3*(t - t_base):12.3f} {base.elapsed_time(b):13.3f} {base.elapsed_time(e):11.3f} {b.elapsed_time(e):7.3f}")
