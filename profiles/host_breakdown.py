"""Where the host time of a drop-in view goes (perf_counter wrappers around the package's own layers; one stream, GPU
busy as in a training step).  Diagnostic only."""
import importlib, os, sys, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("3d-gaussian-splatting_b200")
rz = importlib.import_module("3d-gaussian-splatting_b200.rasterization")
orc = importlib.import_module("3d-gaussian-splatting_b200.synthetic")
N, W, H, deg = bench.CONFIGS["A"]
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
dev = torch.device("cuda", 0)
leaves = {k: act[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
acc = {}
def timed(name, fn):
    def w(*a, **k):
        t = time.perf_counter()
        try:
            return fn(*a, **k)
        finally:
            acc[name] = acc.get(name, 0.0) + time.perf_counter() - t
    return w
rz.make_cam = timed("make_cam", rz.make_cam)
rz.project_and_bin = timed("project_and_bin", rz.project_and_bin)
bk = importlib.import_module("3d-gaussian-splatting_b200.backend")
rz._render_one = bk._render_one = timed("_render_one", rz._render_one)
rz.RasterizeSplats.forward = staticmethod(timed("Function.forward", rz.RasterizeSplats.forward))
rz.RasterizeSplats.backward = staticmethod(timed("Function.backward", rz.RasterizeSplats.backward))
torch.cuda.Event.synchronize = timed("Event.synchronize (M wait)", torch.cuda.Event.synchronize)
be = pkg.B200Backend(sc["K"].to(dev))
v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1)).to(dev)
cams = [c.to(dev) for c in orc.circle_cameras(64)]
bg = torch.zeros(3, device=dev)
rwd = timed("render_with_depth", be.render_with_depth)
def run(n):
    for i in range(n):
        if i % 4 == 0:
            for p in leaves.values(): p.grad = None
        vp = {"world_view_transform": cams[i % 64], "image_width": W, "image_height": H}
        res = rwd(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
        t = time.perf_counter()
        res["render"].backward(v)
        acc["tensor.backward()"] = acc.get("tensor.backward()", 0.0) + time.perf_counter() - t
run(12); torch.cuda.synchronize(); acc.clear()
t0 = time.perf_counter(); n = 200; run(n); t1 = time.perf_counter(); torch.cuda.synchronize()
print("host loop %.1f us per view (GPU-paced where the host waits for M)" % ((t1 - t0) / n * 1e6))
for k, val in acc.items():
    print("%-28s %8.1f us per view" % (k, val / n * 1e6))
