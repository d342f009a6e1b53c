"""Per-stage CUDA-event times of config A/C views through the public backend call (tuning aid for A/B runs of
kernel variants; bench.py remains the source of reported numbers).  usage: stage_times.py [views] [config]"""
import importlib, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("3d-gaussian-splatting_b200")
from oracle import torch_oracle as orc
views = int(sys.argv[1]) if len(sys.argv) > 1 else 40
cfg = sys.argv[2] if len(sys.argv) > 2 else "A"
N, W, H, deg = bench.CONFIGS[cfg]
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
dev = "cuda:0"
leaves = {k: act[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
be = pkg.B200Backend(sc["K"].to(dev))
v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1)).to(dev)
cams = orc.circle_cameras(64)
bg = torch.zeros(3, device=dev)
def run(n, k0=0):
    for i in range(n):
        for p in leaves.values():
            p.grad = None
        vp = {"world_view_transform": cams[(k0 + i) % 64].to(dev), "image_width": W, "image_height": H}
        res = be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
        res["render"].backward(v)
run(5)
torch.cuda.synchronize()
timer = bench.StageTimer(torch)
pkg._lib.stage_timer = timer
timer.enabled = True
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(views, 5); e1.record()
torch.cuda.synchronize()
timer.enabled = False
tot = timer.totals_ms(); cnt = timer.counts()
print("%s ms/view %.4f | " % (os.environ.get("TAG", ""), e0.elapsed_time(e1) / views) +
      " ".join("%s=%.4f" % (k, tot[k] / max(cnt[k], 1)) for k in tot))
