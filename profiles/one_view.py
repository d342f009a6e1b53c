"""Runs a few forward+backward views of a config (default A) through the public backend call — the
program ncu is pointed at for launch lists and --set full captures (never a source of bench numbers)."""
import importlib, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("3d-gaussian-splatting_b200")
from oracle import torch_oracle as orc
cfg = sys.argv[2] if len(sys.argv) > 2 else "A"
N, W, H, deg = bench.CONFIGS[cfg]
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
dev = "cuda:0"
leaves = {k: act[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
be = pkg.B200Backend(sc["K"].to(dev))
v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1)).to(dev)
vp = {"world_view_transform": sc["viewmat"].to(dev), "image_width": W, "image_height": H}
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    res = be.render_with_depth(leaves, vp, bg_color=torch.zeros(3, device=dev), render_mode="RGB", device="cuda", sh_degree=deg)
    res["render"].backward(v)
torch.cuda.synchronize()
print("done", cfg, N, W, H)
