"""Forward projection kernel timed alone at config A (tuning aid: A/B of build variants via profiles/gpu_variants-like loops).
usage: proj_probe.py [reps]"""
import importlib, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("3d-gaussian-splatting_b200")
ras = importlib.import_module("3d-gaussian-splatting_b200.rasterization")
from oracle import torch_oracle as orc
lib, _ptr, check = ras.lib, ras._ptr, ras.check
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
N, W, H, deg = bench.CONFIGS["A"]
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
dev = torch.device("cuda:0")
p = {k: act[k].to(dev).contiguous() for k in ("means3D", "scales", "rotations", "opacities", "shs")}
cam = ras.make_cam(sc["viewmat"].to(dev), sc["K"].to(dev), W, H)
tw, th = (W + 15) // 16, (H + 15) // 16
radii = torch.empty(N, 2, dtype=torch.int32, device=dev); splats = torch.empty(N, 12, device=dev)
tpg = torch.empty(N, dtype=torch.int32, device=dev)
nb = int(lib.b200_project_blocks(N)); bs = torch.empty(nb, dtype=torch.int64, device=dev)
dk = torch.empty(N, dtype=torch.int32, device=dev); gi = torch.empty(N, dtype=torch.int32, device=dev)
grid = torch.empty((tw + 1) * (th + 1), dtype=torch.int32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
def call():
    check(lib.b200_project_sh_fwd(_ptr(p["means3D"]), _ptr(p["rotations"]), _ptr(p["scales"]), _ptr(p["opacities"].reshape(-1)), _ptr(p["shs"]),
                                  16, deg, N, None, _ptr(cam), W, H, _ptr(radii), _ptr(splats), _ptr(tpg), (None if os.environ.get("NO_BS") else _ptr(bs)), _ptr(dk), _ptr(gi), _ptr(grid), None, None, None, st), "proj")
for _ in range(5): call()
tot = 0.0
for _ in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); call(); e1.record(); torch.cuda.synchronize()
    tot += e0.elapsed_time(e1)
print("%s project_sh_fwd alone (L2 flushed): %.4f ms  (memset of the tile grid included)  N_vis=%d M=%d" % (os.environ.get("TAG", ""), tot / reps, int((radii[:, 0] > 0).sum()), int(tpg.sum())))
