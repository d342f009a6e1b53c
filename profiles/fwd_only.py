"""Runs a few forward passes of config A (for ncu launch lists of the binning/sort kernels)."""
import importlib, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d-gaussian-splatting_b200")
from oracle import torch_oracle as orc
N, W, H = 1_000_000, 1920, 1080
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
a = {k: v.cuda() for k, v in act.items()}
cam = pkg.make_cam(sc["viewmat"], sc["K"], W, H)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    b = pkg.project_and_bin(a["means3D"], a["rotations"], a["scales"], a["opacities"].squeeze(-1).contiguous(), a["shs"], 3, cam, W, H)
torch.cuda.synchronize()
print("M", b["M"])
