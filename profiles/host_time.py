"""Host-side enqueue time per view (no device sync inside the measured calls except the library's own M read-back):
tells whether the Python orchestration or the GPU bounds a step.  Diagnostic only."""
import importlib, os, sys, time, torch
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
def run(n):
    tf = tb = 0.0
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n):
        a = time.perf_counter()
        res = be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
        b = time.perf_counter()
        res["render"].backward(v)
        c = time.perf_counter()
        tf += b - a; tb += c - b
    torch.cuda.synchronize(); t1 = time.perf_counter()
    return tf / n * 1e3, tb / n * 1e3, (t1 - t0) / n * 1e3
run(10)
print("host ms/view: forward call %.3f, backward call %.3f; wall per view %.3f" % run(40))
# the forward call includes waiting for M (i.e. for the GPU to drain up to tile_prep); measure pure CPU cost with an idle GPU
tf = 0.0
for _ in range(20):
    torch.cuda.synchronize(); a = time.perf_counter()
    with torch.no_grad():
        res = be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
    tf += time.perf_counter() - a
print("forward call on an idle GPU (includes ~0.25 ms of GPU work before M is known): %.3f ms" % (tf / 20 * 1e3))
import subprocess; print(subprocess.run("lscpu | grep -E 'Model name|MHz' | head -3", shell=True, capture_output=True, text=True).stdout)
