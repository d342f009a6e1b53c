"""cProfile of the host side of config-A views through the public backend call (where do the ~0.65 ms of Python per view
go?).  Diagnostic only."""
import cProfile, importlib, os, pstats, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("3d-gaussian-splatting_b200")
orc = importlib.import_module("3d-gaussian-splatting_b200.synthetic")
N, W, H, deg = bench.CONFIGS["A"]
sc = orc.synthetic_scene(N, W, H, seed=0); act = orc.activate(sc)
dev = torch.device("cuda", 0)
leaves = {k: act[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
be = pkg.B200Backend(sc["K"].to(dev))
v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1)).to(dev)
cams = [c.to(dev) for c in orc.circle_cameras(64)]
bg = torch.zeros(3, device=dev)
def run(n):
    for i in range(n):
        for p in leaves.values(): p.grad = None
        vp = {"world_view_transform": cams[i % 64], "image_width": W, "image_height": H}
        res = be.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
        res["render"].backward(v)
run(10); torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable(); run(100); pr.disable(); torch.cuda.synchronize()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(28)
st.sort_stats("cumulative").print_stats(22)
