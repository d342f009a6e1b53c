"""BASELINE.json configs[4] (config D): densification-heavy run — the population grows 100 k -> >= 4 M by repeated
densify_and_prune, with full optimiser iterations (render fwd+bwd at 1920x1080, fused loss, Adam, grad-stat
accumulation; every view re-bins and re-sorts) in between.  Prints one JSON line; diagnostic, not bench.py."""
import importlib, json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d-gaussian-splatting_b200")
from oracle import torch_oracle as orc
dev = "cuda:0"
W, H, deg = 1920, 1080, 3
sc = orc.synthetic_scene(100_000, W, H, seed=0)
st = pkg.GaussianStore.from_tensors(*[sc[k].to(dev) for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")])
opt = pkg.StoreAdam(st, {"xyz": 1.6e-4, "f_dc": 0.0025, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001})
K = sc["K"].to(dev)
cams = orc.circle_cameras(64).to(dev)
gt = torch.rand(H, W, 3, device=dev)
bg = torch.zeros(3, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
rows, it = [], 0
torch.manual_seed(0)
while True:
    n = len(st)
    steps = 6
    e0, e1 = ev(), ev()
    for i in range(2 + steps):
        if i == 2:
            e0.record()
        st.zero_grad()
        vp = {"world_view_transform": cams[it % 64], "image_width": W, "image_height": H}
        res = st.render(K, vp, bg_color=bg, sh_degree=deg)
        pkg.l1_ssim_loss(res["render"], gt)["loss"].backward()
        opt.step()
        pkg.densify_stats(st.xyz_gradient_accum, st.xyz_gradient_count, st.xyz.grad)
        it += 1
    e1.record()
    d0, d1 = ev(), ev()
    d0.record()
    # threshold 0: every Gaussian above the opacity floor is cloned or split (population roughly doubles)
    stats = st.densify_and_prune(0.0, 0.005, extent=4.0, max_screen_size=None, optimizer=opt, max_gaussians=4_200_000)
    d1.record()
    torch.cuda.synchronize()
    rows.append({"N": n, "ms_per_iteration": e0.elapsed_time(e1) / steps, "densify_ms": d0.elapsed_time(d1), **stats})
    if n >= 4_000_000 or stats["n_after"] == n:
        break
print(json.dumps({"config": "D: 100k -> 4M Gaussians, 1920x1080, one view per optimiser iteration", "rows": rows}))
