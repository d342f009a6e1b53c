"""Diagnostic: the same view rendered + back-propagated on cuda:0, cuda:1, cuda:0 and cuda:1 again in ONE process; prints
how far the gradients of two runs are apart on the same device (atomic-order noise) and across devices."""
import importlib, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("3d-gaussian-splatting_b200")
orc = importlib.import_module("3d-gaussian-splatting_b200.synthetic")
N, W, H = 6000, 640, 360
sc = orc.synthetic_scene(N, W, H, seed=3); act = orc.activate(sc)
outs = []
for d in ("cuda:0", "cuda:1", "cuda:0", "cuda:1"):
    leaves = {k: act[k].to(d).clone().requires_grad_(True) for k in act}
    be = pkg.B200Backend(sc["K"].to(d))
    vp = {"world_view_transform": sc["viewmat"].to(d), "image_width": W, "image_height": H}
    res = be.render_with_depth(leaves, vp, bg_color=torch.zeros(3, device=d), render_mode="RGB", device=d, sh_degree=3)
    res["render"].sum().backward()
    torch.cuda.synchronize(d)
    outs.append({k: v.grad.cpu() for k, v in leaves.items()} | {"render": res["render"].detach().cpu()})
for a, b in ((0, 2), (1, 3), (0, 1), (2, 3)):
    print(f"runs {a} vs {b}:", " ".join(
        "%s maxabs %.3g rel-to-max %.3g allclose(1e-4,1e-6)=%s" % (k, (outs[a][k] - outs[b][k]).abs().max().item(),
        ((outs[a][k] - outs[b][k]).abs().max() / outs[a][k].abs().max().clamp_min(1e-30)).item(),
        torch.allclose(outs[a][k], outs[b][k], rtol=1e-4, atol=1e-6)) for k in ("render", "means3D", "opacities", "shs")))
