"""Oracle pin against the reference's real arithmetic provider, gsplat==1.5.3 (/root/reference requirements.txt:9),
IF that wheel imports on the box (SURVEY §8c "Strengthening on the GPU box").  It is not in the image and there is no
network, so on the driver's boxes this records "gsplat unavailable on this box" (a skip with that reason) and parity
with gsplat stays unpinned; the comparison below is what runs the day the wheel is present.

Call form = src/backends/gsplat_backend.py:168-182.
"""
import importlib.util

import pytest
import torch

from parity_util import CAMS, apply_camera, grad_rows, image_close

pytestmark = pytest.mark.gpu
DEV = "cuda"


def real_gsplat():
    """The real wheel, never this repo's registration shim (dropin.install_shims puts one in sys.modules)."""
    import sys
    mod = sys.modules.get("gsplat")
    if mod is not None and getattr(mod, "__version__", "") == "b200-shim":
        return None
    try:
        if importlib.util.find_spec("gsplat") is None:
            return None
        import gsplat
        return gsplat
    except Exception:  # a wheel built for another torch / CUDA fails at import: unavailable
        return None


def test_gsplat_probe_is_recorded():
    g = real_gsplat()
    msg = f"gsplat {g.__version__} importable: pin tests below run" if g is not None else "gsplat unavailable on this box"
    print(f"oracle_pin: {msg}")
    if g is None:
        pytest.skip("gsplat unavailable on this box (parity with gsplat 1.5.3 stays unpinned; oracle = SURVEY Appendix A)")


@pytest.mark.parametrize("cam", CAMS)
@pytest.mark.parametrize("N,W,H", [(20000, 320, 192), (1_000_000, 1920, 1080)])
def test_b200_matches_gsplat(pkg, oracle, N, W, H, cam):
    gsplat = real_gsplat()
    if gsplat is None:
        pytest.skip("gsplat unavailable on this box")
    sc = apply_camera(oracle.synthetic_scene(N, W, H, seed=0), cam, W, H)
    act = oracle.activate(sc)
    v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1)).to(DEV)
    out = {}
    for name, fn in (("gsplat", gsplat.rasterization), ("b200", pkg.rasterization)):
        leaves = {k: act[k].to(DEV).clone().requires_grad_(True) for k in ("means3D", "rotations", "scales", "opacities", "shs")}
        r, a, meta = fn(means=leaves["means3D"], quats=leaves["rotations"], scales=leaves["scales"],
                        opacities=leaves["opacities"].squeeze(-1), colors=leaves["shs"], viewmats=sc["viewmat"].to(DEV)[None],
                        Ks=sc["K"].to(DEV)[None], width=W, height=H, packed=False, render_mode="RGB", sh_degree=3,
                        backgrounds=torch.zeros(1, 3, device=DEV), **({"meta_isect_ids": True} if name == "b200" else {}))
        (r * v).sum().backward()
        out[name] = (r.detach(), a.detach(), meta, {k: t.grad for k, t in leaves.items()})
    (rg, ag, mg, gg), (rb, ab, mb, gb) = out["gsplat"], out["b200"]
    same_radii = torch.equal(mg["radii"].reshape(-1, 2).cpu(), mb["radii"].reshape(-1, 2).cpu())
    print(f"radii identical: {same_radii}; M gsplat {mg['flatten_ids'].numel()} / b200 {mb['flatten_ids'].numel()}")
    if same_radii:  # same projection bits -> tile keys, sort order and ranges must match bit-exactly
        assert torch.equal(mg["isect_ids"].reshape(-1).cpu(), mb["isect_ids"].reshape(-1).cpu())
        assert torch.equal(mg["flatten_ids"].reshape(-1).cpu(), mb["flatten_ids"].reshape(-1).cpu())
        assert torch.equal(mg["isect_offsets"].reshape(-1).cpu(), mb["isect_offsets"].reshape(-1).cpu())
    image_close(rb[0], rg[0], None, max_marginal=5e-3, what="b200 vs gsplat render")
    for k in gg:
        grad_rows(gb[k], gg[k], f"{k}/vs-gsplat")
