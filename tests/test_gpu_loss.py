"""SURVEY.md §8 f3 — fused L1 + SSIM loss (value and gradient in one kernel) against the reference's formulation
(trainer.py:102-107 _stclamp, :354-371 _compute_ssim, :485-492) restated with the same torch ops."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _reference_loss(render, gt, lam):
    x = render + (render.clamp(0.0, 1.0) - render).detach()
    l1 = F.l1_loss(x, gt)
    # .contiguous(): with the permuted (channels-last) views the reference passes, torch 2.11's CUDA
    # avg_pool2d backward returns gradients that differ from its own CPU / contiguous-CUDA result (forward values
    # agree) — measured here; the checker is the mathematically consistent variant.
    a, b = x.permute(2, 0, 1).unsqueeze(0).contiguous(), gt.permute(2, 0, 1).unsqueeze(0).contiguous()
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    mu1, mu2 = F.avg_pool2d(a, 3, 1, padding=1), F.avg_pool2d(b, 3, 1, padding=1)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.avg_pool2d(a * a, 3, 1, padding=1) - mu1_sq
    s2 = F.avg_pool2d(b * b, 3, 1, padding=1) - mu2_sq
    s12 = F.avg_pool2d(a * b, 3, 1, padding=1) - mu1_mu2
    ssim = (((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()
    return (1.0 - lam) * l1 + lam * (1.0 - ssim), l1, ssim


@pytest.mark.parametrize("H,W", [(16, 32), (37, 53), (270, 480), (1080, 1920), (1, 1), (3, 200)])
@pytest.mark.parametrize("lam", [0.2, 0.0, 1.0])
def test_fused_loss_matches_reference(pkg, H, W, lam):
    g = torch.Generator().manual_seed(H * 1000 + W)
    render = (torch.rand(H, W, 3, generator=g) * 1.4 - 0.2).to(DEV)  # exercises the clamp on both sides
    gt = torch.rand(H, W, 3, generator=g).to(DEV)
    r64 = render.double().requires_grad_(True)
    loss_ref, l1_ref, ssim_ref = _reference_loss(r64, gt.double(), lam)
    loss_ref.backward()
    r = render.clone().requires_grad_(True)
    out = pkg.l1_ssim_loss(r, gt, lam)
    out["loss"].backward()
    assert abs(out["loss"].item() - loss_ref.item()) <= 2e-6 * max(1.0, abs(loss_ref.item()))
    assert abs(out["l1"].item() - l1_ref.item()) <= 2e-6 and abs(out["ssim"].item() - ssim_ref.item()) <= 5e-6
    mse_ref = ((render.double().clamp(0, 1) - gt.double()) ** 2).mean().item()
    assert abs(out["mse"].item() - mse_ref) <= 2e-6
    rel = ((r.grad.double() - r64.grad).norm() / r64.grad.norm().clamp_min(1e-30)).item()
    assert rel <= 2e-4, rel


def test_fused_loss_gradient_matches_cpu_autograd(pkg):
    """Same comparison with the reference chain evaluated by CPU autograd (no CUDA pooling kernels involved)."""
    H, W = 45, 70
    g = torch.Generator().manual_seed(77)
    render = torch.rand(H, W, 3, generator=g) * 1.4 - 0.2
    gt = torch.rand(H, W, 3, generator=g)
    r64 = render.double().requires_grad_(True)
    loss_ref = _reference_loss(r64, gt.double(), 0.2)[0]
    loss_ref.backward()
    r = render.to(DEV).requires_grad_(True)
    out = pkg.l1_ssim_loss(r, gt.to(DEV), 0.2)
    out["loss"].backward()
    assert abs(out["loss"].item() - loss_ref.item()) <= 2e-6
    assert ((r.grad.cpu().double() - r64.grad).norm() / r64.grad.norm()).item() <= 2e-4


def test_loss_scaling_and_no_grad(pkg):
    H, W = 64, 96
    render = torch.rand(H, W, 3, device=DEV, requires_grad=True)
    gt = torch.rand(H, W, 3, device=DEV)
    (0.25 * pkg.l1_ssim_loss(render, gt)["loss"]).backward()
    g1 = render.grad.clone()
    render.grad = None
    pkg.l1_ssim_loss(render, gt)["loss"].backward()
    assert torch.allclose(4.0 * g1, render.grad, rtol=1e-6, atol=1e-12)
    with torch.no_grad():
        out = pkg.l1_ssim_loss(render, gt)
    assert not out["loss"].requires_grad
    with pytest.raises(RuntimeError):
        pkg.l1_ssim_loss(render.detach().cpu(), gt.cpu())


def test_loss_drives_the_rasterizer(pkg, oracle):
    """render -> fused loss -> backward through the rasterizer equals the torch-op loss chain."""
    N, W, H = 6000, 160, 96
    sc = oracle.synthetic_scene(N, W, H, seed=12)
    act = oracle.activate(sc)
    gt = torch.rand(H, W, 3, generator=torch.Generator().manual_seed(9)).to(DEV)
    be = pkg.B200Backend(sc["K"].to(DEV))
    vp = {"world_view_transform": sc["viewmat"].to(DEV), "image_width": W, "image_height": H}
    grads = []
    for fused in (True, False):
        leaves = {k: act[k].to(DEV).clone().requires_grad_(True) for k in act}
        res = be.render_with_depth(leaves, vp, bg_color=torch.zeros(3, device=DEV), render_mode="RGB", sh_degree=3)
        loss = pkg.l1_ssim_loss(res["render"], gt)["loss"] if fused else _reference_loss(res["render"], gt, 0.2)[0]
        loss.backward()
        grads.append({k: t.grad for k, t in leaves.items()})
    for k in grads[0]:
        rel = ((grads[0][k] - grads[1][k]).norm() / grads[1][k].norm()).item()
        assert rel <= 1e-3, (k, rel)


def test_torch_avg_pool2d_backward_on_the_references_permuted_views():
    """DESIGN.md §7 / VERDICT r1 item 3c.  The reference computes SSIM on `x.permute(2,0,1).unsqueeze(0)` — a channels-last
    VIEW of the [H,W,3] render (src/training/trainer.py:487-490).  On this image (torch 2.11.0+cu128) the CUDA backward of
    `avg_pool2d` for such inputs returns a gradient that differs from torch's own CPU result and from its CUDA result on a
    contiguous copy by 57-73 % (L2-relative; forward values agree to rounding) — measured by profiles/avgpool_probe.py on
    the B200 box.  The fused loss kernel (loss.cu) follows the CPU / contiguous result, i.e. it deliberately does NOT
    reproduce that defect.  This test pins both facts: the contiguous formulation is the consistent one, and the
    deviation of the permuted one is either the known defect (> 10 %) or gone (a fixed torch) — never a small drift."""
    from dropin_harness import avgpool_permuted_backward_deviation
    for H, W in ((16, 32), (90, 160)):
        dev = avgpool_permuted_backward_deviation(H, W)
        print(f"{H}x{W}: permuted {dev['permuted']:.3e}  contiguous {dev['contiguous']:.3e}")
        assert dev["contiguous"] <= 1e-9
        assert dev["permuted"] <= 1e-9 or dev["permuted"] >= 0.1
