"""SURVEY.md §8 f4 — one-kernel Adam over the flat store against torch.optim.Adam with the reference's six groups
(trainer.py:331-340)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
LRS = {"xyz": 1.6e-4, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}


@pytest.mark.parametrize("N", [1, 7, 4099])
def test_store_adam_matches_torch_adam(pkg, oracle, N):
    sc = oracle.synthetic_scene(N, 64, 48, seed=21)
    raw = {k: sc[k].to(DEV).clone().requires_grad_(True)
           for k in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")}
    groups = [{"params": [raw["xyz"]], "lr": LRS["xyz"]}, {"params": [raw["features_dc"]], "lr": LRS["f_dc"]},
              {"params": [raw["features_rest"]], "lr": LRS["f_rest"]}, {"params": [raw["opacity"]], "lr": LRS["opacity"]},
              {"params": [raw["scaling"]], "lr": LRS["scaling"]}, {"params": [raw["rotation"]], "lr": LRS["rotation"]}]
    ref = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    st = pkg.GaussianStore.from_tensors(sc["xyz"].to(DEV), sc["features_dc"].to(DEV), sc["features_rest"].to(DEV),
                                        sc["scaling"].to(DEV), sc["rotation"].to(DEV), sc["opacity"].to(DEV))
    opt = pkg.StoreAdam(st, LRS)
    gen = torch.Generator().manual_seed(5)
    for it in range(5):
        scale = 10.0 ** (it - 3)  # gradients from 1e-3 to 10
        st.zero_grad()
        for k, p in raw.items():
            g = (torch.randn(p.shape, generator=gen) * scale).to(DEV)
            p.grad = g.clone()
            getattr(st, k).grad.copy_(g)
        if it == 3:  # the trainer's position-rate schedule
            ref.param_groups[0]["lr"] = 3e-5
            opt.lrs["xyz"] = 3e-5
        ref.step()
        opt.step()
        for k, p in raw.items():
            q = getattr(st, k)
            assert torch.allclose(q.detach(), p.detach(), rtol=1e-5, atol=3e-7), (it, k, (q - p).abs().max().item())
    pad = st.flat.numel() - 59 * N
    assert pad >= 0 and opt.t == 5


def test_scale_ceiling_is_the_reference_clamp(pkg, oracle):
    """trainer.py:854-859: `scaling.data.clamp_(max=scale_ceiling)` after every optimiser step, here folded into the
    Adam kernel; every other block must be untouched by the bound."""
    N = 5000
    sc = oracle.synthetic_scene(N, 64, 48, seed=22)
    raw = {k: sc[k].to(DEV).clone().requires_grad_(True)
           for k in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")}
    groups = [{"params": [raw[k]], "lr": LRS[a]} for k, a in
              (("xyz", "xyz"), ("features_dc", "f_dc"), ("features_rest", "f_rest"), ("opacity", "opacity"),
               ("scaling", "scaling"), ("rotation", "rotation"))]
    ref = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    st = pkg.GaussianStore.from_tensors(sc["xyz"].to(DEV), sc["features_dc"].to(DEV), sc["features_rest"].to(DEV),
                                        sc["scaling"].to(DEV), sc["rotation"].to(DEV), sc["opacity"].to(DEV))
    ceiling = float(sc["scaling"].median())  # about half the log-scales sit above the cap
    opt = pkg.StoreAdam(st, LRS, scale_ceiling=ceiling)
    gen = torch.Generator().manual_seed(6)
    for it in range(3):
        st.zero_grad()
        for k, p in raw.items():
            g = torch.randn(p.shape, generator=gen).to(DEV)
            p.grad = g.clone()
            getattr(st, k).grad.copy_(g)
        ref.step()
        with torch.no_grad():
            raw["scaling"].data.clamp_(max=ceiling)
        opt.step()
        assert st.scaling.max().item() <= ceiling
        for k, p in raw.items():
            q = getattr(st, k)
            assert torch.allclose(q.detach(), p.detach(), rtol=1e-5, atol=3e-7), (it, k, (q - p).abs().max().item())
    opt.scale_ceiling = None  # switching the bound off again
    st.zero_grad()
    st.scaling.grad.fill_(-1.0)
    opt.step()
    assert st.scaling.max().item() > ceiling


@pytest.mark.parametrize("N", [1, 1000, 100_003])
def test_adam_with_fused_densify_statistic(pkg, oracle, N):
    """`StoreAdam.step(densify_stats=True)`: the stage-5 statistic (trainer.py:786-791) taken inside the optimiser pass
    equals `optimizer.step()` followed by the reference's three torch ops, bit for bit on the parameters."""
    sc = oracle.synthetic_scene(N, 64, 48, seed=23)
    mk = lambda: pkg.GaussianStore.from_tensors(sc["xyz"].to(DEV), sc["features_dc"].to(DEV), sc["features_rest"].to(DEV),  # noqa: E731
                                                sc["scaling"].to(DEV), sc["rotation"].to(DEV), sc["opacity"].to(DEV))
    a, b = mk(), mk()
    oa, ob = pkg.StoreAdam(a, LRS), pkg.StoreAdam(b, LRS)
    gen = torch.Generator().manual_seed(8)
    for it in range(3):
        g = torch.randn(a.flat.numel(), generator=gen).to(DEV)
        for st in (a, b):
            st.zero_grad()
            st.flat_grad.copy_(g)
        oa.step(densify_stats=True)
        ob.step()
        with torch.no_grad():  # the reference's ops
            b.xyz_gradient_accum += torch.norm(b.xyz.grad, dim=-1, keepdim=True)
            b.xyz_gradient_count += 1
        assert torch.equal(a.flat, b.flat)
        torch.testing.assert_close(a.xyz_gradient_accum, b.xyz_gradient_accum, rtol=1e-6, atol=1e-7)
        assert torch.equal(a.xyz_gradient_count, b.xyz_gradient_count) and float(a.xyz_gradient_count.min()) == it + 1


def test_store_adam_refuses_a_store_reallocated_behind_its_back(pkg, oracle):
    """`densify_and_prune(optimizer=None)` (the reference's signature default) re-lays the flat buffer; stepping
    the old moments would read out of bounds.  torch's Adam raises in that situation; so must StoreAdam."""
    sc = oracle.synthetic_scene(5000, 64, 48, seed=24)
    st = pkg.GaussianStore.from_tensors(sc["xyz"].to(DEV), sc["features_dc"].to(DEV), sc["features_rest"].to(DEV),
                                        sc["scaling"].to(DEV), sc["rotation"].to(DEV), sc["opacity"].to(DEV))
    opt = pkg.StoreAdam(st, LRS)
    st.xyz_gradient_accum.fill_(1.0)
    st.xyz_gradient_count.fill_(1.0)
    stats = st.densify_and_prune(1e-4, 0.05, 0.2, None, optimizer=None)
    assert stats["n_after"] != stats["n_before"]
    with pytest.raises(RuntimeError, match="re-allocated"):
        opt.step()
    opt2 = pkg.StoreAdam(st, LRS)
    st.zero_grad()
    opt2.step()
    stats = st.densify_and_prune(1e-4, 0.05, 0.2, None, optimizer=opt2)  # spliced: keeps stepping
    opt2.step()
