"""Generates tests/golden/splat_small.pt.

The reference cannot produce vectors for this path (its only backend needs the absent, CUDA-only gsplat
wheel — SURVEY.md F1-F5), so the fixture is made by this repo's own oracle: fp32 forward for the
integer outputs and the image, fp64 autograd for the gradients.  It freezes the oracle's behaviour so a
later edit of oracle/ that changes results is caught; parity with gsplat itself stays "unpinned".
    python tests/golden/make_golden.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import torch_oracle as orc  # noqa: E402

N, W, H, deg = 1500, 96, 64, 3
sc = orc.synthetic_scene(N, W, H, seed=123)
act = orc.activate(sc)
bg = torch.tensor([0.2, 0.4, 0.6])
r, a, m = orc.rasterization(act["means3D"], act["rotations"], act["scales"], act["opacities"].squeeze(-1), act["shs"],
                            sc["viewmat"], sc["K"], W, H, sh_degree=deg, background=bg)
v = torch.randn(H, W, 3, generator=torch.Generator().manual_seed(5))
leaves = {k: t.double().clone().requires_grad_(True) for k, t in act.items()}
r64, _, m64 = orc.rasterization(leaves["means3D"], leaves["rotations"], leaves["scales"], leaves["opacities"].squeeze(-1),
                                leaves["shs"], sc["viewmat"].double(), sc["K"].double(), W, H, sh_degree=deg,
                                background=bg.double())
assert torch.equal(m64["radii"], m["radii"]), "fp32/fp64 binning differ: pick another seed"
(r64 * v.double()).sum().backward()
out = dict(W=W, H=H, deg=deg, bg=bg, act=act, scene={"viewmat": sc["viewmat"], "K": sc["K"]}, render=r, alpha=a,
           radii=m["radii"], isect_ids=m["isect_ids"], flatten_ids=m["flatten_ids"], isect_offsets=m["isect_offsets"],
           v_render=v, grads={k: t.grad.float() for k, t in leaves.items()})
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "splat_small.pt")
torch.save(out, path)
print("wrote", path, os.path.getsize(path), "bytes; M =", m["isect_ids"].numel())
