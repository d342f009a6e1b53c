"""Synthetic workload generator of the bench: scenes of the named sizes, the reference getters'
activations and the config-B camera ring.  Deterministic CPU generator, so every rank, the CPU baseline and the tests see
the same bits.  Pure torch, no kernels; it lives in the package so that ``bench.py``'s B200 arm never imports ``oracle/``
(the oracle keeps its own copy for the tests; ``tests/test_oracle_cpu.py`` checks the two agree bit for bit)."""
from __future__ import annotations

import math
from typing import Dict

import torch


def synthetic_scene(N: int, W: int, H: int, seed: int = 0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    fx = fy = (24.0 / 36.0) * max(W, H)
    cx, cy = W / 2.0, H / 2.0
    z = 2.0 + 8.0 * torch.rand(N, generator=g)
    u = torch.rand(N, generator=g)
    v = torch.rand(N, generator=g)
    x = z * (u - 0.5) * 1.1 * W / fx
    y = z * (v - 0.5) * 1.1 * H / fy
    xyz = torch.stack([x, y, z], dim=-1)
    log_scale = (math.log(0.012) + 0.6 * torch.randn(N, 3, generator=g)).clamp(math.log(1e-4), math.log(0.3))
    rot = torch.randn(N, 4, generator=g)
    opacity_logit = 2.0 * torch.randn(N, 1, generator=g)
    features_dc = (torch.rand(N, 1, 3, generator=g) * 2 - 1) * 1.77
    features_rest = 0.1 * torch.randn(N, 15, 3, generator=g)
    Kmat = torch.tensor([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=torch.float32)
    out = dict(
        xyz=xyz, scaling=log_scale, rotation=rot, opacity=opacity_logit,
        features_dc=features_dc, features_rest=features_rest,
        K=Kmat, viewmat=torch.eye(4),
    )
    return {k: (t.to(dtype) if t.is_floating_point() else t) for k, t in out.items()}


def activate(scene: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Activations exactly as the reference's getters (gaussian_model.py:90-108)."""
    return {
        "means3D": scene["xyz"],
        "scales": torch.exp(scene["scaling"]),
        "rotations": torch.nn.functional.normalize(scene["rotation"], dim=-1),
        "opacities": torch.sigmoid(scene["opacity"]),
        "shs": torch.cat([scene["features_dc"], scene["features_rest"]], dim=1),
    }


def circle_cameras(n: int, radius: float = 1.0, target=(0.0, 0.0, 6.0)) -> torch.Tensor:
    """Config-B cameras: n w2c matrices on a circle in the xz-plane looking at target."""
    mats = []
    tgt = torch.tensor(target, dtype=torch.float64)
    for i in range(n):
        th = 2 * math.pi * i / n
        # camera sits on a circle around a point `radius` in front of the target so that it keeps
        # looking roughly down +z (the scene only exists in front of the origin camera)
        pos = torch.tensor([radius * math.sin(th), 0.0, radius * (1 - math.cos(th)) * 0.5], dtype=torch.float64)
        fwd = tgt - pos
        fwd = fwd / fwd.norm()
        up = torch.tensor([0.0, -1.0, 0.0], dtype=torch.float64)
        right = torch.linalg.cross(fwd, up)
        right = right / right.norm()
        down = torch.linalg.cross(fwd, right)
        Rc = torch.stack([right, down, fwd], dim=0)  # rows = camera axes in world coords
        m = torch.eye(4, dtype=torch.float64)
        m[:3, :3] = Rc
        m[:3, 3] = -Rc @ pos
        mats.append(m)
    return torch.stack(mats).to(torch.float32)
