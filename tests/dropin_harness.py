"""Test / bench harness that drives the reference's OWN, unmodified code on top of a backend.

What it provides (TEST INFRASTRUCTURE — the product path never imports this file):

* ``reference()``: registers B200Backend with the reference tree (``dropin.register``: /root/reference in the
  build container, the sha256-verified copy under baseline/_ref on the GPU box) and returns its modules.
* ``OracleBackend``: the CPU oracle (oracle/cpu_splat.c) behind the interface of
  ``GSplatBackend.render_with_depth`` (src/backends/gsplat_backend.py:129-210) — the comparator arm.
* ``synthetic_capture()``: a seeded ground-truth scene rendered from a ring of cameras, written as a LOSSLESS
  video (what ``VideoLoader`` / ``cv2.VideoCapture`` read, src/video/video_loader.py:4-63) plus the
  ``merged_data`` dict the SfM stage would hand to ``GaussianTrainer.train`` (src/training/trainer.py:653;
  keys read at :228-249, :111-168, :676-703) and the per-video pickle ``src/eval.py:121-128`` loads.
* ``make_trainer()``: a real ``GaussianTrainer`` + ``GaussianModel`` + ``Adam`` (``_setup_optimizer``,
  trainer.py:331-340) + ``GaussianRasterizer(..., backend="gsplat")`` (trainer.py:706-708) on a device.
"""
from __future__ import annotations

import contextlib
import importlib
import os
import pickle
import sys
from typing import Callable, Dict, List, Optional

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG = "3d-gaussian-splatting_b200"
KEYS = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")


_pkg = importlib.import_module(PKG)
reference_available = _pkg.dropin.reference_available
reference = _pkg.dropin.reference_modules          # (rasterizer, trainer, model, config modules) with B200Backend registered
backend_override = _pkg.dropin.backend_override
make_trainer = _pkg.dropin.make_reference_trainer  # a real GaussianTrainer + GaussianModel + Adam + GaussianRasterizer
tensor_views = _pkg.dropin.tensor_views


# ------------------------------------------------------------------------------------------------------------
# comparator arm: the CPU oracle behind the reference's backend interface
# ------------------------------------------------------------------------------------------------------------
class _OracleRender(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means, scales, quats, opac, shs, viewmat, K, W, H, deg, bg):
        from oracle import cpu_port
        act = {"means3D": means, "scales": scales, "rotations": quats, "opacities": opac, "shs": shs}
        fw = cpu_port.forward(act, viewmat, K, W, H, deg, bg=bg)
        ctx.fw, ctx.dims = fw, (W, H, deg, opac.shape)
        return fw["render"]

    @staticmethod
    def backward(ctx, v_render):
        from oracle import cpu_port
        W, H, deg, oshape = ctx.dims
        g = cpu_port.backward(ctx.fw, W, H, deg, v_render.contiguous())
        return (g["means3D"], g["scales"], g["rotations"], g["opacities"].reshape(oshape), g["shs"],
                None, None, None, None, None, None)


class OracleBackend:
    """``render_with_depth`` of gsplat_backend.py:129-210 (RGB mode, the only one train/eval request) on the CPU oracle."""

    def __init__(self, K):
        self.K = K

    def render_with_depth(self, gaussian_params, viewpoint, bg_color=None, render_mode="RGB", device="cpu", sh_degree=None):
        if render_mode != "RGB":
            raise NotImplementedError("oracle comparator arm: RGB only")
        if sh_degree is None:
            sh_degree = 3
        W, H = int(viewpoint["image_width"]), int(viewpoint["image_height"])
        bg = torch.zeros(3) if bg_color is None else bg_color.detach().float().cpu()
        img = _OracleRender.apply(gaussian_params["means3D"], gaussian_params["scales"], gaussian_params["rotations"],
                                  gaussian_params["opacities"], gaussian_params["shs"],
                                  viewpoint["world_view_transform"], self.K, W, H, int(sh_degree), bg)
        return {"render": img}


# ------------------------------------------------------------------------------------------------------------
# synthetic capture: lossless video + what the SfM stage would produce
# ------------------------------------------------------------------------------------------------------------
def ring_cameras(n: int, radius: float = 0.5, roll: float = 0.1) -> torch.Tensor:
    """n world->camera matrices around the scene of ``synthetic_scene`` with yaw, pitch AND roll (general poses)."""
    import math
    mats = []
    for i in range(n):
        th = 2 * math.pi * i / n
        pos = torch.tensor([radius * math.sin(th), 0.3 * radius * math.cos(2 * th), radius * 0.5 * (1 - math.cos(th))],
                           dtype=torch.float64)
        fwd = torch.tensor([0.0, 0.0, 6.0], dtype=torch.float64) - pos
        fwd = fwd / fwd.norm()
        ang = roll * math.sin(3 * th + 0.7)
        up = torch.tensor([math.sin(ang), -math.cos(ang), 0.0], dtype=torch.float64)
        right = torch.linalg.cross(fwd, up)
        right = right / right.norm()
        down = torch.linalg.cross(fwd, right)
        Rc = torch.stack([right, down, fwd], dim=0)
        m = torch.eye(4, dtype=torch.float64)
        m[:3, :3] = Rc
        m[:3, 3] = -Rc @ pos
        mats.append(m)
    return torch.stack(mats).to(torch.float32)


def write_lossless_video(path: str, frames_u8: List[np.ndarray]) -> str:
    """BGR uint8 frames -> a video cv2.VideoCapture decodes bit-exactly (FFV1, else PNG-in-AVI)."""
    import cv2
    h, w = frames_u8[0].shape[:2]
    for fourcc in ("FFV1", "png "):
        wr = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*fourcc), 30, (w, h))
        if not wr.isOpened():
            continue
        for f in frames_u8:
            wr.write(np.ascontiguousarray(f))
        wr.release()
        cap = cv2.VideoCapture(path)
        ok, f0 = cap.read()
        n = int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
        cap.release()
        if ok and n == len(frames_u8) and np.array_equal(f0, frames_u8[0]):
            return fourcc
    raise RuntimeError("no lossless cv2 video codec available (FFV1 / PNG)")


def synthetic_capture(outdir: str, render_fn: Callable, truth: Dict[str, torch.Tensor], act: Dict[str, torch.Tensor],
                      cams: torch.Tensor, W: int, H: int, n_points: int = 4000, seed: int = 3) -> Dict:
    """render_fn(act, viewmat, K, W, H) -> [H,W,3] float image in [0,1] (CPU tensor)."""
    os.makedirs(outdir, exist_ok=True)
    K = truth["K"]
    frames = []
    for vm in cams:
        img = render_fn(act, vm, K, W, H).clamp(0, 1)
        frames.append((img.numpy() * 255.0 + 0.5).astype(np.uint8))
    video = os.path.join(outdir, "capture.avi")
    codec = write_lossless_video(video, frames)
    g = torch.Generator().manual_seed(seed)
    vis = torch.sigmoid(truth["opacity"].reshape(-1)) > 0.3
    idx = torch.nonzero(vis).reshape(-1)
    idx = idx[torch.randperm(idx.numel(), generator=g)[:n_points]]
    pts = truth["xyz"][idx].double().numpy()
    cols = (truth["features_dc"][idx, 0, :] * 0.28209479177387814 + 0.5).clamp(0, 1).double().numpy()
    poses = cams.double().numpy()
    merged = {
        "points_3d": pts, "colors": cols,
        "video_info": [{"path": video}],
        "all_poses": [poses], "all_intrinsics": [K.double().numpy()],
        "frame_indices": [list(range(len(cams)))],
    }
    sfm_pkl = os.path.join(outdir, "capture_sfm.pkl")
    with open(sfm_pkl, "wb") as f:  # the per-video cache src/eval.py:121-128 reads
        pickle.dump({"poses": poses, "intrinsics": K.double().numpy(), "frame_indices": list(range(len(cams))),
                     "video_path": video, "points_3d": pts, "colors": cols}, f)
    return dict(merged_data=merged, video=video, sfm_pkl=sfm_pkl, frames=frames, codec=codec, K=K, cams=cams)


# ------------------------------------------------------------------------------------------------------------
# torch 2.11 CUDA avg_pool2d backward on the permuted views the reference's SSIM passes (trainer.py:487-490)
# ------------------------------------------------------------------------------------------------------------
def avgpool_permuted_backward_deviation(H: int = 90, W: int = 160) -> Dict[str, float]:
    """L2-relative deviation of d(mean(avg_pool2d(x_chw)))-style SSIM gradients on CUDA from the CPU result, for the
    reference's permuted [H,W,3] -> [1,3,H,W] views and for contiguous copies of them (profiles/avgpool_probe.py)."""
    import torch.nn.functional as F

    def ssim(a, b):
        C1, C2 = 0.01 ** 2, 0.03 ** 2
        mu1, mu2 = F.avg_pool2d(a, 3, 1, padding=1), F.avg_pool2d(b, 3, 1, padding=1)
        s1 = F.avg_pool2d(a * a, 3, 1, padding=1) - mu1.pow(2)
        s2 = F.avg_pool2d(b * b, 3, 1, padding=1) - mu2.pow(2)
        s12 = F.avg_pool2d(a * b, 3, 1, padding=1) - mu1 * mu2
        return (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))).mean()

    def grad(x, gt, dev, contiguous):
        x = x.to(dev).double().requires_grad_(True)
        a, b = x.permute(2, 0, 1).unsqueeze(0), gt.to(dev).double().permute(2, 0, 1).unsqueeze(0)
        if contiguous:
            a, b = a.contiguous(), b.contiguous()
        ssim(a, b).backward()
        return x.grad.cpu()

    gen = torch.Generator().manual_seed(H)
    x, gt = torch.rand(H, W, 3, generator=gen), torch.rand(H, W, 3, generator=gen)
    ref = grad(x, gt, "cpu", False)
    out = {}
    for name, contiguous in (("permuted", False), ("contiguous", True)):
        g = grad(x, gt, "cuda", contiguous)
        out[name] = ((g - ref).norm() / ref.norm()).item()
    return out


@contextlib.contextmanager
def contiguous_avg_pool2d():
    """`torch.nn.functional.avg_pool2d` on a contiguous copy of its input for the duration of the block — the
    reference's loss code is untouched; only torch's own CUDA backward defect on channels-last views is bypassed."""
    import torch.nn.functional as F
    orig = F.avg_pool2d

    def fixed(x, *a, **k):
        return orig(x.contiguous(), *a, **k)

    F.avg_pool2d = fixed
    try:
        yield
    finally:
        F.avg_pool2d = orig
