"""Fused photometric loss of the reference's training step.

``l1_ssim_loss(render, gt)`` == trainer.py:485-492 — ``x = _stclamp(render)``, ``(1 - lambda) * F.l1_loss(x, gt) +
lambda * (1 - _compute_ssim(x, gt))`` — with the value AND ``d loss / d render`` produced by one kernel
(``b200_l1_ssim_fwd_bwd``) instead of five ``avg_pool2d`` passes, a dozen elementwise kernels and their autograd
mirror.  Returns the loss and the per-view statistics the trainer logs (l1, ssim, psnr inputs) as device scalars:
nothing here synchronises with the host.
"""
from __future__ import annotations

from typing import Dict

import torch

from ._lib import lib, check


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, render, gt, lambda_dssim):
        if render.device.type != "cuda" or gt.device != render.device:
            raise RuntimeError("l1_ssim_loss: CUDA tensors on one device required (there is no CPU fallback)")
        if render.dim() != 3 or render.shape[-1] != 3 or gt.shape != render.shape:
            raise ValueError("l1_ssim_loss: render and gt must both be [H,W,3]")
        render, gt = render.contiguous(), gt.contiguous()
        H, W = render.shape[0], render.shape[1]
        need_grad = ctx.needs_input_grad[0]
        v_render = torch.empty_like(render) if need_grad else None
        sums = torch.empty(8, dtype=torch.float32, device=render.device)  # 3 sums, counter, (loss, l1, ssim, mse)
        with torch.cuda.device(render.device):
            check(lib.b200_l1_ssim_fwd_bwd(render.data_ptr(), gt.data_ptr(), W, H, float(lambda_dssim), 1.0,
                                           None if v_render is None else v_render.data_ptr(), sums.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream), "b200_l1_ssim_fwd_bwd")
        loss, stats = sums[4], sums[5:8]  # finished by the kernel's last CTA: (l1, ssim, mse of the clamped render)
        ctx.v_render = v_render
        ctx.mark_non_differentiable(stats)
        return loss, stats

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, v_loss, _v_stats):
        if ctx.v_render is None or v_loss is None:
            return None, None, None
        return ctx.v_render * v_loss, None, None


def l1_ssim_loss(render: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2) -> Dict[str, torch.Tensor]:
    """{'loss', 'l1', 'ssim', 'mse'} — 0-d device tensors; only 'loss' is differentiable (w.r.t. ``render``)."""
    loss, stats = _L1SSIM.apply(render, gt, float(lambda_dssim))
    return {"loss": loss, "l1": stats[0], "ssim": stats[1], "mse": stats[2]}
