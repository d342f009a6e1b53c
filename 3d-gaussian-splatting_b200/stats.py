"""Stage 5: fused densification-statistics update (reference src/training/trainer.py:786-791;
buffers src/gaussians/gaussian_model.py:78-80)."""
from __future__ import annotations

import torch

from ._lib import lib, check


def densify_stats(xyz_gradient_accum: torch.Tensor, xyz_gradient_count: torch.Tensor, xyz_grad: torch.Tensor) -> None:
    """In place: ``accum[N,3] += ||grad||_2[:,None]``, ``count[N,1] += 1`` in one kernel."""
    for t in (xyz_gradient_accum, xyz_gradient_count, xyz_grad):
        if t.device.type != "cuda" or t.dtype != torch.float32 or not t.is_contiguous():
            raise RuntimeError("densify_stats: contiguous fp32 CUDA tensors required (no CPU fallback)")
    N = xyz_grad.shape[0]
    if xyz_gradient_accum.shape != (N, 3) or xyz_gradient_count.numel() != N:
        raise RuntimeError("densify_stats: shape mismatch")
    with torch.cuda.device(xyz_grad.device):
        check(lib.b200_densify_stats(xyz_grad.data_ptr(), N, xyz_gradient_accum.data_ptr(),
                                     xyz_gradient_count.data_ptr(), torch.cuda.current_stream().cuda_stream),
              "b200_densify_stats")
