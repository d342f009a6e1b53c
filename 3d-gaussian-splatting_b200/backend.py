"""B200Backend — same method set, argument meaning and result dicts as the reference's
``GSplatBackend`` (src/backends/gsplat_backend.py:7-210), with ``gsplat.rasterization`` replaced by
the sm_100a path in :mod:`rasterization`.  ``render_with_depth`` is the method the trainer and eval
call (src/training/trainer.py:417-424,471-478; src/eval.py:180-183).
"""
from __future__ import annotations

from typing import Dict

import torch

from .rasterization import _render_one, rasterization


class B200Backend:
    def __init__(self, K):
        if not isinstance(K, torch.Tensor) or K.device.type != "cuda":
            raise RuntimeError("B200Backend needs CUDA intrinsics K[3,3]; there is no CPU fallback")
        self.rasterization = rasterization
        self.K = K

    # -- gsplat_backend.py:12-55 (the HWC->CHW slot bug of the reference, SURVEY.md B-3, is fixed by a permute)
    def render_batch(self, gaussian_params: Dict, view_matrices: torch.Tensor, proj_matrices: torch.Tensor,
                     width: int, height: int, bg_color: torch.Tensor, device: str):
        batch_size = view_matrices.shape[0]
        rendered_images = torch.zeros(batch_size, 3, height, width, device=device)
        alphas = torch.zeros(batch_size, 1, height, width, device=device)
        for i in range(batch_size):
            result = self.render_single(
                gaussian_params,
                {
                    "world_view_transform": view_matrices[i],
                    "projection_matrix": proj_matrices[i],
                    "image_width": width,
                    "image_height": height,
                },
                bg_color,
            )
            rendered_images[i] = result["render"].permute(2, 0, 1)
            alphas[i] = result["alpha"].reshape(1, height, width)
        return {"images": rendered_images, "alphas": alphas}

    # -- gsplat_backend.py:57-127 (RGB+ED, sh_degree 3; opacities squeezed here)
    def render_single(self, gaussian_params: Dict, viewpoint: Dict, bg_color: torch.Tensor):
        means3D = gaussian_params["means3D"]
        scales = gaussian_params["scales"]
        quats = gaussian_params["rotations"]
        opacities = gaussian_params["opacities"]
        if opacities.dim() == 2:
            opacities = opacities.squeeze(-1)
        if "shs" in gaussian_params and gaussian_params["shs"] is not None:
            colors = gaussian_params["shs"]
            sh_degree = 3
        else:
            colors = gaussian_params.get("colors", gaussian_params.get("features", None))
            sh_degree = 3 if (colors is not None and colors.dim() == 3) else None
        viewmat = viewpoint["world_view_transform"]
        renders, alphas, meta = self.rasterization(
            means=means3D, quats=quats, scales=scales, opacities=opacities, colors=colors,
            viewmats=viewmat[None, ...], Ks=self.K[None, ...],
            width=viewpoint["image_width"], height=viewpoint["image_height"], packed=False,
            render_mode="RGB+ED", sh_degree=sh_degree,
            backgrounds=bg_color[None, ...] if bg_color.dim() == 1 else bg_color,
        )
        rendered = renders.squeeze(0)  # one camera: squeeze is a view both ways (select's backward = zeros + full-image copy)
        result = {"render": rendered[..., :3], "alpha": alphas.squeeze(0), "depth": rendered[..., 3]}
        result["means2d"] = meta["means2d"][0]
        result["radii"] = meta["radii"][0]
        result["visibility_filter"] = meta["visibility_filter"][0]  # = radii > 0, from the projection kernel
        result["tiles_touched"] = meta["tiles_per_gauss"][0]  # key the reference probes
        return result

    # -- gsplat_backend.py:129-210: THE boundary
    def render_with_depth(self, gaussian_params: Dict, viewpoint: Dict, bg_color: torch.Tensor = None,
                          render_mode: str = "RGB+ED", device: str = "cuda", sh_degree: int = None) -> Dict:
        if str(device).startswith("cpu"):
            raise RuntimeError("B200Backend has no CPU path (device='cpu' requested)")
        if bg_color is None:
            bg_color = torch.zeros(3, device=device)
        means3D = gaussian_params["means3D"]
        scales = gaussian_params["scales"]
        quats = gaussian_params["rotations"]
        opacities = gaussian_params["opacities"]
        if opacities.dim() == 2:
            opacities = opacities.squeeze(-1)
        colors = gaussian_params.get("colors", gaussian_params.get("shs"))
        if sh_degree is None:
            sh_degree = 3 if "shs" in gaussian_params else None
        viewmat = viewpoint["world_view_transform"]
        if render_mode not in ("RGB", "D", "ED", "RGB+D", "RGB+ED"):
            raise ValueError(f"render_mode must be one of RGB, D, ED, RGB+D, RGB+ED, got {render_mode!r}")
        if self.rasterization is rasterization and viewmat.dim() == 2 and bg_color.dim() == 1:
            # the same call without the round trip through the gsplat-shaped signature ([None] on the way in, squeeze(0) /
            # [0] on the way out: a dozen tensor views per view on the host and two extra nodes in the backward graph)
            output, alpha, aux = _render_one(means3D, quats, scales, opacities, colors, viewmat, self.K,
                                             viewpoint["image_width"], viewpoint["image_height"], sh_degree, render_mode,
                                             bg_color, 0.01, 1e10, 0.3, 0.0)
            meta = None
        else:
            renders, alphas, meta = self.rasterization(
                means=means3D, quats=quats, scales=scales, opacities=opacities, colors=colors,
                viewmats=viewmat[None, ...], Ks=self.K[None, ...],
                width=viewpoint["image_width"], height=viewpoint["image_height"], packed=False,
                render_mode=render_mode, sh_degree=sh_degree, backgrounds=bg_color[None, ...],
            )
            output = renders.squeeze(0)  # one camera: squeeze is a view both ways (select's backward = zeros + full-image copy)
            alpha = alphas.squeeze(0)
        result = {"alpha": alpha}
        if render_mode == "RGB":
            result["render"] = output
        elif render_mode == "D":
            result["accumulated_depth"] = output.squeeze(-1)
        elif render_mode == "ED":
            result["expected_depth"] = output.squeeze(-1)
        elif render_mode == "RGB+D":
            result["render"] = output[..., :3]
            result["accumulated_depth"] = output[..., 3]
        elif render_mode == "RGB+ED":
            result["render"] = output[..., :3]
            result["expected_depth"] = output[..., 3]
        if meta is None:
            result["means2d"] = aux["splats"][:, 0:2]
            result["radii"] = aux["radii"]
            result["visibility_filter"] = aux["visibility"]  # = radii > 0, from the projection kernel
            return result
        if "means2d" in meta:
            result["means2d"] = meta["means2d"][0]
        if "radii" in meta:
            result["radii"] = meta["radii"][0]
            result["visibility_filter"] = meta["visibility_filter"][0]  # = radii > 0, from the projection kernel
        return result
