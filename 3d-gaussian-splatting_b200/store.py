"""GaussianStore — the B200 form of the reference's parameter store (src/gaussians/gaussian_model.py:55-108).

Same attribute names, shapes, initial values and getters as the reference's ``GaussianModel`` — ``xyz``,
``features_dc``, ``features_rest``, ``scaling`` (log), ``rotation`` (wxyz, un-normalised), ``opacity`` (logit),
``get_xyz / get_scaling / get_rotation / get_opacity / get_features``, ``xyz_gradient_accum``,
``xyz_gradient_count``, ``max_radii_2D`` — but laid out for the kernels:

* ONE flat fp32 buffer of 59 floats per Gaussian, ``[xyz 3N | scaling 3N | rotation 4N | opacity N | features
  48N]`` (each block 16-byte aligned), and ONE flat gradient buffer of the same layout.  Every parameter is a
  view of the first, its ``.grad`` a view of the second, so the NCCL all-reduce of a step is one call on one
  buffer (``distributed.allreduce_flat``) and ``zero_grad`` is one memset.
* ``features`` is stored as ``[N,16,3]`` (dc and rest contiguous per Gaussian): ``features_dc`` /
  ``features_rest`` are strided views of it, so ``get_features`` is the buffer itself — the reference's
  ``torch.cat`` (192 B/Gaussian copied every step, gaussian_model.py:106-108) disappears, and the projection
  kernel reads the coefficients with 12 aligned 16-byte loads.
* ``render()`` feeds the RAW parameters to the kernels, which apply exp / sigmoid / normalise in registers
  (``b200_project_sh_fwd_raw``), and the backward pass ADDS the raw-parameter gradients straight into the flat
  gradient buffer (``b200_project_sh_bwd_raw(accumulate=1)``): no activated copies, no autograd nodes for the
  getters, no per-view ``grad += new`` passes.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .rasterization import rasterization

_BLOCKS = (("xyz", 3), ("scaling", 3), ("rotation", 4), ("opacity", 1), ("features", 48))
FLOATS_PER_GAUSSIAN = sum(w for _, w in _BLOCKS)  # 59


def _align4(n: int) -> int:
    return (n + 3) // 4 * 4  # floats -> 16-byte multiples


class GaussianStore(nn.Module):
    def __init__(self, n_gaussians: int = 100_000, device: str = "cuda"):
        super().__init__()
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("GaussianStore lives in HBM; there is no CPU path")
        n = int(n_gaussians)
        self._allocate(n, dev)
        with torch.no_grad():  # the reference's initial values (gaussian_model.py:65-84)
            self.xyz.copy_(torch.randn(n, 3, device=dev) * 10)
            self.scaling.fill_(-3.0)
            self.rotation.zero_()
            self.rotation[:, 0] = 1
            self.opacity.zero_()
            self.features.zero_()
        self.xyz_gradient_accum = torch.zeros(n, 3, device=dev)
        self.xyz_gradient_count = torch.zeros(n, 1, device=dev)
        self.max_radii_2D = torch.zeros(n, device=dev)

    # ------------------------------------------------------------------ layout
    def _allocate(self, n: int, dev: torch.device) -> None:
        offs, o = {}, 0
        for name, w in _BLOCKS:
            offs[name] = o
            o += _align4(n * w)
        self._offsets, self._n = offs, n
        self.flat = torch.zeros(o, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(o, dtype=torch.float32, device=dev)

        def view(buf, name, *shape):
            w = dict(_BLOCKS)[name]
            return buf[offs[name]: offs[name] + n * w].view(n, *shape)

        feats = view(self.flat, "features", 16, 3)
        gfeats = view(self.flat_grad, "features", 16, 3)
        self.xyz = nn.Parameter(view(self.flat, "xyz", 3))
        self.scaling = nn.Parameter(view(self.flat, "scaling", 3))
        self.rotation = nn.Parameter(view(self.flat, "rotation", 4))
        self.opacity = nn.Parameter(view(self.flat, "opacity", 1))
        self.features = nn.Parameter(feats)
        # reference names: strided views of the same storage.  They are Parameters (so the reference's six
        # optimiser groups, trainer.py:331-340, can be built from them) but are not registered with the module,
        # so parameters() lists every float once.
        object.__setattr__(self, "features_dc", nn.Parameter(feats[:, 0:1, :]))
        object.__setattr__(self, "features_rest", nn.Parameter(feats[:, 1:, :]))
        self.xyz.grad = view(self.flat_grad, "xyz", 3)
        self.scaling.grad = view(self.flat_grad, "scaling", 3)
        self.rotation.grad = view(self.flat_grad, "rotation", 4)
        self.opacity.grad = view(self.flat_grad, "opacity", 1)
        self.features.grad = gfeats
        self.features_dc.grad = gfeats[:, 0:1, :]
        self.features_rest.grad = gfeats[:, 1:, :]

    @classmethod
    def from_tensors(cls, xyz, features_dc, features_rest, scaling, rotation, opacity) -> "GaussianStore":
        """Pack the reference's six raw parameter tensors (e.g. ``GaussianModel`` attributes or a checkpoint)."""
        st = cls(xyz.shape[0], device=str(xyz.device))
        with torch.no_grad():
            st.xyz.copy_(xyz)
            st.scaling.copy_(scaling)
            st.rotation.copy_(rotation)
            st.opacity.copy_(opacity.reshape(-1, 1))
            st.features[:, 0:1, :].copy_(features_dc.reshape(-1, 1, 3))
            st.features[:, 1:, :].copy_(features_rest.reshape(-1, 15, 3))
        return st

    def __len__(self) -> int:
        return self._n

    # ------------------------------------------------------------------ reference getters (gaussian_model.py:90-108)
    @property
    def get_xyz(self):
        return self.xyz

    @property
    def get_scaling(self):
        return torch.exp(self.scaling)

    @property
    def get_rotation(self):
        return F.normalize(self.rotation, dim=-1)

    @property
    def get_opacity(self):
        return torch.sigmoid(self.opacity)

    @property
    def get_features(self):
        return self.features

    def gaussian_params(self) -> Dict[str, torch.Tensor]:
        """The activated input dict of trainer.py:399-406 (for the unfused backend call / parity tests)."""
        return {"means3D": self.get_xyz, "scales": self.get_scaling, "rotations": self.get_rotation,
                "opacities": self.get_opacity, "shs": self.get_features}

    # ------------------------------------------------------------------ gradients
    _SHAPES = {"xyz": (3,), "scaling": (3,), "rotation": (4,), "opacity": (1,), "features": (16, 3)}

    def _grad_view(self, name: str) -> torch.Tensor:
        w = dict(_BLOCKS)[name]
        o = self._offsets[name]
        return self.flat_grad[o: o + self._n * w].view(self._n, *self._SHAPES[name])

    def _attach_grads(self) -> bool:
        """Make every ``param.grad`` the matching view of ``flat_grad``; True if any had been detached."""
        detached = False
        for name in self._SHAPES:
            p = getattr(self, name)
            want = self._grad_view(name)
            if p.grad is None or p.grad.data_ptr() != want.data_ptr():
                p.grad = want
                detached = True
        gf = self._grad_view("features")
        for p, want in ((self.features_dc, gf[:, 0:1, :]), (self.features_rest, gf[:, 1:, :])):
            if p.grad is None or p.grad.data_ptr() != want.data_ptr():
                p.grad = want
                detached = True
        return detached

    def optimizer_groups(self, lrs: Dict[str, float]):
        """The reference's six parameter groups (trainer.py:331-340), e.g. for ``torch.optim.Adam(..., eps=1e-15)``.
        ``lrs`` maps xyz / features_dc / features_rest / scaling / rotation / opacity to learning rates."""
        return [{"params": [getattr(self, k)], "lr": lrs[k], "name": k}
                for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")]

    def zero_grad(self, set_to_none: bool = False) -> None:
        """One memset of the flat gradient buffer; the ``.grad`` views stay attached (``set_to_none`` is ignored:
        the kernels accumulate into the buffer, not into tensors autograd hands out)."""
        self.flat_grad.zero_()
        self._attach_grads()

    def grad_buffers(self):
        """(g_xyz, g_rotation, g_scaling, g_opacity, g_features) in the kernels' argument order."""
        if self._attach_grads():
            # an optimiser's zero_grad(set_to_none=True) dropped the views: honour it as "gradients are zero"
            self.flat_grad.zero_()
        return (self._grad_view("xyz"), self._grad_view("rotation"), self._grad_view("scaling"),
                self._grad_view("opacity").view(-1), self._grad_view("features"))

    # ------------------------------------------------------------------ rendering
    def render(self, K: torch.Tensor, viewpoint: Dict, bg_color: Optional[torch.Tensor] = None,
               render_mode: str = "RGB", sh_degree: int = 3, accumulate_grads: bool = True) -> Dict[str, torch.Tensor]:
        """``render_with_depth`` of the reference backend (gsplat_backend.py:129-210) on the raw store.

        accumulate_grads=True: ``loss.backward()`` adds the parameter gradients into ``flat_grad`` (i.e. into every
        ``param.grad``) inside the backward kernel.  False: gradients flow through autograd as usual."""
        W, H = int(viewpoint["image_width"]), int(viewpoint["image_height"])
        renders, alphas, meta = rasterization(
            means=self.xyz, quats=self.rotation, scales=self.scaling, opacities=self.opacity.view(-1),
            colors=self.features, viewmats=viewpoint["world_view_transform"][None], Ks=K[None], width=W, height=H,
            sh_degree=sh_degree, render_mode=render_mode,
            backgrounds=None if bg_color is None else bg_color[None], raw_params=True,
            grad_into=self.grad_buffers() if (accumulate_grads and torch.is_grad_enabled()) else None)
        out, alpha = renders[0], alphas[0]
        result = {"alpha": alpha, "means2d": meta["means2d"][0], "radii": meta["radii"][0]}
        result["visibility_filter"] = result["radii"] > 0
        if render_mode == "RGB":
            result["render"] = out
        elif render_mode == "D":
            result["accumulated_depth"] = out.squeeze(-1)
        elif render_mode == "ED":
            result["expected_depth"] = out.squeeze(-1)
        else:
            result["render"] = out[..., :3]
            result["accumulated_depth" if render_mode == "RGB+D" else "expected_depth"] = out[..., 3]
        return result
