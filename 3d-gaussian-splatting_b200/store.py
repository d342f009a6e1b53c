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

from .rasterization import _render_one, rasterization

_BLOCKS = (("xyz", 3), ("scaling", 3), ("rotation", 4), ("opacity", 1), ("features", 48))
FLOATS_PER_GAUSSIAN = sum(w for _, w in _BLOCKS)  # 59


def _align4(n: int) -> int:
    return (n + 3) // 4 * 4  # floats -> 16-byte multiples


class GradState:
    """Per-gradient-buffer state shared with the backward pass (lazy zero flag, event of the last in-place accumulation).
    Deliberately NOT a Mapping / Iterable: under autocast, custom_fwd's argument cast rebuilds dicts, and the backward
    would then flip ``pending_zero`` on a copy."""

    def __init__(self, **kw):
        self._d = dict(kw)

    def get(self, key, default=None):
        return self._d.get(key, default)

    def __getitem__(self, key):
        return self._d[key]

    def __setitem__(self, key, value):
        self._d[key] = value


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
        # bumped on every (re-)allocation: whoever holds buffers laid out like `flat` (StoreAdam's moments, ViewStreams'
        # accumulator nodes, optimisers built from optimizer_groups) can tell that the population changed under them
        self.generation = getattr(self, "generation", -1) + 1
        group = getattr(self, "_sym_group", None)
        if group is None:
            self._sym_flat = self._sym_grad = None
            self.flat = torch.zeros(o, dtype=torch.float32, device=dev)
            self._flat_grad = torch.zeros(o, dtype=torch.float32, device=dev)
        else:
            # parameters and gradients replicated in SYMMETRIC memory (one multicast range over NVSwitch): the gradient
            # all-reduce / the fused reduce-scatter + Adam + all-gather are this library's multimem kernels
            # (distributed.SymmetricBuffer, csrc/collective.cu).  Collective: every rank allocates the same size.
            from .distributed import SymmetricBuffer
            self._sym_flat = SymmetricBuffer(o, dev, group)
            self._sym_grad = SymmetricBuffer(o, dev, group)
            self.flat, self._flat_grad = self._sym_flat.tensor, self._sym_grad.tensor
        # shared with the backward pass (rasterization.RasterizeSplats): True after zero_grad(lazy=True) until the first
        # backward has WRITTEN the buffer instead of adding into it
        self._grad_state = GradState(pending_zero=False)

        def view(buf, name, *shape):
            w = dict(_BLOCKS)[name]
            return buf[offs[name]: offs[name] + n * w].view(n, *shape)

        feats = view(self.flat, "features", 16, 3)
        gfeats = view(self._flat_grad, "features", 16, 3)
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
        self.xyz.grad = view(self._flat_grad, "xyz", 3)
        self.scaling.grad = view(self._flat_grad, "scaling", 3)
        self.rotation.grad = view(self._flat_grad, "rotation", 4)
        self.opacity.grad = view(self._flat_grad, "opacity", 1)
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

    @torch.no_grad()
    def make_symmetric(self, group=None) -> "GaussianStore":
        """Move the flat parameter and gradient buffers into symmetric (multicast) memory shared with the ranks of
        ``group`` — collective, every rank with the same population.  Afterwards ``allreduce_grads`` and
        ``StoreAdam(sharded=True)`` use this library's multimem kernels instead of NCCL; ``densify_and_prune`` re-allocates
        symmetrically (call it on every rank, as the replicated training loop does anyway).  The ``Parameter`` objects are
        re-created (``generation`` is bumped)."""
        import torch.distributed as dist
        old = self.flat
        self._sym_group = group if group is not None else dist.group.WORLD
        stats = (self.xyz_gradient_accum, self.xyz_gradient_count, self.max_radii_2D)
        self._allocate(self._n, old.device)
        self.flat.copy_(old)
        self.xyz_gradient_accum, self.xyz_gradient_count, self.max_radii_2D = stats
        return self

    def allreduce_grads(self, scale: float) -> torch.Tensor:
        """flat_grad <- scale * sum over ranks: the library's multimem kernel when the store is symmetric, NCCL otherwise."""
        from .distributed import allreduce_flat
        g = self.flat_grad  # (resolves a pending lazy zero)
        if self._sym_grad is not None:
            return self._sym_grad.allreduce_(scale)
        return allreduce_flat(g, scale)

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
        return self._flat_grad[o: o + self._n * w].view(self._n, *self._SHAPES[name])

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

    @property
    def flat_grad(self) -> torch.Tensor:
        """The flat gradient buffer (what the all-reduce and the optimiser take).  Reading it after ``zero_grad(lazy=True)``
        with no backward in between performs the postponed memset, so a reader never sees stale values."""
        if self._grad_state["pending_zero"]:
            self._flat_grad.zero_()
            self._grad_state["pending_zero"] = False
        return self._flat_grad

    def zero_grad(self, set_to_none: bool = False, lazy: bool = False) -> None:
        """One memset of the flat gradient buffer; the ``.grad`` views stay attached (``set_to_none`` is ignored:
        the kernels accumulate into the buffer, not into tensors autograd hands out).

        lazy=True postpones the memset: the first backward of a ``render(accumulate_grads=True)`` that runs afterwards
        WRITES the buffer (zeros for culled Gaussians and inactive SH coefficients included) and the following ones add
        to it — no 236 B/Gaussian fill and no read half of the first view's read-modify-write.  Until then the
        ``param.grad`` views hold the previous step's values; ``flat_grad`` (property) materialises the zeros on demand.
        Use it only when a ``render`` backward is the first writer of the step: plain autograd contributions to a
        ``param.grad`` (e.g. a regulariser's ``backward()``) must come after it, or they are overwritten."""
        self._attach_grads()
        if lazy:
            self._grad_state["pending_zero"] = True
            return
        self._flat_grad.zero_()
        self._grad_state["pending_zero"] = False

    def grad_buffers(self):
        """(g_xyz, g_rotation, g_scaling, g_opacity, g_features, state) — the kernels' argument order, then the
        lazy-zero state the backward consults (see ``zero_grad``)."""
        if self._attach_grads():
            # an optimiser's zero_grad(set_to_none=True) dropped the views: honour it as "gradients are zero"
            self._flat_grad.zero_()
            self._grad_state["pending_zero"] = False
        return (self._grad_view("xyz"), self._grad_view("rotation"), self._grad_view("scaling"),
                self._grad_view("opacity").view(-1), self._grad_view("features"), self._grad_state)

    # ------------------------------------------------------------------ rendering
    def render(self, K: torch.Tensor, viewpoint: Dict, bg_color: Optional[torch.Tensor] = None,
               render_mode: str = "RGB", sh_degree: int = 3, accumulate_grads: bool = True) -> Dict[str, torch.Tensor]:
        """``render_with_depth`` of the reference backend (gsplat_backend.py:129-210) on the raw store.

        accumulate_grads=True: ``loss.backward()`` adds the parameter gradients into ``flat_grad`` (i.e. into every
        ``param.grad``) inside the backward kernel.  False: gradients flow through autograd as usual."""
        W, H = int(viewpoint["image_width"]), int(viewpoint["image_height"])
        viewmat = viewpoint["world_view_transform"]
        grad_into = self.grad_buffers() if (accumulate_grads and torch.is_grad_enabled()) else None
        if render_mode not in ("RGB", "D", "ED", "RGB+D", "RGB+ED"):
            raise ValueError(f"render_mode must be one of RGB, D, ED, RGB+D, RGB+ED, got {render_mode!r}")
        if viewmat.dim() == 2 and (bg_color is None or bg_color.dim() == 1):
            # the same call without the round trip through the gsplat-shaped signature ([None] in, squeeze(0) / [0] out)
            out, alpha, aux = _render_one(self.xyz, self.rotation, self.scaling, self.opacity.view(-1), self.features, viewmat, K,
                                          W, H, sh_degree, render_mode, bg_color, 0.01, 1e10, 0.3, 0.0, False, "auto", True,
                                          grad_into)
            result = {"alpha": alpha, "means2d": aux["splats"][:, 0:2], "radii": aux["radii"],
                      "visibility_filter": aux["visibility"]}  # = radii > 0, from the projection kernel
        else:
            renders, alphas, meta = rasterization(
                means=self.xyz, quats=self.rotation, scales=self.scaling, opacities=self.opacity.view(-1),
                colors=self.features, viewmats=viewmat[None], Ks=K[None], width=W, height=H,
                sh_degree=sh_degree, render_mode=render_mode,
                backgrounds=None if bg_color is None else bg_color[None], raw_params=True, grad_into=grad_into)
            out, alpha = renders.squeeze(0), alphas.squeeze(0)  # views both ways (select's backward = zeros + full-image copy)
            result = {"alpha": alpha, "means2d": meta["means2d"][0], "radii": meta["radii"][0]}
            result["visibility_filter"] = meta["visibility_filter"][0]  # = radii > 0, from the projection kernel
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

    # ------------------------------------------------------------------ densification
    def _layout_array(self):
        import ctypes as C
        o = self._offsets
        return (C.c_int64 * 5)(o["xyz"], o["scaling"], o["rotation"], o["opacity"], o["features"])

    @torch.no_grad()
    def _rebuild(self, src_rows: torch.Tensor, keep_state: torch.Tensor, optimizer=None) -> None:
        """New population: row j <- old row src_rows[j]; optimiser moments follow when keep_state[j] (else zero)."""
        from ._lib import lib, check
        dev = self.flat.device
        old_flat, old_layout = self.flat, self._layout_array()
        old_radii = self.max_radii_2D
        n_new = int(src_rows.numel())
        m_old = v_old = None
        if optimizer is not None:
            if getattr(optimizer, "sharded", False):
                optimizer.gather_moments()  # every rank holds only its slice of the moments: complete them first
            m_old, v_old = optimizer.exp_avg, optimizer.exp_avg_sq
        self._allocate(n_new, dev)  # fresh flat / flat_grad, parameters re-created as views
        new_moment = getattr(optimizer, "_new_moment", None) or (lambda flat: torch.zeros_like(flat))
        m_new = new_moment(self.flat) if optimizer is not None else None
        v_new = new_moment(self.flat) if optimizer is not None else None
        src_rows = src_rows.to(torch.int64).contiguous()
        ks = keep_state.to(torch.uint8).contiguous()
        ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        with torch.cuda.device(dev):
            check(lib.b200_store_gather(old_flat.data_ptr(), ptr(m_old), ptr(v_old), old_layout, self.flat.data_ptr(),
                                        ptr(m_new), ptr(v_new), self._layout_array(), src_rows.data_ptr(), ks.data_ptr(),
                                        n_new, torch.cuda.current_stream().cuda_stream), "b200_store_gather")
        if optimizer is not None:
            optimizer.exp_avg, optimizer.exp_avg_sq = m_new, v_new
            optimizer._generation = self.generation
        self.max_radii_2D = torch.where(keep_state.bool(), old_radii[src_rows], torch.zeros((), device=dev))
        self.xyz_gradient_accum = torch.zeros(n_new, 3, device=dev)
        self.xyz_gradient_count = torch.zeros(n_new, 1, device=dev)

    @torch.no_grad()
    def densify_and_prune(self, grads_threshold: float, min_opacity: float, extent: float, max_screen_size: float = None,
                          optimizer=None, clone_extent_ratio: float = 0.1, prune_extent_ratio: float = 2.0,
                          max_gaussians: int = None) -> Dict:
        """gaussian_model.py:110-231 — same arguments, same statistics dict, same resulting population (row order,
        values, spliced Adam moments; the split noise is drawn with the same ``torch.randn_like`` call on the same
        shape), but decided and moved ON THE DEVICE with ONE host read-back:

        * the clone / split / prune decisions are element-wise masks over the N Gaussians; the population cap
          (``_top_k_mask_by_grad``, gaussian_model.py:233-249) is a rank threshold from one stable descending sort of
          the candidates' gradient norms — the reference's ``torch.topk`` with k in the hundred-thousands is a 40-60 ms
          call at 3 M Gaussians, the sort ~1 ms;
        * the reference builds the new population in three rounds ([originals | clones] -> minus the split ones, plus
          2 x children -> minus the pruned ones), each a chain of ``cat`` / mask copies over six tensors and two
          optimiser buffers.  Here the final population is addressed directly: a prefix sum over the four candidate
          segments [kept originals | clones | children #1 | children #2] with the prune decision folded in gives every
          survivor its final row, the sizes come back in one 7-integer read, and ``b200_store_gather`` moves
          parameters and both moments once.

        ``optimizer``: a :class:`StoreAdam`."""
        n0 = self._n
        dev = self.flat.device
        stats = {"cloned": 0, "split": 0, "pruned": 0, "n_before": n0, "capped": False}
        grads = self.xyz_gradient_accum / (self.xyz_gradient_count + 1e-8)
        grads_norm = torch.norm(grads, dim=-1)
        max_scale = self.get_scaling.amax(dim=-1)
        opacity = self.get_opacity.reshape(-1)
        hot = (grads_norm >= grads_threshold) & (opacity > min_opacity)
        clone_mask = hot & (max_scale <= extent * clone_extent_ratio)
        split_mask = hot & (max_scale > extent * clone_extent_ratio)
        n_comb = torch.zeros((), dtype=torch.int64, device=dev)
        n_allowed = None
        if max_gaussians is not None and n0 >= int(max_gaussians):
            clone_mask = torch.zeros_like(clone_mask)
            split_mask = torch.zeros_like(split_mask)
            stats["capped"] = True
        elif max_gaussians is not None:
            # keep the n_allowed candidates with the largest gradient norm.  When there are no more than n_allowed of
            # them every candidate ranks below n_allowed and the masks are unchanged, so no branch on a device value
            n_allowed = int(max_gaussians) - n0
            combined = clone_mask | split_mask
            n_comb = combined.sum()
            order = torch.argsort(torch.where(combined, grads_norm, torch.full_like(grads_norm, float("-inf"))),
                                  descending=True, stable=True)
            keep = torch.zeros_like(combined)
            keep[order[:max(n_allowed, 0)]] = True
            clone_mask, split_mask = clone_mask & keep, split_mask & keep
        # prune decision on the post-densify population (gaussian_model.py:190-213): clones carry their source's values,
        # children the source's opacity and its log-scales minus log 1.6 (gaussian_model.py:299-300)
        child_scaling_all = self.scaling - torch.log(torch.tensor(1.6, device=dev))
        child_scales_all = torch.exp(child_scaling_all)
        low_op = opacity < min_opacity
        big_o = max_scale > extent * prune_extent_ratio
        big_c = child_scales_all.amax(dim=-1) > extent * prune_extent_ratio
        prune_o, prune_c = low_op | big_o, low_op | big_c
        stay = ~split_mask
        segs = (stay, clone_mask, split_mask, split_mask)  # [kept originals | clones | children #1 | children #2]
        prunes = (prune_o, prune_o, prune_c, prune_c)
        flags = torch.cat([m & ~p for m, p in zip(segs, prunes)])
        csum = torch.cumsum(flags.to(torch.int32), dim=0)
        cnt = lambda m: m.sum()  # noqa: E731
        counts = torch.stack([
            csum[-1].to(torch.int64) if n0 > 0 else torch.zeros((), dtype=torch.int64, device=dev),
            cnt(clone_mask), cnt(split_mask),
            cnt(stay & prune_o) + cnt(clone_mask & prune_o) + 2 * cnt(split_mask & prune_c),
            cnt(stay & low_op) + cnt(clone_mask & low_op) + 2 * cnt(split_mask & low_op),
            cnt(stay & big_o) + cnt(clone_mask & big_o) + 2 * cnt(split_mask & big_c),
            n_comb,
        ])
        n_new, n_c, n_s, n_pruned, n_low, n_big, n_cand = (int(x) for x in counts.tolist())  # THE host read-back
        stats["cloned"], stats["split"], stats["pruned"] = n_c, n_s, n_pruned
        stats["pruned_low_op"] = n_low if n_pruned else 0
        stats["pruned_too_big"] = n_big if n_pruned else 0
        if n_allowed is not None and n_cand > n_allowed:
            stats["capped"] = True
        if n_c == 0 and n_s == 0 and n_pruned == 0:
            self.xyz_gradient_accum.zero_()
            self.xyz_gradient_count.zero_()
            stats["n_after"] = n0
            return stats
        # final row of every survivor (dump slot n_new for the rest), then one gather
        idx4 = torch.arange(n0, device=dev).repeat(4)
        dst = torch.where(flags, csum.to(torch.int64) - 1, torch.full((), n_new, dtype=torch.int64, device=dev))
        src_rows = torch.empty(n_new + 1, dtype=torch.int64, device=dev).scatter_(0, dst, idx4)[:n_new]
        seg4 = torch.arange(4, device=dev, dtype=torch.uint8).repeat_interleave(n0)
        seg_new = torch.empty(n_new + 1, dtype=torch.uint8, device=dev).scatter_(0, dst, seg4)[:n_new]
        old_xyz = self.xyz.detach()
        split_ord = torch.cumsum(split_mask.to(torch.int64), dim=0) - 1  # j-th split Gaussian in index order
        self._rebuild(src_rows, seg_new == 0, optimizer)
        if n_s > 0:  # _split_gaussians (gaussian_model.py:278-312): same arithmetic, the same randn call
            noise = torch.randn(2 * n_s, 3, device=dev, dtype=old_xyz.dtype)  # == randn_like(xyz[idx].repeat(2, 1))
            is_child = seg_new >= 2
            row = (split_ord[src_rows] + (seg_new.to(torch.int64) - 2).clamp_(min=0) * n_s).clamp_(min=0)
            child_xyz = old_xyz[src_rows] + noise[row] * child_scales_all[src_rows]
            self.xyz.data.copy_(torch.where(is_child[:, None], child_xyz, self.xyz.data))
            self.scaling.data.copy_(torch.where(is_child[:, None], child_scaling_all[src_rows], self.scaling.data))
        stats["n_after"] = self._n
        return stats

    # ------------------------------------------------------------------ periodic opacity reset
    @torch.no_grad()
    def reset_opacity(self, ceiling: float = 0.01) -> None:
        """trainer.py:841-850: every ``opacity_reset_interval`` iterations every opacity is pulled down to at most
        ``ceiling`` (``logit(clamp(min(sigmoid(o), ceiling), 1e-6, 1 - 1e-6))``), in place on the flat buffer; the
        optimiser moments are left alone, as in the reference."""
        target = torch.minimum(self.get_opacity, torch.full_like(self.opacity, ceiling)).clamp(min=1e-6, max=1.0 - 1e-6)
        self.opacity.copy_(torch.logit(target))

    # ------------------------------------------------------------------ checkpoint / PLY wire format (§8 "later" row)
    _REF_KEYS = ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")

    def reference_state_dict(self) -> Dict[str, torch.Tensor]:
        """The reference's ``GaussianModel.state_dict()`` (``model_state`` of trainer.py:384-395): six tensors under the
        reference's names and shapes, detached copies."""
        return {k: getattr(self, k).detach().clone() for k in self._REF_KEYS}

    @classmethod
    def from_reference_state_dict(cls, sd: Dict[str, torch.Tensor], device: str = "cuda") -> "GaussianStore":
        """Build a store from a reference checkpoint's ``model_state`` (or a dict wrapping it, export.py:15-23)."""
        for k in ("model_state", "state_dict"):
            if k in sd and isinstance(sd[k], dict):
                sd = sd[k]
        missing = [k for k in cls._REF_KEYS if k not in sd]
        if missing:
            raise KeyError(f"reference state dict lacks {missing}")
        return cls.from_tensors(*(sd[k].to(device=device, dtype=torch.float32) for k in
                                  ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")))

    _PLY_PROPS = (["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2", "opacity", "scale_0", "scale_1",
                   "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"] + [f"f_rest_{i}" for i in range(45)])

    def export_ply(self, path: str) -> None:
        """export.py:53-139 — binary little-endian PLY, one ``vertex`` element of 62 float32 properties per
        Gaussian in the reference's order: position, zero normal, ``features_dc`` (3), opacity logit, log-scales (3),
        quaternion (4), then ``features_rest.reshape(N, 45)`` (coefficient-major, channel innermost — the reference's
        own flattening, export.py:33-42).  The header is what ``plyfile`` writes for that structured array."""
        import numpy as np
        n = self._n
        with torch.no_grad():
            rows = torch.cat([self.xyz, torch.zeros_like(self.xyz), self.features[:, 0, :], self.opacity, self.scaling,
                              self.rotation, self.features[:, 1:, :].reshape(n, 45)], dim=1)
        data = np.ascontiguousarray(rows.to(torch.float32).cpu().numpy().astype("<f4"))
        header = ["ply", "format binary_little_endian 1.0", f"element vertex {n}"]
        header += [f"property float {p}" for p in self._PLY_PROPS] + ["end_header"]
        with open(path, "wb") as f:
            f.write(("\n".join(header) + "\n").encode("ascii"))
            f.write(data.tobytes())

    @classmethod
    def from_ply(cls, path: str, device: str = "cuda") -> "GaussianStore":
        """Inverse of :meth:`export_ply` (also reads files without the optional ``f_rest_*`` block)."""
        import numpy as np
        with open(path, "rb") as f:
            if f.readline().strip() != b"ply":
                raise ValueError("not a PLY file")
            props, n = [], None
            while True:
                line = f.readline()
                if not line:
                    raise ValueError("PLY header has no end_header")
                tok = line.decode("ascii").split()
                if tok[:1] == ["format"] and tok[1] != "binary_little_endian":
                    raise ValueError(f"unsupported PLY format {tok[1]}")
                if tok[:2] == ["element", "vertex"]:
                    n = int(tok[2])
                elif tok[:1] == ["property"]:
                    if tok[1] not in ("float", "float32"):
                        raise ValueError(f"unsupported PLY property type {tok[1]}")
                    props.append(tok[2])
                elif tok[:1] == ["end_header"]:
                    break
            if n is None:
                raise ValueError("PLY file has no vertex element")
            raw = np.frombuffer(f.read(n * len(props) * 4), dtype="<f4")
        if raw.size != n * len(props):
            raise ValueError("PLY vertex data is truncated")
        t = torch.from_numpy(raw.reshape(n, len(props)).copy())
        col = {p: i for i, p in enumerate(props)}
        take = lambda names: t[:, [col[p] for p in names]]  # noqa: E731
        rest = take([f"f_rest_{i}" for i in range(45)]) if "f_rest_0" in col else torch.zeros(n, 45)
        return cls.from_tensors(take(["x", "y", "z"]).to(device), take(["f_dc_0", "f_dc_1", "f_dc_2"]).to(device),
                                rest.to(device), take(["scale_0", "scale_1", "scale_2"]).to(device),
                                take(["rot_0", "rot_1", "rot_2", "rot_3"]).to(device), take(["opacity"]).to(device))
