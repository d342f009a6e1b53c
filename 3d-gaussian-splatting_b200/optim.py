"""StoreAdam — the reference's optimiser (``Adam(params, lr=0.0, eps=1e-15)`` over six parameter groups,
src/training/trainer.py:331-340) as one kernel over the flat GaussianStore buffers.

``lrs`` uses the reference's group names (``xyz, f_dc, f_rest, opacity, scaling, rotation``; the store's attribute
names ``features_dc`` / ``features_rest`` are accepted too) and stays mutable, so the trainer's per-iteration
position-rate schedule (``_update_learning_rate``, trainer.py:373-383) is ``opt.lrs["xyz"] = lr``.

``scale_ceiling`` (mutable attribute, log-space, None = off) folds the reference's per-iteration hard cap on the
scales, ``gaussians.scaling.data.clamp_(max=math.log(scene_extent * scale_clamp_ratio + 1e-9))``
(trainer.py:854-859), into the same kernel: the updated log-scale is ``min(p, scale_ceiling)``.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import torch

from ._lib import lib, check

_ALIASES = {"features_dc": "f_dc", "features_rest": "f_rest"}


class StoreAdam:
    def __init__(self, store, lrs: Dict[str, float], betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-15,
                 scale_ceiling: Optional[float] = None, sharded: bool = False):
        """sharded=True (multi-GPU, store.make_symmetric() first): ``step_allreduce`` replaces "all-reduce the gradient, then
        every rank runs the same Adam over all 59 N parameters" by ONE kernel per rank over NVSwitch multicast memory —
        reduce-scatter of the gradient (``multimem.ld_reduce``), Adam on the rank's 1/world slice, all-gather of the updated
        PARAMETERS (``multimem.st``).  Each rank then maintains only its slice of the moments (``gather_moments`` completes
        them, e.g. before ``densify_and_prune`` or a checkpoint)."""
        self.store = store
        self.sharded = bool(sharded)
        if self.sharded and getattr(store, "_sym_grad", None) is None:
            raise RuntimeError("StoreAdam(sharded=True) needs a symmetric store: call store.make_symmetric() first")
        self.scale_ceiling = scale_ceiling
        self.lrs = {_ALIASES.get(k, k): float(v) for k, v in lrs.items()}
        missing = {"xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"} - set(self.lrs)
        if missing:
            raise ValueError(f"StoreAdam: missing learning rates for {sorted(missing)}")
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.exp_avg = self._new_moment(store.flat)
        self.exp_avg_sq = self._new_moment(store.flat)
        self._generation = store.generation
        self.t = 0

    def _new_moment(self, flat: torch.Tensor) -> torch.Tensor:
        """A zeroed moment buffer laid out like ``flat`` (sharded: padded to world equal slices, so that the slices can be
        all-gathered in place)."""
        if not self.sharded:
            return torch.zeros_like(flat)
        sb = self.store._sym_grad
        b, e = C.c_int64(0), C.c_int64(0)
        lib.b200_multimem_slice(flat.numel(), 0, sb.world, C.byref(b), C.byref(e))
        return torch.zeros(max(sb.world * (e.value - b.value), flat.numel()), dtype=flat.dtype, device=flat.device)

    def _slice(self):
        sb = self.store._sym_grad
        b, e = C.c_int64(0), C.c_int64(0)
        lib.b200_multimem_slice(self.store.flat.numel(), sb.rank, sb.world, C.byref(b), C.byref(e))
        return b.value, e.value

    @torch.no_grad()
    def gather_moments(self) -> None:
        """sharded: hand every rank's slice of both moments to all ranks (in-place all-gather of equal slices)."""
        if not self.sharded:
            return
        import torch.distributed as dist
        sb = self.store._sym_grad
        b0, e0 = C.c_int64(0), C.c_int64(0)
        lib.b200_multimem_slice(self.store.flat.numel(), 0, sb.world, C.byref(b0), C.byref(e0))
        chunk = e0.value - b0.value
        for buf in (self.exp_avg, self.exp_avg_sq):
            mine = buf[sb.rank * chunk: (sb.rank + 1) * chunk].clone()
            dist.all_gather_into_tensor(buf[: sb.world * chunk], mine, group=sb.group)

    @torch.no_grad()
    def step_allreduce(self, grad_scale: float, densify_stats: bool = False) -> None:
        """sharded: gradient reduce-scatter + Adam on this rank's slice + parameter all-gather in one kernel
        (``b200_allreduce_adam_multimem``), bracketed by cross-rank barriers; ``grad_scale`` = 1 / global number of views (the
        reference's MEAN over the batch, trainer.py:499-500).  Afterwards ``flat_grad`` holds the REDUCED gradient in the
        position block only (what the densification statistic needs); the other blocks keep the rank's local sums.
        Not sharded: ``store.allreduce_grads(grad_scale)`` then ``step``."""
        st = self.store
        if not self.sharded:
            st.allreduce_grads(grad_scale)
            self.step(densify_stats=densify_stats)
            return
        if self._generation != st.generation:
            raise RuntimeError("StoreAdam: the store was re-allocated (densify_and_prune) without optimizer=this StoreAdam")
        self.t += 1
        segs, n_seg = self._segments()
        g = st.flat_grad  # (resolves a pending lazy zero)
        sg, sp = st._sym_grad, st._sym_flat
        n = len(st)
        keep_b = st._offsets["xyz"]
        keep_e = keep_b + (3 * n + 3) // 4 * 4
        with torch.cuda.device(st.flat.device):
            sg.handle.barrier(channel=0)  # every rank's backward has finished writing its gradient replica
            check(lib.b200_allreduce_adam_multimem(sg.mc_ptr, sp.mc_ptr, st.flat.data_ptr(), self.exp_avg.data_ptr(),
                                                   self.exp_avg_sq.data_ptr(), st.flat.numel(), sg.rank, sg.world, float(grad_scale),
                                                   segs, n_seg, self.t, self.betas[0], self.betas[1], self.eps, keep_b, keep_e, 0,
                                                   torch.cuda.current_stream().cuda_stream), "b200_allreduce_adam_multimem")
            sg.handle.barrier(channel=1)  # every slice of the parameters (and of the reduced position gradient) has landed
            if densify_stats:
                check(lib.b200_densify_stats(g.data_ptr() + 4 * keep_b, n, st.xyz_gradient_accum.data_ptr(),
                                             st.xyz_gradient_count.data_ptr(), torch.cuda.current_stream().cuda_stream),
                      "b200_densify_stats")

    def _segments(self):
        st, n = self.store, len(self.store)
        o = st._offsets
        inf = float("inf")
        cap = inf if self.scale_ceiling is None else float(self.scale_ceiling)
        rows = [
            (o["xyz"], o["xyz"] + 3 * n, self.lrs["xyz"], 0.0, 0, 0, inf),
            (o["scaling"], o["scaling"] + 3 * n, self.lrs["scaling"], 0.0, 0, 0, cap),
            (o["rotation"], o["rotation"] + 4 * n, self.lrs["rotation"], 0.0, 0, 0, inf),
            (o["opacity"], o["opacity"] + n, self.lrs["opacity"], 0.0, 0, 0, inf),
            (o["features"], o["features"] + 48 * n, self.lrs["f_dc"], self.lrs["f_rest"], 48, 3, inf),
        ]
        flat = [float(x) for r in rows for x in r]
        return (C.c_double * len(flat))(*flat), len(rows)

    @torch.no_grad()
    def step(self, densify_stats: bool = False) -> None:
        """One Adam step over the flat store.  densify_stats=True also takes the reference's densification statistic
        (trainer.py:786-791: ``accum += ||xyz.grad||``, ``count += 1``, right after ``optimizer.step()``) in the same
        kernel — the position gradient is already being streamed — instead of a separate ``densify_stats`` launch."""
        st = self.store
        # the moments are laid out like `store.flat`: a store that was re-allocated without this optimiser
        # (densify_and_prune(optimizer=None) — the reference's signature default) would be stepped out of bounds or with
        # the wrong Gaussians' moments.  torch's Adam raises in that situation; so does this.
        if self._generation != st.generation or self.exp_avg.numel() != st.flat.numel():
            raise RuntimeError("StoreAdam: the store was re-allocated (densify_and_prune) without optimizer=this StoreAdam; "
                               "pass the optimiser so its moments are spliced, or build a new StoreAdam")
        self.t += 1
        segs, n_seg = self._segments()
        with torch.cuda.device(st.flat.device):
            g = st.flat_grad
            if densify_stats:
                check(lib.b200_adam_step_stats(st.flat.data_ptr(), g.data_ptr(), self.exp_avg.data_ptr(),
                                               self.exp_avg_sq.data_ptr(), st.flat.numel(), segs, n_seg, self.t, self.betas[0],
                                               self.betas[1], self.eps, g.data_ptr() + 4 * st._offsets["xyz"], len(st),
                                               st.xyz_gradient_accum.data_ptr(), st.xyz_gradient_count.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream), "b200_adam_step_stats")
            else:
                check(lib.b200_adam_step(st.flat.data_ptr(), g.data_ptr(), self.exp_avg.data_ptr(),
                                         self.exp_avg_sq.data_ptr(), st.flat.numel(), segs, n_seg, self.t, self.betas[0],
                                         self.betas[1], self.eps, torch.cuda.current_stream().cuda_stream), "b200_adam_step")

    def zero_grad(self, set_to_none: bool = False, lazy: bool = False) -> None:
        self.store.zero_grad(lazy=lazy)  # lazy: see GaussianStore.zero_grad
