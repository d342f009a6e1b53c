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
                 scale_ceiling: Optional[float] = None):
        self.store = store
        self.scale_ceiling = scale_ceiling
        self.lrs = {_ALIASES.get(k, k): float(v) for k, v in lrs.items()}
        missing = {"xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"} - set(self.lrs)
        if missing:
            raise ValueError(f"StoreAdam: missing learning rates for {sorted(missing)}")
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.exp_avg = torch.zeros_like(store.flat)
        self.exp_avg_sq = torch.zeros_like(store.flat)
        self._generation = store.generation
        self.t = 0

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
