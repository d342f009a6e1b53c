"""Multi-GPU data path: camera views sharded over ranks, Gaussians replicated, ONE all-reduce of a flat
gradient buffer per step (BASELINE.json north_star).

The reference has no distributed code; the contract here is "the reduced gradient equals
the single-process gradient of the mean loss over the same global batch of views"
(src/training/trainer.py:458-500 sums per-view losses and divides by the batch size).  One process
per GPU, ``torch.distributed`` (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Callable, Iterable, List, Sequence

import torch
import torch.distributed as dist


def shard_views(n_views: int, rank: int, world_size: int) -> List[int]:
    """Round-robin view indices of this rank (balanced to within one view for any n_views)."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return list(range(rank, n_views, world_size))


def balance_views(costs: Sequence[float], world_size: int, per_rank: int = 0) -> List[List[int]]:
    """Assign views to ranks so that the per-rank sums of ``costs`` (e.g. last step's intersection count M of each
    camera) are as equal as greedy longest-processing-time-first gets them; ``per_rank`` > 0 caps the number of views a
    rank may take (weak scaling: exactly ``per_rank`` each when ``len(costs) == world_size * per_rank``).  Deterministic:
    every rank computes the same assignment from the same costs.  A step ends at the slowest rank, so the spread of the
    per-rank sums is what the all-reduce waits for."""
    n = len(costs)
    cap = per_rank if per_rank > 0 else n
    if world_size * cap < n:
        raise ValueError("balance_views: world_size * per_rank < number of views")
    order = sorted(range(n), key=lambda i: (-float(costs[i]), i))
    load = [0.0] * world_size
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min((r for r in range(world_size) if len(out[r]) < cap), key=lambda r: (load[r], r))
        out[r].append(i)
        load[r] += float(costs[i])
    return [sorted(v) for v in out]


def allreduce_gradients(params: Iterable[torch.Tensor], n_global_views: int, group=None,
                        flat: torch.Tensor = None) -> torch.Tensor:
    """Sum-all-reduce every ``p.grad`` and scale by 1/n_global_views, so the result is the gradient of the MEAN
    per-view loss.  Local gradients must be those of the SUM of the rank's per-view losses.

    When the gradients are views of ONE flat buffer (``GaussianStore.flat_grad``, or ``flat=`` from a previous call
    whose views are still attached) that buffer is reduced by a single collective with no copies.  Otherwise the
    per-attribute gradient tensors are reduced in place by collectives issued back to back (one NCCL group on the
    wire) and scaled by one fused multiply: no flatten / unflatten passes over the 59*N floats.
    Returns the flat buffer if one was used, else None."""
    params = [p for p in params if p.requires_grad]
    for p in params:
        if p.grad is None:
            p.grad = torch.zeros_like(p)
    grads = [p.grad for p in params]
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    scale = 1.0 / float(n_global_views)
    op = _scaled_sum_op(scale, grads[0], group) if multi else None  # NCCL: the 1/V scale rides inside the collective
    if flat is not None and _views_of(flat, grads):
        if multi:
            dist.all_reduce(flat, op=op or dist.ReduceOp.SUM, group=group)
        if op is None:
            flat.mul_(scale)
        return flat
    if multi:
        _allreduce_many(grads, group, op or dist.ReduceOp.SUM)
    if op is None:
        torch._foreach_mul_(grads, scale)
    return None


class SymmetricBuffer:
    """A flat fp32 buffer replicated in SYMMETRIC memory on every rank of ``group`` (one multicast address range over
    NVSwitch) whose sum-all-reduce is this library's own kernel (``b200_allreduce_multimem``: ``multimem.ld_reduce`` /
    ``multimem.st``, csrc/collective.cu) instead of an NCCL call.  Allocation, rendezvous and the cross-rank barriers are
    host plumbing from ``torch.distributed._symmetric_memory``; construction is collective (every rank, same ``numel``).
    Raises if the fabric has no multicast support — there is no silent fall-back; callers choose NCCL explicitly."""

    def __init__(self, numel: int, device, group=None, max_ctas: int = 0):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        n = (int(numel) + 3) // 4 * 4
        self.tensor = symm_mem.empty(n, dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.tensor, self.group)
        self.mc_ptr = int(self.handle.multicast_ptr)
        if self.mc_ptr == 0:
            raise RuntimeError("SymmetricBuffer: no multicast (NVLink SHARP) support on this fabric")
        self.rank, self.world, self.max_ctas = dist.get_rank(self.group), dist.get_world_size(self.group), int(max_ctas)
        self.tensor.zero_()

    def allreduce_(self, scale: float = 1.0) -> torch.Tensor:
        """tensor <- scale * sum over ranks, on every rank (bit-identical replicas), on the current stream."""
        from ._lib import lib, check
        t = self.tensor
        with torch.cuda.device(t.device):
            self.handle.barrier(channel=0)  # every rank's backward has finished writing its replica
            check(lib.b200_allreduce_multimem(self.mc_ptr, t.numel(), self.rank, self.world, float(scale), self.max_ctas,
                                              torch.cuda.current_stream().cuda_stream), "b200_allreduce_multimem")
            self.handle.barrier(channel=1)  # every slice is stored before anyone reads its replica
        return t


class GradientArena:
    """Drop-in path, N GPUs: make autograd's own `.grad` tensors live inside ONE flat buffer in symmetric memory WITHOUT the
    bookkeeping cost of :class:`FlatGradients` (no memset, no extra in-place add of the first view).

    The rasterizer's backward writes every row of its five gradient outputs (zeros for culled Gaussians included).  Once per
    step — for whichever view's backward runs first — ``take`` hands it slices of the arena as those outputs; autograd adopts
    a fresh, contiguous incoming gradient of a leaf as its ``.grad`` without copying, and accumulates the later views into it
    in place.  ``allreduce`` then checks that every ``p.grad`` really is the arena's slice (it is not if the inputs were not
    leaves, or a hook cloned the gradient) and reduces the whole buffer with ONE multimem kernel; anything that did not land
    in the arena falls back to NCCL.  Activate with ``rasterization.set_gradient_arena(arena)``."""

    def __init__(self, like: Sequence[torch.Tensor], group=None):
        """like: tensors with the shapes of (means, quats, scales, opacities, colors) — e.g. the five leaves themselves."""
        self.shapes = [tuple(t.shape) for t in like]
        sizes = [(t.numel() + 3) // 4 * 4 for t in like]
        self.sym = SymmetricBuffer(sum(sizes), like[0].device, group)
        self.views, self._spans, o = [], [], 0
        for t, n in zip(like, sizes):
            self.views.append(self.sym.tensor[o: o + t.numel()].view(t.shape))
            self._spans.append((o, t.numel(), tuple(t.shape)))
            o += n
        self._available = False

    def begin_step(self) -> None:
        self._available = True

    def take(self, inputs: Sequence[torch.Tensor]):
        if not self._available or [tuple(t.shape) for t in inputs] != self.shapes or inputs[0].device != self.sym.tensor.device:
            return None
        self._available = False
        # FRESH view objects: autograd adopts an incoming gradient as .grad without a copy only if nothing else references
        # that tensor object (AccumulateGrad's use-count test), so the arena's own `views` must not be what is handed out
        return tuple(self.sym.tensor[o: o + n].view(shape) for o, n, shape in self._spans)

    def allreduce(self, params: Sequence[torch.Tensor], n_global_views: int, group=None) -> bool:
        """params in the order (means, quats, scales, opacities, colors).  True if the own collective was used."""
        inside = all(p.grad is not None and p.grad.data_ptr() == v.data_ptr() and p.grad.is_contiguous()
                     for p, v in zip(params, self.views))
        if inside:
            self.sym.allreduce_(1.0 / float(n_global_views))
            return True
        allreduce_gradients(list(params), n_global_views, group)
        return False


class FlatGradients:
    """The parameters' ``.grad`` as views of ONE flat buffer (what DDP calls ``gradient_as_bucket_view``), for the drop-in
    path, whose five gradient tensors autograd otherwise allocates separately: ``begin_step()`` zeroes the buffer and
    re-attaches the views, autograd accumulates every view's gradients into them in place, ``allreduce(n_views)`` is then
    ONE collective instead of five back-to-back ones (three of which — 4 to 16 MB — are latency-bound at 8 ranks)."""

    def __init__(self, params: Sequence[torch.Tensor], symmetric: bool = False, group=None):
        """symmetric=True: the buffer lives in symmetric memory and is reduced by this library's multimem kernel
        (:class:`SymmetricBuffer`) instead of NCCL — collective construction."""
        self.params = [p for p in params if p.requires_grad]
        sizes = [(p.numel() + 3) // 4 * 4 for p in self.params]  # 16-byte aligned slices
        self.sym = SymmetricBuffer(sum(sizes), self.params[0].device, group) if symmetric else None
        self.flat = self.sym.tensor if symmetric else torch.zeros(sum(sizes), dtype=self.params[0].dtype,
                                                                   device=self.params[0].device)
        self.views, o = [], 0
        for p, n in zip(self.params, sizes):
            self.views.append(self.flat[o: o + p.numel()].view_as(p))
            o += n

    def begin_step(self) -> None:
        self.flat.zero_()
        for p, v in zip(self.params, self.views):
            p.grad = v

    def allreduce(self, n_global_views: int, group=None) -> torch.Tensor:
        for p, v in zip(self.params, self.views):
            if p.grad is None or p.grad.data_ptr() != v.data_ptr():
                raise RuntimeError("FlatGradients: a .grad was replaced since begin_step() (zero_grad(set_to_none=True)?)")
        if self.sym is not None:
            return self.sym.allreduce_(1.0 / float(n_global_views))
        return allreduce_flat(self.flat, 1.0 / float(n_global_views), group)


def allreduce_flat(flat: torch.Tensor, scale: float, group=None) -> torch.Tensor:
    """``flat <- scale * sum_over_ranks(flat)`` in place with ONE collective (``GaussianStore.flat_grad``)."""
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    op = _scaled_sum_op(scale, flat, group) if multi else None
    if multi:
        dist.all_reduce(flat, op=op or dist.ReduceOp.SUM, group=group)
    if op is None and scale != 1.0:
        flat.mul_(scale)
    return flat


def _scaled_sum_op(scale: float, like: torch.Tensor, group=None):
    """NCCL's pre-multiplied sum (every rank's contribution is multiplied by ``scale`` on its way into the
    reduction), which removes the separate pass over the 59*N floats that scales the reduced gradient.  Other
    backends (gloo in the CPU tests) have no such operator: None, and the caller multiplies afterwards."""
    if not like.is_cuda or dist.get_backend(group) != "nccl" or not hasattr(dist, "_make_nccl_premul_sum"):
        return None
    return dist._make_nccl_premul_sum(float(scale))


def _allreduce_many(tensors: Sequence[torch.Tensor], group=None, op=dist.ReduceOp.SUM) -> None:
    """In-place all-reduce of several tensors: asynchronous collectives issued back to back, no flatten copy.
    (torch's coalescing manager was measured 2x slower per step here than five plain collectives at N=2.)"""
    works = [dist.all_reduce(t, op=op, group=group, async_op=True) for t in tensors]
    for w in works:
        w.wait()


def _views_of(flat: torch.Tensor, grads: Sequence[torch.Tensor]) -> bool:
    """True if every gradient is a contiguous, non-overlapping slice of ``flat`` (padding between slices allowed —
    GaussianStore aligns its blocks to 16 bytes; the padding holds zeros on every rank, so reducing it is harmless)."""
    es = flat.element_size()
    base, end = flat.data_ptr(), flat.data_ptr() + flat.numel() * es
    spans = []
    for g in grads:
        if not g.is_contiguous() or g.dtype != flat.dtype:
            return False
        lo, hi = g.data_ptr(), g.data_ptr() + g.numel() * es
        if not (base <= lo and hi <= end):
            return False
        spans.append((lo, hi))
    spans.sort()
    return all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))


class ViewShardedStep:
    """One optimisation step over a global batch of views, this rank rendering only its shard.

    ``render_loss(view) -> scalar loss`` must build the autograd graph from ``params``.  After
    ``__call__`` every rank holds identical ``p.grad`` = d(mean_v loss_v)/dp, so identical optimiser
    steps keep the replicas bit-identical without a parameter broadcast.  The densification statistic
    is a norm of the REDUCED gradient and therefore must be taken after this call."""

    def __init__(self, params: Sequence[torch.Tensor], group=None):
        self.params = list(params)
        self.group = group

    def __call__(self, views: Sequence, render_loss: Callable[[object], torch.Tensor]) -> float:
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        for p in self.params:
            p.grad = None
        local_sum = 0.0
        mine = shard_views(len(views), rank, world)
        for i in mine:
            loss = render_loss(views[i])
            loss.backward()
            local_sum += float(loss.detach())
        allreduce_gradients(self.params, len(views), self.group)
        t = torch.tensor([local_sum], dtype=torch.float64, device=self.params[0].device)
        if world > 1:
            dist.all_reduce(t, group=self.group)
        return float(t.item()) / len(views)
