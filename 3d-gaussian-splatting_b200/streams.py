"""Views of one batch in flight on alternating CUDA streams.

The reference renders the ``batch_size`` views of a step one after the other on one stream
(src/training/trainer.py:458-500).  The views are independent until their gradients are summed, and a view's
pipeline is a chain of short latency-bound kernels (tile_prep, the depth-sort passes, the scans) followed by two
long issue-bound ones (the blends) whose last wave leaves SMs idle.  Putting consecutive views on alternating
streams lets the short kernels and the drain phases of one view run under the long kernels of the other.
Everything the rasterizer allocates comes from torch's stream-aware caching allocator, the camera constant block
is event-guarded across streams (csrc/project_sh.cu: load_cam / cam_used), and autograd replays every backward
on its forward's stream and orders the gradient accumulation itself, so no kernel changes are involved.
"""
from __future__ import annotations

import contextlib
from typing import List

import torch


class ViewStreams:
    """``with vs.view(i): render + loss + backward`` for the i-th view of the batch; ``vs.begin()`` before the
    first view and ``vs.join()`` after the last make the caller's stream the fork and join point, so whatever
    follows (all-reduce, optimiser, statistics) sees every view's gradients.  ``n_streams <= 1`` is a no-op."""

    def __init__(self, device, n_streams: int = 2, params=()):
        self.device = torch.device(device)
        self.streams: List[torch.cuda.Stream] = (
            [torch.cuda.Stream(device=self.device) for _ in range(n_streams)] if n_streams > 1 else [])
        # autograd runs a leaf's ``grad +=`` on the stream that was current when the leaf's accumulator node was
        # created, and makes that stream wait for the producing backward.  Created lazily inside view 0 that would
        # be a VIEW stream, and every second view's forward would queue behind the other stream's backward.  Create
        # the accumulators here, on the caller's stream, and keep them alive (the leaves hold them weakly).
        self._accumulators = []
        if self.streams:
            for p in params:
                if p.requires_grad and p.is_leaf:
                    self._accumulators.append(p.expand_as(p).grad_fn.next_functions[0][0])

    def begin(self) -> None:
        main = torch.cuda.current_stream(self.device)
        for s in self.streams:
            s.wait_stream(main)

    def view(self, i: int):
        if not self.streams:
            return contextlib.nullcontext()
        return torch.cuda.stream(self.streams[i % len(self.streams)])

    def join(self) -> None:
        main = torch.cuda.current_stream(self.device)
        for s in self.streams:
            main.wait_stream(s)
