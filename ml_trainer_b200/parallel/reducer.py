"""Gradient bucket plan + ready tracking (replaces the C++ DDP Reducer).

The optimizer arena is already in gradient-ready order, so a bucket is a contiguous
element range ``[lo, hi)`` covering consecutive parameters, capped at ``bucket_bytes``
of gradient (DDP's default 25 MiB; its small first bucket is kept so communication
starts early).  When the last gradient of a bucket lands, the bucket's fused
all-reduce + update is launched on a side stream behind an event, overlapping the rest
of backward.  Buckets launch strictly in index order on every rank: the P2P kernels of
different ranks wait for each other, so a rank-dependent order could deadlock.

Gradients reach the arena two ways:
  * native ops write straight into ``param._mlt_grad_sink`` and call ``param._mlt_ready()``;
  * any other module goes through ``register_post_accumulate_grad_hook``: the hook copies
    ``param.grad`` into its slot and clears it (DDP's non-bucket-view behaviour).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import torch

from .comm import Comm, LocalComm


@dataclass
class Bucket:
    index: int
    first_param: int
    last_param: int      # exclusive
    lo: int              # arena element range
    hi: int
    pending: int = 0
    launched: bool = False


def plan_buckets(sizes_bytes: List[int], cap_bytes: int, first_cap_bytes: Optional[int] = None) -> List[tuple]:
    """Greedy contiguous bucketing.  Returns ``[(first_param, last_param_exclusive), ...]``.

    Mirrors DDP's ``_compute_bucket_assignment_by_size`` shape: a small first bucket,
    then buckets that close as soon as they reach the cap (a parameter larger than the
    cap gets a bucket of its own).
    """
    try:
        from ..ops._lib import lib_or_none

        lib = lib_or_none()
        if lib is not None and hasattr(lib, "mlt_plan_buckets") and sizes_bytes:
            import ctypes

            n = len(sizes_bytes)
            arr = (ctypes.c_int64 * n)(*sizes_bytes)
            out = (ctypes.c_int32 * (n + 1))()
            nb = lib.mlt_plan_buckets(arr, n, int(cap_bytes), int(first_cap_bytes or cap_bytes), out)
            return [(out[i], out[i + 1]) for i in range(nb)]
    except Exception:
        pass
    plans, start, acc = [], 0, 0
    cap = first_cap_bytes or cap_bytes
    for i, sz in enumerate(sizes_bytes):
        acc += sz
        if acc >= cap:
            plans.append((start, i + 1))
            start, acc, cap = i + 1, 0, cap_bytes
    if start < len(sizes_bytes):
        plans.append((start, len(sizes_bytes)))
    return plans


class Reducer:
    def __init__(self, optimizer, comm: Optional[Comm] = None, bucket_mb: float = 25.0,
                 first_bucket_mb: float = 1.0, overlap: bool = True):
        self.opt = optimizer
        self.comm = comm or LocalComm()
        self.world = self.comm.world_size
        # fused backends fold the 1/world of the mean into the update kernel; others divide in all_reduce_mean
        self.grad_scale = 1.0 / self.world if (self.world > 1 and self.comm.fused) else 1.0
        esize = optimizer.grad.element_size()
        sizes = [n * esize for n in optimizer.sizes]
        cap = int(bucket_mb * 2 ** 20)
        first = int(min(first_bucket_mb, bucket_mb) * 2 ** 20) if self.world > 1 else cap
        self.buckets: List[Bucket] = []
        for bi, (a, b) in enumerate(plan_buckets(sizes, cap, first)):
            lo, hi = optimizer.span(a, b)
            self.buckets.append(Bucket(bi, a, b, lo, hi))
        self.bucket_of = [0] * len(sizes)
        for bk in self.buckets:
            for i in range(bk.first_param, bk.last_param):
                self.bucket_of[i] = bk.index
        self.cuda = optimizer.device.type == "cuda"
        self.overlap = overlap and self.cuda
        self.side = torch.cuda.Stream(device=optimizer.device) if self.overlap else None
        self._next = 0
        self._armed = False
        self._hooks = []
        self._install()

    # ------------------------------------------------------------------ hooks
    def _install(self):
        for i, p in enumerate(self.opt.params_in_order):
            p._mlt_ready = (lambda i=i: self.mark_ready(i))
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))

    def _make_hook(self, i):
        def hook(p):
            if p.grad is not None:
                self.opt.grad_view(i).copy_(p.grad)
                p.grad = None
            self.mark_ready(i)
        return hook

    def remove(self):
        for h in self._hooks:
            h.remove()
        self._hooks.clear()
        for p in self.opt.params_in_order:
            if hasattr(p, "_mlt_ready"):
                del p._mlt_ready

    # ------------------------------------------------------------- step cycle
    def begin_step(self):
        """Arm the buckets and publish this step's hyper-parameters (eager convenience)."""
        self.opt.begin_step(grad_scale=self.grad_scale)
        self.arm()

    def arm(self):
        """Call before backward.  Pure host bookkeeping, so it is safe inside graph capture."""
        for bk in self.buckets:
            bk.pending = bk.last_param - bk.first_param
            bk.launched = False
        self._next = 0
        self._armed = True
        self._seen = [False] * len(self.bucket_of)

    def mark_ready(self, i: int):
        if not self._armed or self._seen[i]:
            return
        self._seen[i] = True
        bk = self.buckets[self.bucket_of[i]]
        bk.pending -= 1
        while self._next < len(self.buckets) and self.buckets[self._next].pending <= 0:
            self._launch(self.buckets[self._next])
            self._next += 1

    def _launch(self, bk: Bucket):
        bk.launched = True
        if self.overlap:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                self._reduce_update(bk)
        else:
            self._reduce_update(bk)

    def _reduce_update(self, bk: Bucket):
        if self.world > 1:
            self.comm.allreduce_update(self.opt, bk.lo, bk.hi)
        else:
            self.opt.step_range(bk.lo, bk.hi)

    def finish_step(self):
        """Call after backward: flushes buckets whose params got no gradient and joins the side stream."""
        if not self._armed:
            return
        while self._next < len(self.buckets):       # parameters that took no part in this step
            bk = self.buckets[self._next]
            for i in range(bk.first_param, bk.last_param):
                if not self._seen[i]:               # slots are overwritten, not accumulated: clear stale data
                    self.opt.grad_view(i).zero_()
            self._launch(bk)
            self._next += 1
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self.side)
        self._armed = False
