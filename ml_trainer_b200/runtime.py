"""Python face of the native epoch executor (csrc/runtime.cpp).

``NativeLoop`` owns the C++ ring of pinned staging + device buffers, the copy stream and the gather thread.  The
Trainer hands it a whole epoch at once - the sampler's index order, the per-step hyper-parameter rows (so any LR
schedule is honoured) and the step to run - and gets the per-step losses back.  Per step the host then costs a few
microseconds of C++ instead of a pass through the interpreter (reference src/trainer.py:180-197).
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence

import torch

from .ops import _lib
from .ops._lib import check, launches, ptr, stream_handle
from .parallel.p2p import P2PCtx

_c_int, _c_ll, _c_vp = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p


class LeNetCall(ctypes.Structure):
    _fields_ = [("master", _c_vp), ("s1", _c_vp), ("s2", _c_vp), ("hyper", _c_vp), ("numel", _c_ll),
                ("param_off", ctypes.POINTER(_c_ll)), ("rule", _c_int), ("train", _c_int), ("want_acc", _c_int),
                ("acc3", _c_vp), ("loss_out", _c_vp), ("scratch", _c_vp), ("scratch_bytes", _c_ll),
                ("ctx", ctypes.POINTER(P2PCtx))]


class NativeLoop:
    def __init__(self, device: torch.device, max_batch: int, x_row_bytes: int, y_row_bytes: int, slots: int = 3):
        self.L = _lib.lib()
        self.L.mlt_loop_create.restype = _c_vp
        self.L.mlt_loop_slot_x.restype = _c_vp
        self.L.mlt_loop_slot_y.restype = _c_vp
        self.device, self.max_batch = device, max_batch
        self.x_row, self.y_row = x_row_bytes, y_row_bytes
        h = self.L.mlt_loop_create(_c_int(device.index), _c_int(slots), _c_ll(max_batch), _c_ll(x_row_bytes), _c_ll(y_row_bytes))
        if not h:
            raise RuntimeError(f"mlt_loop_create failed: {self.L.mlt_last_error().decode()}")
        self.h = _c_vp(h)
        self.slots = int(self.L.mlt_loop_num_slots(self.h))
        self.last_stats = {}

    def close(self):
        if self.h:
            self.L.mlt_loop_destroy(self.h)
            self.h = None

    def __del__(self):  # best effort
        try:
            self.close()
        except Exception:
            pass

    def slot_buffers(self, slot: int):
        return int(self.L.mlt_loop_slot_x(self.h, _c_int(slot))), int(self.L.mlt_loop_slot_y(self.h, _c_int(slot)))

    def _finish(self, steps):
        a, b, c = _c_ll(0), _c_ll(0), _c_ll(0)
        self.L.mlt_loop_stats(self.h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
        t = (ctypes.c_double * 3)()
        self.L.mlt_loop_timing(self.h, t)
        self.L.mlt_loop_gather_wait_us.restype = ctypes.c_double
        gw = self.L.mlt_loop_gather_wait_us(self.h)
        self.last_stats = {"h2d_bytes": a.value, "d2h_bytes": b.value, "launches": c.value, "steps": steps,
                           "host_us": {"first_enqueued": t[0], "all_enqueued": t[1], "drained": t[2], "gather_wait": gw}}

    def _lenet_call(self, fused) -> "LeNetCall":
        """The fixed arguments of the step kernel, built once per (optimizer arena, scratch buffer)."""
        o = fused.opt
        key = (id(fused), fused._scratch.data_ptr(), o.master.data_ptr())
        if getattr(self, "_call_key", None) != key:
            call = LeNetCall()
            call.master, call.s1 = o.master.data_ptr(), o.s1.data_ptr()
            call.s2 = o.s2.data_ptr() if o.s2.numel() > 1 else None
            call.hyper, call.numel = o.hyper.data_ptr(), o.numel
            call.param_off = ctypes.cast(fused.off, ctypes.POINTER(_c_ll))
            call.rule = fused.rule
            call.scratch, call.scratch_bytes = fused._scratch.data_ptr(), fused._scratch.numel()
            call.ctx = ctypes.pointer(fused.ctx) if fused.ctx is not None else None
            self._call, self._call_key = call, key
        return self._call

    def run_lenet(self, fused, host_x: torch.Tensor, host_y: torch.Tensor, indices: torch.Tensor, batch: int,
                  batch_sizes: Optional[Sequence[int]], hyper_rows: Optional[torch.Tensor], train: bool,
                  acc3: Optional[torch.Tensor], want_acc: bool, loss_buf: torch.Tensor) -> torch.Tensor:
        """Run one epoch of single-kernel LeNet steps.  Returns the per-step losses (CPU float32 tensor).

        ``batch_sizes``: None when every step is a full batch, else the per-step sizes."""
        steps = len(batch_sizes) if batch_sizes is not None else indices.numel() // batch
        fused.ensure_scratch(max(batch_sizes) if batch_sizes is not None else batch, self.device)
        call = self._lenet_call(fused)
        call.train, call.want_acc = int(train), int(want_acc)
        call.acc3 = acc3.data_ptr() if acc3 is not None else None
        call.loss_out = loss_buf.data_ptr()
        losses = torch.empty(steps, dtype=torch.float32)
        idx = indices if indices.is_contiguous() else indices.contiguous()
        bs = (ctypes.c_int * steps)(*batch_sizes) if batch_sizes is not None else None
        hr = hyper_rows.contiguous() if hyper_rows is not None else None
        # kind 0: one launch per step.  kind 2: the whole run is ONE persistent launch of the step kernel fed through the
        # ring (the copy stream publishes "batch i has landed" in a device word the kernel polls).  One GPU: kind 0 is a little
        # faster (37.6 vs 40.5 us/step on B200: back-to-back launches already hide the launch gap).  Several GPUs: the ranks'
        # kernels wait for each other inside every step, so with one launch per step every step pays the slowest rank's launch
        # jitter; the persistent form removes the per-step launches from that critical path.  MLT_LENET_PERSISTENT=0/1 overrides.
        env = os.environ.get("MLT_LENET_PERSISTENT", "auto")
        want_persistent = False if env == "auto" else env == "1"      # "auto" stays on kind 0 until the N >= 2 measurement says otherwise
        persistent = self.slots <= 8 and want_persistent and steps >= 2
        check(self.L.mlt_loop_run(self.h, ptr(host_x), ptr(host_y), ptr(idx), _c_ll(steps), _c_int(batch), bs,
                                  ptr(hr), ptr(fused.opt.hyper), _c_int(2 if persistent else 0), ctypes.byref(call), None, ptr(loss_buf),
                                  ptr(losses), stream_handle()), "mlt_loop_run")
        self.last_kind = 2 if persistent else 0
        launches.bump("lenet_step", 1 if persistent else steps)
        self._finish(steps)
        return losses

    # ------------------------------------------------------------------ generic (CUDA-graph) steps
    def slot_tensors(self, slot: int, x_shape, x_dtype, y_shape, y_dtype):
        """Zero-copy torch views of a ring slot's device buffers (what the captured graphs read)."""
        from .parallel.p2p import _DevMem

        xp, yp = self.slot_buffers(slot)
        xb = int(torch.Size(x_shape).numel()) * torch.empty((), dtype=x_dtype).element_size()
        yb = int(torch.Size(y_shape).numel()) * torch.empty((), dtype=y_dtype).element_size()
        x = torch.as_tensor(_DevMem(xp, xb, self), device=self.device).view(x_dtype).view(x_shape)
        y = torch.as_tensor(_DevMem(yp, yb, self), device=self.device).view(y_dtype).view(y_shape)
        return x, y

    def set_augment(self, aug, out: Optional[torch.Tensor] = None, pos0: int = 0, train: bool = True):
        """uint8 image rows: every step first runs the augmentation kernel from the slot's rows into ``out``."""
        if aug is None:
            if getattr(self, "_aug_set", False):
                self.L.mlt_loop_set_augment(self.h, None, None, _c_int(0), ctypes.c_ulonglong(0), _c_int(0))
                self._aug_set = False
            return
        self._aug_keep = (aug, out)
        self.L.mlt_loop_set_augment(self.h, ctypes.byref(aug.cfg), _c_vp(out.data_ptr()), _c_int(int(out.dtype == torch.bfloat16)),
                                    ctypes.c_ulonglong(int(pos0)), _c_int(int(train)))
        self._aug_set = True

    def set_direct_rows(self, on: bool):
        self.L.mlt_loop_set_direct_rows(self.h, _c_int(int(on)))

    def run_graphs(self, graph_execs: Sequence[int], hyper_dev: torch.Tensor, host_x, host_y, indices, batch: int,
                   hyper_rows: Optional[torch.Tensor], loss_dev: torch.Tensor, n_launches_per_step: int = 0) -> torch.Tensor:
        """Run full-batch steps whose work is one captured CUDA graph per ring slot."""
        steps = indices.numel() // batch
        execs = (_c_vp * len(graph_execs))(*graph_execs)
        losses = torch.empty(steps, dtype=torch.float32)
        hr = hyper_rows.contiguous() if hyper_rows is not None else None
        idx = indices.contiguous()
        check(self.L.mlt_loop_run(self.h, ptr(host_x), ptr(host_y), ptr(idx), _c_ll(steps), _c_int(batch), None, ptr(hr),
                                  ptr(hyper_dev), _c_int(1), None, execs, ptr(loss_dev), ptr(losses), stream_handle()),
              "mlt_loop_run")
        if n_launches_per_step:
            launches.bump("graph_replay_kernels", n_launches_per_step * steps)
        self._finish(steps)
        return losses


def _sampler_order(sampler) -> torch.Tensor:
    """The index order a stock sampler would yield, as one tensor.  The common samplers are reproduced with the SAME
    random draws their ``__iter__`` makes (so a native epoch sees exactly the order a Python epoch would); anything else
    is simply iterated."""
    from torch.utils.data import RandomSampler, SequentialSampler
    from torch.utils.data.distributed import DistributedSampler

    if type(sampler) is SequentialSampler:
        return torch.arange(len(sampler.data_source), dtype=torch.int64)
    if type(sampler) is RandomSampler and not sampler.replacement and sampler.num_samples == len(sampler.data_source):
        g = sampler.generator
        if g is None:
            g = torch.Generator()
            g.manual_seed(int(torch.empty((), dtype=torch.int64).random_().item()))
        return torch.randperm(len(sampler.data_source), generator=g)
    if type(sampler) is DistributedSampler:
        n = len(sampler.dataset)
        if sampler.shuffle:
            g = torch.Generator()
            g.manual_seed(sampler.seed + sampler.epoch)
            idx = torch.randperm(n, generator=g)
        else:
            idx = torch.arange(n, dtype=torch.int64)
        if not sampler.drop_last:
            pad = sampler.total_size - n
            if pad > 0:
                reps = -(-pad // n)
                idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        else:
            idx = idx[:sampler.total_size]
        return idx[sampler.rank:sampler.total_size:sampler.num_replicas].contiguous()
    return torch.tensor(list(sampler), dtype=torch.int64)


def epoch_indices(loader) -> tuple:
    """Flatten a Loader's batch order into (int64 index tensor, per-step batch sizes)."""
    bs = loader.batch_sampler
    if type(bs) is torch.utils.data.BatchSampler:
        # stock batching: the order IS the sampler's order, cut every batch_size (one pass, no list per batch)
        flat = _sampler_order(bs.sampler)
        n, B = flat.numel(), bs.batch_size
        full, rem = divmod(n, B)
        sizes = [B] * full + ([rem] if rem and not bs.drop_last else [])
        flat = flat[:sum(sizes)]
    else:
        batches: List[List[int]] = [list(b) for b in bs]
        sizes = [len(b) for b in batches]
        flat = torch.tensor([i for b in batches for i in b], dtype=torch.int64)
    pool = getattr(loader, "_pool", None)
    if pool is not None:
        flat = flat % pool
    return flat, sizes
