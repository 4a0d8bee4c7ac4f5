"""Native NVLink communicator: symmetric arena + fused P2P / NVLS all-reduce-update kernels.

Replaces ``init_process_group(backend)`` + ``DistributedDataParallel`` (reference src/trainer.py:59,98).  The torch
process group (gloo) is used only to agree on phases of the bootstrap; every byte of gradient and parameter
traffic moves in kernels of ``csrc/comm_p2p.cu`` over memory mapped by ``csrc/symm.cpp``.

Arena layout (identical offsets on every rank):
    [flags]   per-CTA tickets/flags        (FlagBlock)
    [scalars] 2 x world x 32 floats        (loss / metric reduction, double-buffered)
    [recv]    2 x world x slot bytes       (one-shot receive slots, double-buffered)
    [heap]    bump-allocated flat buffers  (gradients, compute-dtype weights, fp32 master)
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import check, launches, ptr, stream_handle
from ..optim.rules import RULE_IDS
from ..utils.logging import get_logger
from .comm import Comm

logger = get_logger(__name__)

ALGO_IDS = {"oneshot": 0, "twoshot": 1, "nvls": 2}

# Per-world crossover points of the three fused all-reduce + update variants: the largest gradient bucket, in bytes,
# for which one-shot is the fastest, and the largest for which two-shot still beats NVLS.  Measured with bench/allreduce_sweep.py
# (bf16 gradients, SGD, 20 back-to-back launches captured in one CUDA graph, max over ranks; records under profiles/).  One-shot
# moves (N-1) S bytes per GPU and updates the whole bucket on every rank, so its range shrinks as the world grows.
CROSSOVER = {
    # 2 GPUs: one-shot 7-16 us up to 1 MiB (two-shot / NVLS 10-26 us), NVLS ahead from 8 MiB (38 us vs 39 / -)
    2: {"oneshot": 2 << 20, "twoshot": 4 << 20, "source": "profiles/allreduce_sweep_2gpu_r2.json"},
    4: {"oneshot": 128 << 10, "twoshot": 128 << 10, "source": "between the 2- and 8-GPU sweeps (no 4-GPU box was measured)"},
    # 8 GPUs: one-shot only wins up to 4 KiB (10 us vs 12-13); NVLS stays at 12-24 us up to 1 MiB while one-shot grows to 33 us
    # and two-shot sits at 23-70 us, so two-shot is never the choice when the multicast object exists
    8: {"oneshot": 4 << 10, "twoshot": 4 << 10, "source": "profiles/allreduce_sweep_8gpu_r2.json"},
}


def crossover_for(world: int) -> dict:
    keys = sorted(CROSSOVER)
    return CROSSOVER[min((k for k in keys if k >= world), default=keys[-1])]
_c_int, _c_ll, _c_vp = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p


class P2PCtx(ctypes.Structure):
    _fields_ = [("base", _c_vp * 8), ("mc_base", _c_vp), ("world", _c_int), ("rank", _c_int),
                ("flags_off", _c_ll), ("recv_off", _c_ll), ("recv_slot_bytes", _c_ll), ("mc_flags", _c_int),
                ("reserved", _c_int)]


class _DevMem:
    """Zero-copy tensor view of raw device memory (``__cuda_array_interface__``)."""

    def __init__(self, addr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (addr, False), "version": 3}
        self._owner = owner


def _align(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class SymmArena:
    def __init__(self, rank: int, world: int, device: torch.device, nbytes: int, tag: str, group=None):
        L = _lib.lib()
        L.mlt_symm_create.restype = _c_vp
        L.mlt_symm_peer_ptr.restype = _c_vp
        L.mlt_symm_multicast_ptr.restype = _c_vp
        L.mlt_symm_bytes.restype = _c_ll
        self.L, self.rank, self.world, self.device, self.group = L, rank, world, device, group
        self.h = L.mlt_symm_create(_c_int(rank), _c_int(world), _c_int(device.index), _c_ll(nbytes), tag.encode())
        if not self.h:
            raise RuntimeError(f"symmetric arena creation failed: {L.mlt_last_error().decode()}")
        self.h = _c_vp(self.h)
        self._barrier()
        check(L.mlt_symm_exchange(self.h), "mlt_symm_exchange")
        self._barrier()
        self.multicast = False
        if world > 1 and os.environ.get("MLT_DISABLE_NVLS", "0") != "1":
            oks = []
            for phase in (0, 1, 2):
                rc = L.mlt_symm_multicast_setup(self.h, _c_int(phase))
                if rc > 0:
                    logger.warning(f"NVLS multicast setup failed in phase {phase}: {L.mlt_last_error().decode()}")
                oks.append(rc == 0)
                ok_all = self._all_true(rc == 0)
                if not ok_all:
                    break
            else:
                self.multicast = True
        self.nbytes = int(L.mlt_symm_bytes(self.h))
        self.bases: List[int] = [int(L.mlt_symm_peer_ptr(self.h, _c_int(p)) or 0) for p in range(world)]
        self.mc_base = int(L.mlt_symm_multicast_ptr(self.h) or 0) if self.multicast else 0
        self.base = self.bases[rank]
        self._mem = torch.as_tensor(_DevMem(self.base, self.nbytes, self), device=device)

    def _barrier(self):
        if self.world > 1:
            dist.barrier(group=self.group)

    def _all_true(self, flag: bool) -> bool:
        if self.world == 1:
            return flag
        t = torch.tensor([1 if flag else 0], dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return bool(t.item())

    def view(self, offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        return self._mem[offset:offset + nbytes].view(dtype)

    def offset_of(self, t: torch.Tensor) -> Optional[int]:
        off = t.data_ptr() - self.base
        return off if 0 <= off < self.nbytes else None

    def close(self):
        if self.h:
            self._mem = None
            self.L.mlt_symm_destroy(self.h)
            self.h = None


class P2PComm(Comm):
    name = "p2p"
    fused = True
    _instances = 0          # communicators built so far in this process (same count on every rank: part of the arena's name)

    def __init__(self, device: torch.device, oneshot_max_bytes: Optional[int] = None, algo: str = "auto"):
        self.device = device
        self.rank = dist.get_rank()
        self.world_size = dist.get_world_size()
        self.local_rank = int(os.environ.get("LOCAL_RANK", self.rank))
        if self.world_size > 8:
            raise RuntimeError("the native P2P backend covers one NVSwitch node (<= 8 GPUs)")
        self.algo = algo
        self.cross = dict(crossover_for(self.world_size))
        if os.environ.get("MLT_ONESHOT_MAX_BYTES"):
            self.cross["oneshot"] = int(os.environ["MLT_ONESHOT_MAX_BYTES"])
        # receive-slot size: the largest bucket the one-shot variant can take (auto mode never asks for more than the crossover)
        self.oneshot_max = int(oneshot_max_bytes or max(self.cross["oneshot"], 2 << 20))
        self.cpu_group = dist.new_group(backend="gloo")      # host-side phases of the bootstrap only
        P2PComm._instances += 1
        self._serial = P2PComm._instances
        self.arena: Optional[SymmArena] = None
        self._heap = 0
        self._vals = None
        self.L = _lib.lib()
        self.L.mlt_p2p_flag_block_bytes.restype = _c_ll

    # ------------------------------------------------------------------ arena
    def prepare(self, param_elems: int, compute_dtype: torch.dtype, n_states: int = 2):
        """Create the arena sized for ``param_elems`` parameters (collective; same arguments on every rank).

        The heap holds the fp32 master, the compute-dtype weights and gradients AND the fp32 optimizer state: the sharded
        collectives update master / state only on the owning rank, and a checkpoint pulls the owners' slices over NVLink
        (``consolidate``), which needs them in peer-addressable memory."""
        if self.arena is not None:
            return
        esize = 2 if compute_dtype == torch.bfloat16 else 4
        flags = _align(int(self.L.mlt_p2p_flag_block_bytes()), 4096)
        scalars = _align(2 * 8 * 32 * 4, 4096)
        recv = 2 * self.world_size * self.oneshot_max
        heap = param_elems * (4 + esize + (esize if compute_dtype == torch.bfloat16 else 0) + 4 * max(int(n_states), 0)) + (1 << 20)
        self.flags_off, self.scalar_off = 0, flags
        self.recv_off = flags + scalars
        self.heap_off = self._heap = _align(self.recv_off + recv, 4096)
        total = self.heap_off + _align(heap, 1 << 21)
        tag = (os.environ.get("MLT_JOB_TAG") or f"{os.environ.get('MASTER_PORT', '0')}-{os.getppid()}") + f"-{self._serial}"
        tags = [None] * self.world_size
        dist.all_gather_object(tags, tag if self.rank == 0 else None, group=self.cpu_group)
        tag = tags[0]
        self.arena = SymmArena(self.rank, self.world_size, self.device, total, tag, group=self.cpu_group)
        self.ctx = P2PCtx()
        for p in range(8):
            self.ctx.base[p] = self.arena.bases[p] if p < self.world_size else 0
        self.ctx.mc_base = self.arena.mc_base or None
        self.ctx.world, self.ctx.rank = self.world_size, self.rank
        self.ctx.flags_off, self.ctx.recv_off, self.ctx.recv_slot_bytes = self.flags_off, self.recv_off, self.oneshot_max
        self.ctx.mc_flags = int(self.arena.multicast and os.environ.get("MLT_MC_FLAGS", "1") != "0")
        timeout_s = float(os.environ.get("MLT_P2P_TIMEOUT_S", "20"))
        check(self.L.mlt_p2p_set_timeout(ctypes.byref(self.ctx), ctypes.c_double(timeout_s)), "mlt_p2p_set_timeout")
        self._oneshot_off = {}            # (lo, hi) of a bucket -> its byte offset inside the one-shot receive slots
        self._oneshot_used = 0
        logger.info(f"symmetric arena: {total >> 20} MiB per rank, multicast={'yes' if self.arena.multicast else 'no'}")

    def alloc(self, numel, dtype, device):
        if self.arena is None:
            raise RuntimeError("P2PComm.prepare() must run before buffers are allocated")
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        off = _align(self._heap, 256)
        if off + nbytes > self.arena.nbytes:
            raise RuntimeError("symmetric arena exhausted")
        self._heap = off + nbytes
        return self.arena.view(off, numel, dtype)        # arena memory is zero-filled at creation

    # ------------------------------------------------------------ collectives
    def pick_algo(self, nbytes: int) -> str:
        if self.algo != "auto":
            if self.algo == "nvls" and not self.arena.multicast:
                return "twoshot"               # no multicast object on this node: same ownership, P2P data path
            return self.algo
        if not self.arena.multicast:               # no NVLS: one-shot beats two-shot up to 2 MiB at 2 and at 8 GPUs (same sweeps)
            return "oneshot" if nbytes <= min(self.oneshot_max, 2 << 20) else "twoshot"
        if nbytes <= self.oneshot_max and nbytes <= self.cross["oneshot"]:
            return "oneshot"
        if nbytes <= self.cross["twoshot"]:
            return "twoshot"
        return "nvls"

    def allreduce_update(self, opt, lo: int, hi: int, stream=None):
        n = hi - lo
        gsz = opt.grad.element_size()
        algo = self.algo_for(lo, hi, gsz)
        recv_off = self._oneshot_off.get((lo, hi), 0) if algo == "oneshot" else 0
        grad_off = self.arena.offset_of(opt.grad) + lo * gsz
        model_off = self.arena.offset_of(opt.model) + lo * opt.model.element_size()
        one = lambda t: (_c_vp * 1)(t.data_ptr() if t is not None else 0)
        s2 = opt.s2[lo:hi] if opt.s2.numel() > 1 else None
        check(self.L.mlt_p2p_allreduce_update_at(
            ctypes.byref(self.ctx), _c_int(ALGO_IDS[algo]), _c_int(RULE_IDS[opt.kind]),
            _c_int(int(opt.grad.dtype == torch.bfloat16)), _c_ll(grad_off), _c_ll(model_off),
            _c_int(int(opt.model is opt.master)), one(opt.master[lo:hi]), one(opt.s1[lo:hi]), one(s2), one(opt.hyper),
            _c_ll(n), _c_int(0), _c_int(1), _c_ll(recv_off), stream_handle(stream)), "mlt_p2p_allreduce_update")
        launches.bump(f"p2p_allreduce_update_{algo}")

    def algo_for(self, lo: int, hi: int, gsz: int) -> str:
        """The variant bucket ``[lo, hi)`` runs with.  Deterministic in (lo, hi) and identical on every rank: a one-shot
        bucket claims its own byte range of the receive slots the first time it is seen (buckets of one step must not share
        slot bytes); a bucket that no longer fits falls back to the sharded variants."""
        nbytes = (hi - lo) * gsz
        algo = self.pick_algo(nbytes)
        if algo == "oneshot":
            off = self._oneshot_off.get((lo, hi))
            if off is None:
                off = _align(self._oneshot_used, 256)
                if off + nbytes <= self.oneshot_max:
                    self._oneshot_off[(lo, hi)] = off
                    self._oneshot_used = off + nbytes
                else:
                    self._oneshot_off[(lo, hi)] = -1
                    off = -1
            if off < 0:
                algo = "nvls" if self.arena.multicast else "twoshot"
        return algo

    def owned_slices(self, opt, lo: int, hi: int):
        gsz = opt.grad.element_size()
        if self.world_size == 1 or self.algo_for(lo, hi, gsz) == "oneshot":
            return []
        E = 16 // gsz                                   # elements per 16-byte vector of the gradient dtype
        nvec = (hi - lo) // E
        per_rank = -(-nvec // self.world_size)
        out = []
        for r in range(self.world_size):
            a = min(nvec, r * per_rank)
            b = min(nvec, a + per_rank)
            if b > a:
                out.append((r, lo + a * E, lo + b * E))
        return out

    def consolidate(self, opt, buckets):
        """Every rank ends up with the complete fp32 master weights and optimizer state.

        The two-shot / NVLS kernels update master and state only on the rank that owns a slice (peers receive the
        compute-dtype weights).  Before a checkpoint each owner broadcasts its slices with the P2P store kernel."""
        if self.world_size == 1:
            return
        tensors = [opt.master, opt.s1] + ([opt.s2] if opt.s2.numel() > 1 else [])
        for bk in buckets:
            for r, a, b in self.owned_slices(opt, bk.lo, bk.hi):
                for t in tensors:
                    self.broadcast(t[a:b], src=r)
        torch.cuda.current_stream().synchronize()

    def all_reduce_mean(self, flat, stream=None):
        """Unfused in-place mean all-reduce (the reference's ``_average_gradients``), through the same P2P / NVLS
        kernels with the "mean only" rule: the reduced values are written back over the gradients on every rank.
        A tensor outside the arena (or oddly aligned) is staged through the arena's receive slots in chunks."""
        if self.world_size == 1:
            return
        off = self.arena.offset_of(flat) if self.arena is not None else None
        n = flat.numel()
        if flat.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("P2PComm.all_reduce_mean handles fp32 and bf16 tensors")
        if off is None or off % 16 != 0 or n % 8 != 0 or not flat.is_contiguous():
            self._staged_all_reduce_mean(flat)
            return
        self._mean_kernel(off, n, flat.dtype, stream, oneshot_ok=True)

    def _mean_kernel(self, off: int, n: int, dtype: torch.dtype, stream=None, oneshot_ok: bool = True):
        esz = 2 if dtype == torch.bfloat16 else 4
        algo = self.pick_algo(n * esz)
        if algo == "oneshot" and (not oneshot_ok or n * esz > self.oneshot_max):
            algo = "twoshot"
        if not hasattr(self, "_mean_tmp") or self._mean_tmp[0].numel() < n:
            self._mean_tmp = (torch.empty(n, dtype=torch.float32, device=self.device), torch.empty(n, dtype=torch.float32, device=self.device))
            self._mean_hyper = torch.zeros(8, dtype=torch.float32, device=self.device)
            self._mean_hyper[7] = 1.0 / self.world_size
        tmp_p, tmp_s = self._mean_tmp
        one = lambda t: (_c_vp * 1)(t.data_ptr() if t is not None else 0)
        check(self.L.mlt_p2p_allreduce_update(
            ctypes.byref(self.ctx), _c_int(ALGO_IDS[algo]), _c_int(5), _c_int(int(dtype == torch.bfloat16)), _c_ll(off),
            _c_ll(off), _c_int(0), one(tmp_p), one(tmp_s), one(None), one(self._mean_hyper), _c_ll(n), _c_int(0), _c_int(1),
            stream_handle(stream)), "mlt_p2p_allreduce_update(mean only)")
        launches.bump(f"p2p_allreduce_{algo}")

    def _staged_all_reduce_mean(self, flat):
        """Tensors the kernels cannot address directly: copy chunks into the (otherwise idle) receive-slot region of the arena,
        reduce there with the sharded kernel (each rank reads and writes only its own slice of every peer's copy), copy back."""
        esz = flat.element_size()
        region = 2 * self.world_size * self.oneshot_max
        cap = region // esz // 1024 * 1024                       # elements per chunk (a multiple of world * 8)
        src = flat.reshape(-1) if flat.is_contiguous() else flat.contiguous().view(-1)
        n = src.numel()
        for lo in range(0, n, cap):
            m = min(cap, n - lo)
            mp = (m + 1023) // 1024 * 1024
            stage = self.arena.view(self.recv_off, mp, flat.dtype)
            stage[:m].copy_(src[lo:lo + m])
            if mp > m:
                stage[m:].zero_()
            self._mean_kernel(self.recv_off, mp, flat.dtype, None, oneshot_ok=False)     # one-shot would receive into this very region
            src[lo:lo + m].copy_(stage[:m])
            self.device_barrier()                                 # nobody may restage before every rank has copied out
        if src.data_ptr() != flat.data_ptr():
            flat.copy_(src.view(flat.shape))

    def broadcast(self, tensor, src=0):
        if self.world_size == 1:
            return
        off = self.arena.offset_of(tensor)
        nbytes = tensor.numel() * tensor.element_size()
        if off is not None and off % 16 == 0 and nbytes % 16 == 0 and tensor.is_contiguous():
            check(self.L.mlt_p2p_broadcast(ctypes.byref(self.ctx), _c_ll(off), _c_ll(nbytes), _c_int(src), _c_int(1),
                                           stream_handle()), "mlt_p2p_broadcast")
            launches.bump("p2p_broadcast")
            return
        # tensors outside the arena (module buffers): bounce through the receive-slot region, in chunks when they are large
        region = 2 * self.world_size * self.oneshot_max
        flat = tensor.contiguous().view(-1).view(torch.uint8)
        for lo in range(0, nbytes, region):
            m = min(region, nbytes - lo)
            mp = _align(m, 16)
            stage = self.arena.view(self.recv_off, mp, torch.uint8)
            if self.rank == src:                      # only the source writes the region the broadcast pushes from (and into)
                stage[:m].copy_(flat[lo:lo + m])
            check(self.L.mlt_p2p_broadcast(ctypes.byref(self.ctx), _c_ll(self.recv_off), _c_ll(mp), _c_int(src), _c_int(1),
                                           stream_handle()), "mlt_p2p_broadcast")
            launches.bump("p2p_broadcast")
            flat[lo:lo + m].copy_(stage[:m])
            self.device_barrier()                     # peers must not reuse the region before everyone copied out
        if flat.data_ptr() != tensor.data_ptr():
            tensor.copy_(flat.view(tensor.dtype).view(tensor.shape))

    def reduce_scalars(self, values):
        if self.world_size == 1:
            return values
        k = values.numel()
        if self._vals is None:
            self._vals = torch.zeros(32, dtype=torch.float32, device=self.device)
            self._vals_ptr = torch.tensor([self._vals.data_ptr()], dtype=torch.int64, device=self.device)
        self._vals[:k].copy_(values.float().view(-1))
        check(self.L.mlt_p2p_reduce_scalars(ctypes.byref(self.ctx), _c_ll(self.scalar_off), ptr(self._vals_ptr), _c_int(k),
                                            _c_int(1), stream_handle()), "mlt_p2p_reduce_scalars")
        launches.bump("p2p_reduce_scalars")
        return self._vals[:k].clone().view(values.shape)

    def reduce_scalars_host(self, values):
        """Sum a short list of host floats over all ranks AND read this rank's timeout diagnostic, in one round trip to the
        device (what the Trainer does at the end of every epoch so that every rank logs the same loss / metric; the reference
        leaves these per-rank, its only cross-rank arithmetic is the sum-then-divide of /root/reference/src/trainer.py:152-158).
        Raises like :meth:`check_errors`."""
        if self.world_size == 1 or self.arena is None:
            return [float(v) for v in values]
        k = len(values)
        if self._vals is None:
            self._vals = torch.zeros(32, dtype=torch.float32, device=self.device)
            self._vals_ptr = torch.tensor([self._vals.data_ptr()], dtype=torch.int64, device=self.device)
        if getattr(self, "_pin_vals", None) is None:
            self._pin_vals = torch.zeros(40, dtype=torch.float32).pin_memory()
        pin = self._pin_vals
        for i, v in enumerate(values):
            pin[i] = float(v)
        check(self.L.mlt_p2p_reduce_scalars_host(ctypes.byref(self.ctx), _c_ll(self.scalar_off), ptr(self._vals_ptr), ptr(self._vals),
                                                 ctypes.c_void_p(pin.data_ptr()), _c_int(k), stream_handle()),
              "mlt_p2p_reduce_scalars_host")
        launches.bump("p2p_reduce_scalars")
        return pin[:k].tolist()

    def device_barrier(self):
        """All ranks have reached this point (and finished the work queued before it on the current stream): the scalar
        all-reduce kernel is a device-side barrier - its flag exchange completes only when every rank's kernel runs.  For
        waits that are short by construction (staging buffers); it is subject to the kernels' wait limit."""
        if self.world_size > 1:
            if not hasattr(self, "_bar_val"):
                self._bar_val = torch.zeros(1, dtype=torch.float32, device=self.device)
            self.reduce_scalars(self._bar_val)
            torch.cuda.current_stream().synchronize()

    def barrier(self):
        """Host-level barrier (the bootstrap's gloo group): what the Trainer uses around file writes, where one rank may
        legitimately keep the others waiting for longer than any kernel should spin."""
        if self.world_size > 1:
            torch.cuda.current_stream().synchronize()
            dist.barrier(group=self.cpu_group)

    def check_errors(self):
        """Raises ``RuntimeError`` with the rank / peer / CTA / ticket diagnostic if a flag wait of any collective on this rank
        ran past the limit (``MLT_P2P_TIMEOUT_S``, default 20 s), or a peer reported that it did."""
        if self.arena is None:
            return
        out = (ctypes.c_uint * 6)()
        check(self.L.mlt_p2p_check_error(ctypes.byref(self.ctx), out), "mlt_p2p_check_error")

    def close(self):
        if self.arena is not None:
            torch.cuda.synchronize(self.device)
            self.barrier()
            self.arena.close()
            self.arena = None
