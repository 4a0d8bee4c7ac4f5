"""Communicator API used by the reducer and the trainer.

The reference hands a backend *string* to ``dist.init_process_group`` and lets
``DistributedDataParallel`` issue every collective (reference src/trainer.py:40,59,98;
SURVEY §2.3 N10-N13).  Here the collectives the trainer needs are an explicit, small
interface with three implementations:

``LocalComm``   world size 1, everything is a no-op.
``TorchComm``   ``torch.distributed`` collectives (gloo on CPU, nccl for the library
                baseline in sweeps).  Semantics: SUM all-reduce then divide by world
                size, i.e. the reference's ``_average_gradients`` (src/trainer.py:152-158).
``P2PComm``     the B200 product path (``parallel/p2p.py``): symmetric memory mapped
                across ranks, hand-written NVLink load/store + NVLS multimem kernels,
                fused with the optimizer update.

Backend names accepted from the Trainer: ``smddp`` (the reference default; mapped to
the native path on GPUs, gloo on CPU), ``p2p``/``native``, ``gloo``, ``nccl``.
"""
from __future__ import annotations

import datetime
import os
from typing import Optional

import torch
import torch.distributed as dist

from ..utils.logging import get_logger

logger = get_logger(__name__)

NATIVE_NAMES = ("smddp", "p2p", "native", "auto")


class Comm:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    name: str = "local"
    fused: bool = False          # True when allreduce_update also applies the optimizer rule

    def prepare(self, param_elems: int, compute_dtype: torch.dtype, n_states: int = 2) -> None:
        """Size any backend-owned memory for a model of ``param_elems`` (padded) parameters and ``n_states`` fp32
        optimizer-state vectors.  Collective."""

    def reduce_scalars_host(self, values):
        """Host floats summed over all ranks, plus the error check; back ends with a cheaper fused form override this."""
        dev = getattr(self, "device", None)
        if dev is None or getattr(self, "name", "") == "gloo":
            dev = "cpu"
        out = self.reduce_scalars(torch.tensor([float(v) for v in values], dtype=torch.float32, device=dev)).tolist()
        self.check_errors()
        return out

    def check_errors(self) -> None:
        """Raise if a collective recorded a failure (a peer that stopped answering).  Host-side, cheap."""

    def owned_slices(self, optimizer, lo: int, hi: int):
        """Element ranges ``[(rank, lo, hi), ...]`` of bucket ``[lo, hi)`` whose fp32 master / optimizer state is only
        up to date on ``rank`` (sharded-update collectives); empty when every rank updates the whole bucket."""
        return []

    def consolidate(self, optimizer, buckets) -> None:
        """Make every rank's fp32 master weights and optimizer state complete (before a checkpoint)."""

    # -- collectives ---------------------------------------------------------
    def all_reduce_mean(self, flat: torch.Tensor, stream=None) -> None:
        """In-place mean across ranks of a flat tensor."""

    def allreduce_update(self, optimizer, lo: int, hi: int, stream=None) -> None:
        """Average ``optimizer.grad[lo:hi]`` across ranks and apply the update rule to that slice."""
        self.all_reduce_mean(optimizer.grad[lo:hi], stream)
        optimizer.step_range(lo, hi)

    def broadcast(self, tensor: torch.Tensor, src: int = 0) -> None:
        pass

    def reduce_scalars(self, values: torch.Tensor) -> torch.Tensor:
        """Element-wise SUM of a small fp32 vector across ranks (returned on the same device)."""
        return values

    def barrier(self) -> None:
        pass

    def alloc(self, numel: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
        """Allocate a flat buffer peers can address (plain memory for non-P2P backends)."""
        return torch.zeros(numel, dtype=dtype, device=device)

    def close(self) -> None:
        pass


class LocalComm(Comm):
    pass


class TorchComm(Comm):
    """``torch.distributed`` backed communicator (gloo / nccl)."""

    def __init__(self, backend: str, device: torch.device):
        self.name = backend
        self.rank = dist.get_rank()
        self.world_size = dist.get_world_size()
        self.local_rank = int(os.environ.get("LOCAL_RANK", self.rank))
        self.device = device

    def all_reduce_mean(self, flat, stream=None):
        if self.world_size == 1:
            return
        if flat.dtype == torch.bfloat16 and self.name == "gloo":
            tmp = flat.float()
            dist.all_reduce(tmp, op=dist.ReduceOp.SUM)
            flat.copy_(tmp.div_(self.world_size))
            return
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(self.world_size)

    def broadcast(self, tensor, src=0):
        if self.world_size > 1:
            dist.broadcast(tensor, src=src)

    def reduce_scalars(self, values):
        if self.world_size > 1:
            values = values.clone()
            dist.all_reduce(values, op=dist.ReduceOp.SUM)
        return values

    def barrier(self):
        if self.world_size > 1:
            dist.barrier()


def _env_world() -> int:
    return int(os.environ.get("WORLD_SIZE", "1"))


def init_process_group(backend: str, device: torch.device, timeout_s: int = 600) -> str:
    """env:// rendezvous (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT), same contract as the reference.

    Returns the torch backend actually initialised.  The torch process group is only
    used for bootstrap (handle exchange) and for the gloo/nccl fallback collectives.
    """
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    if backend in NATIVE_NAMES:
        torch_backend = "nccl" if device.type == "cuda" else "gloo"
        # the native path needs only a CPU-side store/gloo group for bootstrap
        if device.type == "cuda":
            torch_backend = "cpu:gloo,cuda:nccl"
    else:
        torch_backend = backend
    if not dist.is_initialized():
        dist.init_process_group(backend=torch_backend, timeout=datetime.timedelta(seconds=timeout_s))
    return torch_backend


def make_comm(backend: str, device: torch.device, is_parallel: bool, algo: str = "auto") -> Comm:
    """Build the communicator for ``backend``; initialises the process group when parallel."""
    if not is_parallel:
        return LocalComm()
    init_process_group(backend, device)
    if dist.get_world_size() == 1 and backend not in ("p2p", "native"):
        c = TorchComm("gloo" if device.type == "cpu" else "nccl", device)
        return c
    if backend in NATIVE_NAMES and device.type == "cuda":
        from .p2p import P2PComm

        return P2PComm(device, algo=algo)
    if backend in NATIVE_NAMES:
        logger.info("native P2P backend needs CUDA; using gloo collectives on CPU")
        return TorchComm("gloo", device)
    return TorchComm(backend, device)
