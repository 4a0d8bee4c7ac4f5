"""Checkpoint writing without disturbing the live model.

The reference saves with ``torch.save(self.model.cpu().state_dict(), ...)`` which moves
the live model off the GPU every epoch and re-uploads it afterwards (reference
src/trainer.py:232-235,176).  Here the fp32 master weights are snapshotted with an
asynchronous D2H copy into pinned staging buffers on a side stream; the file content
stays identical: an ``OrderedDict[str, fp32 CPU tensor]`` keyed by the module's own
names, ``module.``-prefixed when written from parallel mode.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, Optional

import torch


class SnapshotWriter:
    def __init__(self, device: torch.device):
        self.device = device
        self.stream = torch.cuda.Stream(device=device) if device.type == "cuda" else None
        self._staging: Dict[str, torch.Tensor] = {}
        self.d2h_bytes = 0

    def snapshot(self, tensors: "OrderedDict[str, torch.Tensor]") -> "OrderedDict[str, torch.Tensor]":
        """Return CPU copies of ``tensors`` (fp32 for floating types), synchronised."""
        out = OrderedDict()
        if self.stream is None:
            for k, t in tensors.items():
                t = t.detach()
                out[k] = (t.float() if t.is_floating_point() else t).cpu().clone()
            return out
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            for k, t in tensors.items():
                t = t.detach()
                if t.is_floating_point() and t.dtype != torch.float32:
                    t = t.float()
                buf = self._staging.get(k)
                if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
                    buf = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
                    self._staging[k] = buf
                buf.copy_(t, non_blocking=True)
                self.d2h_bytes += buf.numel() * buf.element_size()
        self.stream.synchronize()
        for k in tensors:
            out[k] = self._staging[k].clone()      # un-pinned, owned by the checkpoint
        return out


def atomic_torch_save(obj, path: str) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(obj, tmp)
    os.replace(tmp, path)


def save_state_dict(state: "OrderedDict[str, torch.Tensor]", model_dir: str, name: str = "model.pth",
                    writer: Optional[SnapshotWriter] = None) -> str:
    os.makedirs(model_dir, exist_ok=True)
    path = os.path.join(model_dir, name)
    cpu_state = writer.snapshot(state) if writer is not None else OrderedDict(
        (k, v.detach().float().cpu() if v.is_floating_point() else v.detach().cpu()) for k, v in state.items())
    atomic_torch_save(cpu_state, path)
    return path
