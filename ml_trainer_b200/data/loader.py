"""Data pipeline: ``Loader``, synthetic datasets and the pinned H2D prefetcher.

Reference: ``class Loader(DataLoader): pass`` (src/dataloader.py:5-6), driven with
``num_workers=0``, pageable memory and blocking ``.to(device)`` copies
(src/trainer.py:77-79,182-183).  ``Loader`` here is still a drop-in ``DataLoader``
subclass with the same constructor, plus:

* a vectorised fast path for in-memory tensor datasets (``TensorDataset`` /
  :class:`SyntheticDataset`): a batch is one ``index_select`` into a pinned slot instead
  of ``batch_size`` Python ``__getitem__`` calls and a collate;
* :class:`PinnedPrefetcher`: pinned staging slots, ``cudaMemcpyAsync`` on a dedicated
  copy stream and an event hand-off to the compute stream, so the upload of batch
  i+1 overlaps the compute of batch i.
"""
from __future__ import annotations

from typing import Iterator, Optional, Sequence, Tuple

import torch
from torch.utils.data import DataLoader, Dataset, TensorDataset


class SyntheticDataset(Dataset):
    """Random inputs of a given shape with integer class labels.

    A pool of ``pool`` distinct samples is generated once (seeded) and indexed modulo the
    pool size, so arbitrarily long epochs cost no extra memory.  ``kind='image'`` gives
    N(0,1) float inputs; ``kind='tokens'`` gives int64 token ids in ``[0, vocab)``.
    """

    def __init__(self, shape: Sequence[int], num_classes: int = 10, length: int = 1024,
                 dtype: torch.dtype = torch.float32, seed: int = 0, pool: Optional[int] = None,
                 kind: str = "image", vocab: int = 30522, target_shape: Optional[Sequence[int]] = None):
        g = torch.Generator().manual_seed(seed)
        self.length = int(length)
        n = min(self.length, pool or self.length)
        if kind == "tokens":
            self.inputs = torch.randint(0, vocab, (n, *shape), generator=g, dtype=torch.int64)
        else:
            self.inputs = torch.randn((n, *shape), generator=g, dtype=torch.float32).to(dtype)
        if target_shape is not None:           # regression targets (l1 / l2 / mcrmse)
            self.targets = torch.randn((n, *target_shape), generator=g)
        else:
            self.targets = torch.randint(0, num_classes, (n,), generator=g, dtype=torch.int64)
        self.tensors = (self.inputs, self.targets)
        self._pool = n

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        j = i % self._pool
        return self.inputs[j], self.targets[j]


def _tensor_backing(dataset) -> Optional[Tuple[torch.Tensor, ...]]:
    if isinstance(dataset, SyntheticDataset):
        return dataset.tensors
    from .augment import U8ImageDataset

    if isinstance(dataset, U8ImageDataset):
        return dataset.tensors
    if isinstance(dataset, TensorDataset) and all(isinstance(t, torch.Tensor) and t.device.type == "cpu"
                                                  for t in dataset.tensors):
        return dataset.tensors
    return None


class Loader(DataLoader):
    """``DataLoader`` with a vectorised path for in-memory tensor datasets."""

    def __init__(self, dataset, batch_size=1, shuffle=None, sampler=None, **kw):
        super().__init__(dataset, batch_size=batch_size, shuffle=shuffle, sampler=sampler, **kw)
        backing = _tensor_backing(dataset)
        self._fast = (backing is not None and self.num_workers == 0
                      and kw.get("collate_fn") is None and self.batch_sampler is not None)
        self._backing = backing
        self._pool = getattr(dataset, "_pool", None)
        # uint8 image datasets: batches travel as uint8 and are augmented on the device (data/augment.py)
        self._augment = getattr(dataset, "augment", None) if backing is not None and backing[0].dtype == torch.uint8 else None
        self._augment_train = bool(getattr(dataset, "train", True))

    def __iter__(self):
        if not self._fast:
            return super().__iter__()
        self.draw_base_seed()
        return self._fast_iter()

    def draw_base_seed(self) -> int:
        """Every ``iter(DataLoader)`` draws one 64-bit base seed from the loader's generator (the global RNG when there is
        none) before its sampler draws anything (torch/utils/data/dataloader.py, ``_BaseDataLoaderIter.__init__``).  The
        vectorised path and the native epoch executors make the same draw, so a run consumes the random stream exactly as the
        reference's stock ``DataLoader`` does: same seed, same shuffles, same loss curves
        (tests/test_reference_parity_cpu.py)."""
        return int(torch.empty((), dtype=torch.int64).random_(generator=self.generator).item())

    def _fast_iter(self) -> Iterator[Tuple[torch.Tensor, ...]]:
        pin = bool(self.pin_memory) and torch.cuda.is_available()
        for idx in self.batch_sampler:
            index = torch.as_tensor(idx, dtype=torch.int64)
            if self._pool is not None:
                index = index % self._pool
            out = []
            for t in self._backing:
                buf = torch.empty((index.numel(), *t.shape[1:]), dtype=t.dtype, pin_memory=pin)
                torch.index_select(t, 0, index, out=buf)
                out.append(buf)
            aug = self._augment
            if aug is not None and not getattr(self, "_augment_on_device", False):
                # host statement of the augmentation (CPU training, or a consumer that does not augment on the device)
                pos0, train = aug.position, self._augment_train
                out[0] = torch.stack([aug.apply_cpu(out[0][k], pos0 + k, train) for k in range(index.numel())])
                if train:
                    aug.position += index.numel()
            yield tuple(out)


class PinnedPrefetcher:
    """Double-buffered H2D pipeline over any iterable of ``(inputs, targets)`` CPU batches.

    ``slots`` device buffers and matching pinned host buffers are allocated lazily from the
    first batch's shapes.  ``__next__`` returns *device* tensors whose contents are
    guaranteed ready on the current (compute) stream: the copy runs on ``copy_stream``
    and the compute stream waits on the slot's event.  A slot is reused only after the
    compute stream has passed the point where it was consumed (``release`` event).
    """

    def __init__(self, loader, device: torch.device, slots: int = 2, cast_inputs: Optional[torch.dtype] = None):
        self.loader = loader
        self.device = device
        self.slots = max(int(slots), 1)
        self.cast_inputs = cast_inputs
        self.enabled = device.type == "cuda"
        self.copy_stream = torch.cuda.Stream(device=device) if self.enabled else None
        self._host = [None] * self.slots
        self._dev = [None] * self.slots
        self._ready = [None] * self.slots
        self._released = [None] * self.slots
        self.h2d_bytes = 0

    def __len__(self):
        return len(self.loader)

    def _stage(self, slot: int, batch):
        tensors = [torch.as_tensor(t) for t in batch]
        if self._dev[slot] is None or any(d.shape != t.shape or h.dtype != t.dtype
                                          for d, h, t in zip(self._dev[slot], self._host[slot], tensors)):
            self._host[slot] = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in tensors]
            self._dev[slot] = [torch.empty(t.shape, dtype=t.dtype, device=self.device) for t in tensors]
            self._ready[slot] = torch.cuda.Event()
            self._released[slot] = None
        if self._released[slot] is not None:
            self._released[slot].synchronize()          # host must not overwrite pinned data still in flight
        for h, t in zip(self._host[slot], tensors):
            if t.is_pinned() and t.shape == h.shape:
                continue
            h.copy_(t)
        from ..utils.tracing import phase

        with phase("h2d"), torch.cuda.stream(self.copy_stream):
            if self._released[slot] is not None:
                self.copy_stream.wait_event(self._released[slot])
            for d, h, t in zip(self._dev[slot], self._host[slot], tensors):
                src = t if (t.is_pinned() and t.shape == h.shape) else h
                d.copy_(src, non_blocking=True)
                self.h2d_bytes += src.numel() * src.element_size()
            self._ready[slot].record(self.copy_stream)

    def __iter__(self):
        if not self.enabled:
            for batch in self.loader:
                yield tuple(torch.as_tensor(t) for t in batch)
            return
        it = iter(self.loader)
        queue = []
        slot = 0
        try:
            for _ in range(self.slots):
                self._stage(slot, next(it))
                queue.append(slot)
                slot = (slot + 1) % self.slots
        except StopIteration:
            pass
        while queue:
            cur = queue.pop(0)
            torch.cuda.current_stream().wait_event(self._ready[cur])
            out = tuple(self._dev[cur])
            if self.cast_inputs is not None and out[0].is_floating_point():
                out = (out[0].to(self.cast_inputs),) + out[1:]
            yield out
            rel = torch.cuda.Event()
            rel.record(torch.cuda.current_stream())
            self._released[cur] = rel
            try:
                self._stage(cur, next(it))
                queue.append(cur)
            except StopIteration:
                pass
