"""uint8 image datasets augmented on the device (the reference's CIFAR recipe without PIL in the loop).

Reference data flow (main.py:12-28 + src/utils/functions.py:5-12): ``torchvision.datasets.CIFAR10`` yields PIL images, each
goes through ``RandomCrop(32, padding=4) -> RandomHorizontalFlip -> ToTensor -> Normalize`` on the host, ``default_collate``
stacks 32 of them, and the fp32 batch is copied to the GPU from pageable memory.

Here the images stay uint8 in one tensor (``U8ImageDataset``, built straight from ``CIFAR10.data`` / ``.targets``): a batch is a
row gather of 3 KB samples, goes over PCIe as uint8 (4x fewer bytes) and ONE kernel (csrc/augment.cu) pads, crops, flips,
scales and normalises it into the step's input buffer.  The random draw of every sample is a counter-based function of
``(seed, position in the training stream)``; :meth:`DeviceAugment.params` states it on the host, which is how the CPU data
path (``U8ImageDataset.__getitem__``, any ``DataLoader``) produces bit-identical images and how the tests check the kernel.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch
from torch.utils.data import Dataset

from ..utils.functions import CIFAR_MEAN, CIFAR_STD


class AugmentCfg(ctypes.Structure):
    """Mirror of ``mlt_augment_cfg`` (include/mlt_b200.h)."""

    _fields_ = [("C", ctypes.c_int), ("H", ctypes.c_int), ("W", ctypes.c_int), ("out_h", ctypes.c_int), ("out_w", ctypes.c_int),
                ("pad", ctypes.c_int), ("flip", ctypes.c_int), ("hwc", ctypes.c_int), ("mean", ctypes.c_float * 4),
                ("std", ctypes.c_float * 4), ("seed", ctypes.c_ulonglong)]


def _philox4x32_10(c, k0, k1):
    M0, M1, W0, W1, MASK = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85, 0xFFFFFFFF
    c = list(c)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [((p1 >> 32) ^ c[1] ^ k0) & MASK, p1 & MASK, ((p0 >> 32) ^ c[3] ^ k1) & MASK, p0 & MASK]
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c


class DeviceAugment:
    """``RandomCrop(size, padding) + RandomHorizontalFlip + ToTensor + Normalize(mean, std)`` for uint8 images."""

    def __init__(self, shape: Sequence[int] = (3, 32, 32), padding: int = 4, flip: bool = True, mean=CIFAR_MEAN, std=CIFAR_STD,
                 seed: int = 0, layout: str = "hwc"):
        C, H, W = (int(v) for v in shape)
        if not 1 <= C <= 4:
            raise ValueError("DeviceAugment handles 1 to 4 channels")
        if layout not in ("hwc", "chw"):
            raise ValueError("layout must be 'hwc' or 'chw'")
        self.shape, self.padding, self.flip, self.layout, self.seed = (C, H, W), int(padding), bool(flip), layout, int(seed)
        self.mean, self.std = tuple(float(m) for m in mean)[:C], tuple(float(s) for s in std)[:C]
        self.position = 0                    # samples drawn so far in training (advanced by whoever feeds batches)
        cfg = AugmentCfg()
        cfg.C, cfg.H, cfg.W, cfg.out_h, cfg.out_w = C, H, W, H, W
        cfg.pad, cfg.flip, cfg.hwc, cfg.seed = self.padding, int(self.flip), int(layout == "hwc"), self.seed & (2 ** 64 - 1)
        for i in range(4):
            cfg.mean[i] = self.mean[i] if i < C else 0.0
            cfg.std[i] = self.std[i] if i < C else 1.0
        self.cfg = cfg

    @property
    def out_shape(self):
        return self.shape

    # ---- the random draw -------------------------------------------------------------------------------------------
    def params(self, pos: int):
        """``(top, left, flip)`` of the sample at stream position ``pos`` - the same draw the kernel makes."""
        c = _philox4x32_10([pos & 0xFFFFFFFF, (pos >> 32) & 0xFFFFFFFF, 0x61756730, 0], self.seed & 0xFFFFFFFF, (self.seed >> 32) & 0xFFFFFFFF)
        span = 2 * self.padding + 1
        return (c[0] * span) >> 32, (c[1] * span) >> 32, (c[2] >> 31) if self.flip else 0

    # ---- host statement (CPU data path and test oracle) ---------------------------------------------------------------
    def apply_cpu(self, img_u8: torch.Tensor, pos: Optional[int], train: bool = True) -> torch.Tensor:
        """One uint8 image ([H, W, C] or [C, H, W] per ``layout``) -> normalised fp32 [C, H, W]."""
        C, H, W = self.shape
        x = img_u8.reshape(H, W, C).permute(2, 0, 1) if self.layout == "hwc" else img_u8.reshape(C, H, W)
        x = x.float() * (1.0 / 255.0)
        p = self.padding
        if train and pos is not None:
            top, left, flip = self.params(pos)
            x = torch.nn.functional.pad(x, (p, p, p, p))[:, top:top + H, left:left + W]
            if flip:
                x = x.flip(-1)
        mean = torch.tensor(self.mean).view(C, 1, 1)
        inv = 1.0 / torch.tensor(self.std).view(C, 1, 1)
        return (x - mean) * inv

    # ---- device path -------------------------------------------------------------------------------------------------
    def apply_cuda(self, batch_u8: torch.Tensor, out: Optional[torch.Tensor] = None, pos0: Optional[int] = None, train: bool = True,
                   dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """uint8 batch on the device -> ``[B, C, H, W]`` ``dtype``.  ``pos0`` defaults to (and then advances) ``self.position``."""
        from ..ops import _lib
        from ..ops._lib import check, launches, ptr, stream_handle

        B = batch_u8.shape[0]
        C, H, W = self.shape
        if out is None:
            out = torch.empty((B, C, H, W), dtype=dtype, device=batch_u8.device)
        if pos0 is None:
            pos0 = self.position
            if train:
                self.position += B
        check(_lib.lib().mlt_augment_u8(ptr(batch_u8.contiguous()), ptr(out), ctypes.c_int(int(out.dtype == torch.bfloat16)), ctypes.c_int(B),
                                        ctypes.byref(self.cfg), ctypes.c_ulonglong(pos0), ctypes.c_int(int(train)), stream_handle()),
              "mlt_augment_u8")
        launches.bump("augment_u8")
        return out


class U8ImageDataset(Dataset):
    """uint8 images + integer labels held in two tensors, augmented by a :class:`DeviceAugment`.

    ``data``: ``[N, H, W, C]`` (what ``torchvision.datasets.CIFAR10.data`` is) or ``[N, C, H, W]`` uint8 array / tensor;
    ``targets``: N labels.  ``__getitem__`` is the host statement of the pipeline (any ``DataLoader`` works, CPU training
    works); the Trainer's native epoch loop reads ``tensors`` directly and augments on the device."""

    def __init__(self, data, targets, augment: Optional[DeviceAugment] = None, train: bool = True, layout: str = "hwc"):
        data = torch.as_tensor(data)
        if data.dtype != torch.uint8 or data.dim() != 4:
            raise ValueError("U8ImageDataset wants a uint8 [N, H, W, C] or [N, C, H, W] array")
        if augment is None:
            shp = (data.shape[3], data.shape[1], data.shape[2]) if layout == "hwc" else tuple(data.shape[1:])
            augment = DeviceAugment(shp, padding=0, flip=False, layout=layout)
        self.augment, self.train = augment, bool(train)
        self.inputs = data.reshape(data.shape[0], -1).contiguous()
        self.targets = torch.as_tensor(targets, dtype=torch.int64).contiguous()
        self.tensors = (self.inputs, self.targets)
        self._served = 0

    @classmethod
    def from_torchvision(cls, ds, augment: Optional[DeviceAugment] = None, train: bool = True):
        """Wrap a ``torchvision.datasets.CIFAR10``-style dataset (``.data`` uint8 NHWC array, ``.targets`` list)."""
        return cls(ds.data, ds.targets, augment=augment, train=train, layout="hwc")

    def __len__(self):
        return self.inputs.shape[0]

    def __getitem__(self, i):
        pos = self._served
        self._served += 1
        return self.augment.apply_cpu(self.inputs[i], pos if self.train else None, self.train), self.targets[i]


CIFAR10_CLASSES = ("airplane", "automobile", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck")


def synthetic_cifar(n: int, seed: int = 0):
    """A CIFAR-10-shaped stand-in for boxes without the dataset: ``(data uint8 [n, 32, 32, 3], labels int64 [n])``.

    Every class has its own colour and a bright square whose position depends on the class, under per-image noise, so a
    small CNN can actually learn it (the notebooks and examples show a falling loss, not a flat one)."""
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, 10, (n,), generator=g)
    base = torch.randint(0, 256, (10, 3), generator=torch.Generator().manual_seed(1234)).float()
    img = base[labels].view(n, 1, 1, 3).expand(n, 32, 32, 3).clone()
    img += torch.randn(n, 32, 32, 3, generator=g) * 40.0
    for c in range(10):
        idx = (labels == c).nonzero().flatten()
        y0, x0 = 2 + 5 * (c // 5) * 3, 2 + 6 * (c % 5)
        img[idx, y0:y0 + 8, x0:x0 + 6, :] = 255.0 - base[c]
    return img.clamp_(0, 255).to(torch.uint8), labels
