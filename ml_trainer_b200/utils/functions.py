"""User-pluggable pre-processing and loss hooks (reference src/utils/functions.py:5-17).

``custom_pre_process_function`` returns the CIFAR augmentation pipeline and
``custom_loss_function`` is the mean-squared-error example loss.  torchvision is
imported lazily so the loss hook is usable without it.  A tensor-only variant of
the augmentation is provided for data that already lives in (pinned) tensors, which
is what the synthetic/benchmark pipeline feeds.
"""
from __future__ import annotations

import torch

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)


def custom_pre_process_function():
    """PIL -> normalised tensor pipeline with crop/flip augmentation."""
    from torchvision import transforms

    return transforms.Compose(
        [
            transforms.RandomCrop(32, padding=4),
            transforms.RandomHorizontalFlip(),
            transforms.ToTensor(),
            transforms.Normalize(CIFAR_MEAN, CIFAR_STD),
        ]
    )


class TensorAugment:
    """Same augmentation for CHW float tensors in [0, 1] (no PIL round trip)."""

    def __init__(self, size: int = 32, padding: int = 4, mean=CIFAR_MEAN, std=CIFAR_STD):
        self.size, self.padding = size, padding
        self.mean = torch.tensor(mean).view(-1, 1, 1)
        self.std = torch.tensor(std).view(-1, 1, 1)

    def __call__(self, img: torch.Tensor) -> torch.Tensor:
        p = self.padding
        img = torch.nn.functional.pad(img, (p, p, p, p))
        i, j = (int(v) for v in torch.randint(0, 2 * p + 1, (2,)))
        img = img[:, i : i + self.size, j : j + self.size]
        if torch.rand(()) < 0.5:
            img = img.flip(-1)
        return (img - self.mean) / self.std


def custom_loss_function(output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return torch.mean((output - target) ** 2)
