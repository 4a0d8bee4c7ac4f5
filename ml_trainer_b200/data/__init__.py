from .augment import CIFAR10_CLASSES, DeviceAugment, U8ImageDataset, synthetic_cifar
from .loader import Loader, PinnedPrefetcher, SyntheticDataset

__all__ = ["CIFAR10_CLASSES", "synthetic_cifar", "DeviceAugment", "Loader", "PinnedPrefetcher", "SyntheticDataset", "U8ImageDataset"]
