from .augment import DeviceAugment, U8ImageDataset
from .loader import Loader, PinnedPrefetcher, SyntheticDataset

__all__ = ["DeviceAugment", "Loader", "PinnedPrefetcher", "SyntheticDataset", "U8ImageDataset"]
