from .loader import Loader, PinnedPrefetcher, SyntheticDataset

__all__ = ["Loader", "PinnedPrefetcher", "SyntheticDataset"]
