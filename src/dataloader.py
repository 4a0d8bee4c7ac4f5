"""``src.dataloader`` of the reference (src/dataloader.py) -> :mod:`ml_trainer_b200.data.loader`."""
from ml_trainer_b200.data.loader import Loader, PinnedPrefetcher, SyntheticDataset  # noqa: F401
