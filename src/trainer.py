"""``src.trainer`` of the reference (src/trainer.py) -> :mod:`ml_trainer_b200.trainer`."""
from ml_trainer_b200.trainer import Trainer, logger  # noqa: F401
