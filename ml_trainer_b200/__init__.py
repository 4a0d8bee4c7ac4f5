"""ml_trainer_b200: a B200-native training framework with the capabilities of abbomarengo/ml-trainer.

Public surface (mirrors the reference's ``src`` package, reference src/trainer.py, src/model.py,
src/dataloader.py, src/utils/*):

    from ml_trainer_b200 import Trainer, MLModel, Loader
    Trainer(model, datasets=(train, val), epochs=E, batch_size=B, is_parallel=False, **config).fit()
"""
from .config import TrainerConfig, validate_kwargs
from .data.augment import DeviceAugment, U8ImageDataset
from .data.loader import Loader, PinnedPrefetcher, SyntheticDataset
from .models import MLModel, available_models, build_model
from .trainer import Trainer
from .utils.functions import custom_loss_function, custom_pre_process_function
from .utils.utils import load_history, load_model, plot_history

__version__ = "0.1.0"

__all__ = [
    "Trainer", "TrainerConfig", "validate_kwargs", "MLModel", "build_model", "available_models",
    "Loader", "PinnedPrefetcher", "SyntheticDataset", "DeviceAugment", "U8ImageDataset", "load_model", "load_history", "plot_history",
    "custom_loss_function", "custom_pre_process_function",
]
