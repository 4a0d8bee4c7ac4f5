"""Local-mode smoke run (reference test.py:1-43, which is broken twice over: its transform
helper returns ``None`` and it passes the config dict positionally into ``epochs``).

This is the working version of that script: CIFAR-10 from ``cifar10-dataset/`` when it is on
disk, otherwise synthetic data of the same shape, trained in single-process mode.
"""
import os

from src.dataloader import SyntheticDataset
from src.model import MLModel
from src.trainer import Trainer

training_dir = "cifar10-dataset"


def _get_transforms():
    from src.utils.functions import custom_pre_process_function

    return custom_pre_process_function()


def _datasets():
    if os.path.isdir(training_dir):
        import torchvision

        train_set = torchvision.datasets.CIFAR10(root=training_dir, train=True, download=False, transform=_get_transforms())
        val_set = torchvision.datasets.CIFAR10(root=training_dir, train=False, download=False, transform=_get_transforms())
        return train_set, val_set
    return (SyntheticDataset((3, 32, 32), 10, length=512, seed=32),
            SyntheticDataset((3, 32, 32), 10, length=128, seed=33))


config = {
    "seed": 32,
    "scheduler": None,
    "optimizer": "sgd",
    "momentum": 0.9,
    "lr": 0.001,
    "criterion": "cross_entropy",
    "metric": "accuracy",
    "model_dir": os.environ.get("MLT_MODEL_DIR", "model_output"),
}

if __name__ == "__main__":
    model = MLModel()
    trainer = Trainer(model, _datasets(), epochs=int(os.environ.get("MLT_EPOCHS", "2")), batch_size=32,
                      is_parallel=False, **config)
    trainer.fit()
