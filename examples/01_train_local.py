"""Local training walkthrough (the flow of the reference's 01_ML_Training_local.ipynb).

transforms -> dataset -> Trainer(model, datasets, epochs, batch_size, **config) -> fit() -> save_history_ ->
load_history / plot_history -> load_model + trainer.test().  Uses CIFAR-10 from ./cifar10-dataset when present,
otherwise synthetic data of the same shape.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from src.dataloader import Loader, SyntheticDataset
from src.model import MLModel
from src.trainer import Trainer
from src.utils.utils import load_history, load_model, plot_history


def datasets(root="cifar10-dataset"):
    if os.path.isdir(root):
        import torchvision
        from src.utils.functions import custom_pre_process_function

        t = custom_pre_process_function()
        return (torchvision.datasets.CIFAR10(root, train=True, download=False, transform=t),
                torchvision.datasets.CIFAR10(root, train=False, download=False, transform=t))
    return SyntheticDataset((3, 32, 32), 10, length=4096, seed=0), SyntheticDataset((3, 32, 32), 10, length=1024, seed=1)


if __name__ == "__main__":
    config = {"seed": 32, "scheduler": None, "optimizer": "sgd", "momentum": 0.9, "lr": 0.001,
              "criterion": "cross_entropy", "metric": "accuracy", "model_dir": "model_output"}
    train_set, val_set = datasets()
    trainer = Trainer(MLModel(), datasets=(train_set, val_set), epochs=int(os.environ.get("EPOCHS", 6)), batch_size=32, **config)
    trainer.fit()
    trainer.save_history_("model_output")
    history = load_history("model_output")
    print({k: (v[-1] if isinstance(v, list) and v else v) for k, v in history.items()})
    try:
        plot_history(history, show=False, save_path="model_output/history.png")
    except Exception as e:  # matplotlib is optional
        print("plot skipped:", e)
    model = load_model(MLModel(), "model_output/model.pth")
    print("test:", trainer.test(model, Loader(val_set, batch_size=32, shuffle=True)))
