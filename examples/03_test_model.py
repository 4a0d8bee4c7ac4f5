"""Inference-only flow (the reference's 03_ML_Testing.ipynb): Loader -> load_model -> Trainer(model) -> test()."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from src.dataloader import Loader, SyntheticDataset
from src.model import MLModel
from src.trainer import Trainer
from src.utils.utils import load_model

if __name__ == "__main__":
    val_set = SyntheticDataset((3, 32, 32), 10, length=1024, seed=1)
    test_loader = Loader(val_set, batch_size=32, shuffle=True)
    model = load_model(MLModel(), sys.argv[1] if len(sys.argv) > 1 else "model_output/model.pth")
    trainer = Trainer(model)                 # no datasets: testing only
    print(trainer.test(model, test_loader))
