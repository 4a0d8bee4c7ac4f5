"""Multi-GPU launch (the role of the reference's 02_ML_Training_SageMaker_distributed.ipynb).

The reference hands main.py to a SageMaker PyTorch estimator with smdistributed data-parallel enabled; here the same
main.py is started as one process per GPU of the node by the in-tree launcher (torchrun works too), with the same
hyper-parameters-as-flags contract:

    python examples/02_train_distributed.py --nproc 8
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from ml_trainer_b200.parallel.launch import launch  # noqa: E402

hyperparameters = {"seed": 32, "optimizer": "sgd", "momentum": 0.9, "lr": 0.001, "criterion": "cross_entropy",
                   "pred_function": "softmax", "metric": "accuracy", "custom_function": True, "backend": "smddp",
                   "epochs": 2, "batch_size": 256}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--nproc", type=int, default=0)
    ap.add_argument("--data_dir", default=None, help="CIFAR-10 root; omitted -> synthetic data")
    a = ap.parse_args()
    import torch

    n = a.nproc or max(torch.cuda.device_count(), 1)
    flags = [f"--{k}={v}" for k, v in hyperparameters.items()]
    flags += ["--data_dir", a.data_dir] if a.data_dir else ["--synthetic"]
    raise SystemExit(launch([sys.executable, os.path.join(ROOT, "main.py"), *flags, "--no_progress"], n))
