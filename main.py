"""Training entry point (reference main.py:1-84).

Same flags as the reference.  Differences, all deliberate:
``--epochs`` / ``--batch_size`` are honoured (the reference parses then ignores them,
main.py:44-45); the ``SM_*`` environment variables are optional; ``--custom_function``
parses real booleans; ``--synthetic`` trains without a dataset on disk; ``--model`` picks
from the model zoo; ``--local`` forces single-process mode.  Under ``torchrun`` /
``python -m ml_trainer_b200.parallel.launch`` one process per GPU runs this file with the
env:// contract (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT).
"""
import argparse
import json
import os

from src.model import build_model
from src.trainer import Trainer


def _str2bool(v):
    if isinstance(v, bool):
        return v
    return str(v).strip().lower() in ("1", "true", "t", "yes", "y")


def _none_or_str(v):
    return None if v in (None, "None", "none", "") else v


def build_datasets(args):
    if args.synthetic:
        from src.dataloader import SyntheticDataset

        if args.model == "bert_base":
            mk = lambda n, s: SyntheticDataset((args.seq_len,), num_classes=2, length=n, seed=s, kind="tokens", pool=512)
        elif args.model == "resnet50":
            mk = lambda n, s: SyntheticDataset((3, 224, 224), num_classes=1000, length=n, seed=s, pool=256)
        else:
            mk = lambda n, s: SyntheticDataset((3, 32, 32), num_classes=10, length=n, seed=s)
        return mk(args.synthetic_train, args.seed), mk(args.synthetic_val, args.seed + 1)
    import torchvision

    if args.custom_function and args.device_augment:
        # The reference's recipe (RandomCrop(32, 4) + RandomHorizontalFlip + ToTensor + Normalize, src/utils/functions.py:5-12)
        # without PIL in the loop: the images stay uint8 in one tensor, batches go to the GPU as uint8 and one kernel
        # crops / flips / normalises them (ml_trainer_b200/data/augment.py).  On a CPU-only box the same class applies
        # the same arithmetic on the host.
        from ml_trainer_b200 import DeviceAugment, U8ImageDataset

        raw_train = torchvision.datasets.CIFAR10(root=args.data_dir, train=True, download=False)
        raw_val = torchvision.datasets.CIFAR10(root=args.data_dir, train=False, download=False)
        aug = DeviceAugment((3, 32, 32), padding=4, flip=True, seed=args.seed)
        return (U8ImageDataset.from_torchvision(raw_train, aug, train=True),
                U8ImageDataset.from_torchvision(raw_val, DeviceAugment((3, 32, 32), padding=0, flip=False), train=False))
    if args.custom_function:
        from src.utils.functions import custom_pre_process_function

        transform = custom_pre_process_function()
    else:
        from torchvision import transforms

        transform = transforms.ToTensor()       # the reference passes no transform here and then fails to collate PIL images
    train_set = torchvision.datasets.CIFAR10(root=args.data_dir, train=True, download=False, transform=transform)
    val_set = torchvision.datasets.CIFAR10(root=args.data_dir, train=False, download=False, transform=transform)
    return train_set, val_set


def main(args):
    datasets = build_datasets(args)
    model = build_model(args.model)
    config = {
        "seed": args.seed,
        "scheduler": args.scheduler,
        "optimizer": args.optimizer,
        "momentum": args.momentum,
        "weight_decay": args.weight_decay,
        "lr": args.lr,
        "criterion": args.criterion,
        "pred_function": args.pred_function,
        "metric": args.metric,
        "model_dir": args.model_dir,
        "backend": args.backend,
        "dtype": args.dtype,
        "graph": args.graph,
        "progress": not args.no_progress,
    }
    is_parallel = (not args.local) and int(os.environ.get("WORLD_SIZE", "1")) >= 1 and \
        ("RANK" in os.environ or args.is_parallel)
    trainer = Trainer(model, datasets=datasets, epochs=args.epochs, batch_size=args.batch_size,
                      is_parallel=is_parallel, save_history=True, **config)
    trainer.fit()
    return trainer


def get_parser():
    parser = argparse.ArgumentParser()
    # Pytorch environments (reference main.py:52-77)
    parser.add_argument("--batch_size", type=int, default=32, help="input (global) batch size for training (default: 32)")
    parser.add_argument("--epochs", type=int, default=10, help="number of epochs to train (default: 10)")
    parser.add_argument("--optimizer", type=str, default="sgd", help="optimizer for backward pass (default: sgd)")
    parser.add_argument("--lr", type=float, default=0.001, help="learning rate (default: 0.001)")
    parser.add_argument("--momentum", type=float, default=0.9, help="Optimizer momentum (default: 0.9)")
    parser.add_argument("--weight_decay", type=float, default=0.0, help="Optimizer weight decay (default: 0.0)")
    parser.add_argument("--seed", type=int, default=32, help="random seed (default: 32)")
    parser.add_argument("--scheduler", type=_none_or_str, default=None, help="apply scheduler for learning rate (default: None)")
    parser.add_argument("--criterion", type=str, default="cross_entropy", help="loss function to apply (default: cross_entropy)")
    parser.add_argument("--metric", type=_none_or_str, default=None, help="metric for model evaluation (default: None)")
    parser.add_argument("--backend", type=str, default="smddp",
                        help="distributed backend: smddp/p2p (native NVLink kernels on GPUs), gloo, nccl")
    parser.add_argument("--custom_function", type=_str2bool, default=False, help="apply a pre-processing function (default: False)")
    parser.add_argument("--pred_function", type=_none_or_str, default=None,
                        help="probability function to apply to make predictions (default: None)")
    # SageMaker environment (reference main.py:80-83) - optional here
    parser.add_argument("--hosts", type=json.loads, default=json.loads(os.environ.get("SM_HOSTS", '["localhost"]')))
    parser.add_argument("--current-host", type=str, default=os.environ.get("SM_CURRENT_HOST", "localhost"))
    parser.add_argument("--model_dir", type=str, default=os.environ.get("SM_MODEL_DIR", "model_output"))
    parser.add_argument("--data_dir", type=str, default=os.environ.get("SM_CHANNEL_TRAIN", "cifar10-dataset"))
    # B200 additions
    parser.add_argument("--model", type=str, default="mlmodel", choices=["mlmodel", "resnet50", "bert_base"])
    parser.add_argument("--dtype", type=str, default="fp32", choices=["fp32", "bf16"])
    parser.add_argument("--graph", type=str, default="auto", help="CUDA-graph the train step: auto|true|false")
    parser.add_argument("--is_parallel", action="store_true", help="force parallel mode (implied under torchrun)")
    parser.add_argument("--local", action="store_true", help="force single-process mode")
    parser.add_argument("--nproc", type=int, default=0, help="spawn this many local ranks (0 = do not spawn)")
    parser.add_argument("--synthetic", action="store_true", help="train on synthetic data of the model's input shape")
    parser.add_argument("--synthetic_train", type=int, default=2048)
    parser.add_argument("--synthetic_val", type=int, default=512)
    parser.add_argument("--seq_len", type=int, default=512)
    parser.add_argument("--no_progress", action="store_true")
    parser.add_argument("--device_augment", type=_str2bool, default=True,
                        help="with --custom_function: keep CIFAR uint8 and crop / flip / normalise in a GPU kernel (default: True); "
                             "False uses the torchvision PIL pipeline exactly as the reference does")
    return parser


if __name__ == "__main__":
    args = get_parser().parse_args()
    if args.nproc and "RANK" not in os.environ:
        from ml_trainer_b200.parallel.launch import spawn_self

        raise SystemExit(spawn_self(args.nproc))
    main(args)
