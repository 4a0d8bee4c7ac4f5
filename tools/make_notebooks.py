#!/usr/bin/env python
"""Write (and run) the three walkthrough notebooks that mirror the reference's (01 local training, 02 distributed launch,
03 testing).  There is no Jupyter on the build box, so this script is the kernel: every code cell is executed in one shared
namespace from the ``notebooks/`` directory, its stdout / last expression / PIL images are stored as the cell's outputs, and
the result is written as nbformat-4 JSON.  ``python tools/make_notebooks.py [01 02 03]``."""
import base64
import contextlib
import io
import json
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NB_DIR = os.path.join(ROOT, "notebooks")


def md(text):
    return {"cell_type": "markdown", "metadata": {}, "source": text.strip("\n").splitlines(keepends=True)}


def code(text):
    return {"cell_type": "code", "metadata": {}, "execution_count": None, "outputs": [], "source": text.strip("\n").splitlines(keepends=True)}


def run_cells(cells):
    import ast

    ns = {"__name__": "__main__"}
    n = 0
    cwd = os.getcwd()
    os.chdir(NB_DIR)
    try:
        for cell in cells:
            if cell["cell_type"] != "code":
                continue
            n += 1
            cell["execution_count"] = n
            src = "".join(cell["source"])
            tree = ast.parse(src)
            last = None
            if tree.body and isinstance(tree.body[-1], ast.Expr):
                last = ast.Expression(tree.body.pop().value)
            buf = io.StringIO()
            value = None
            ns["_images"] = []
            try:
                with contextlib.redirect_stdout(buf), contextlib.redirect_stderr(buf):
                    exec(compile(tree, f"<cell {n}>", "exec"), ns)
                    if last is not None:
                        value = eval(compile(last, f"<cell {n}>", "eval"), ns)
            except Exception:
                buf.write(traceback.format_exc())
                print(f"cell {n} raised:\n{traceback.format_exc()}", file=sys.stderr)
            text = buf.getvalue()
            if text:
                cell["outputs"].append({"output_type": "stream", "name": "stdout", "text": text.splitlines(keepends=True)})
            for img in ns.get("_images", []):
                b = io.BytesIO()
                img.save(b, format="PNG")
                cell["outputs"].append({"output_type": "display_data", "metadata": {},
                                        "data": {"image/png": base64.b64encode(b.getvalue()).decode(), "text/plain": ["<PIL image>"]}})
            if value is not None:
                cell["outputs"].append({"output_type": "execute_result", "execution_count": n, "metadata": {},
                                        "data": {"text/plain": repr(value).splitlines(keepends=True)}})
    finally:
        os.chdir(cwd)
    return cells


def write(name, cells):
    nb = {"cells": cells, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                                       "language_info": {"name": "python"}}, "nbformat": 4, "nbformat_minor": 5}
    with open(os.path.join(NB_DIR, name), "w") as f:
        json.dump(nb, f, indent=1)
    print("wrote", name)


SHOW = '''
def show(img):
    """Display helper: a PIL image becomes the cell's picture (matplotlib when it is installed)."""
    try:
        from matplotlib import pyplot as plt
        plt.figure(figsize=(16, 8)); plt.axis('off'); plt.imshow(img); plt.show()
    except ImportError:
        globals().setdefault('_images', []).append(img)
'''


def nb01():
    return [
        md("""
# 01 - Local training (ml_trainer_b200)

The reference's `01_ML_Training_local.ipynb`, cell for cell: transforms -> CIFAR-10 -> `Trainer` -> class histogram ->
`make_grid` preview -> `fit()` -> `save_history_` -> `load_history` / `plot_history` -> `load_model` + `trainer.test`.

What differs underneath: the images stay **uint8** (`U8ImageDataset`), the reference's
`RandomCrop(32, 4) + RandomHorizontalFlip + ToTensor + Normalize` runs as one GPU kernel per batch (the same arithmetic on the
host when there is no GPU), and on a B200 the whole training step of `MLModel` is a single kernel driven by a C++ epoch loop.
If `cifar10-dataset/` is not on disk (this box has no network) a CIFAR-shaped synthetic set stands in.
"""),
        code("""
import os, sys
sys.path.insert(0, '..')
import torch
import torchvision
from torchvision.utils import make_grid
from src.model import MLModel
from src.trainer import Trainer
from src.dataloader import Loader
from src.utils.utils import plot_history, load_model, load_history
from ml_trainer_b200 import DeviceAugment, U8ImageDataset
from ml_trainer_b200.data import CIFAR10_CLASSES, synthetic_cifar
""" + SHOW),
        code("training_dir = 'cifar10-dataset'"),
        code("""
# Define data augmentation: the reference's transforms.Compose([RandomCrop(32, padding=4), RandomHorizontalFlip(),
# ToTensor(), Normalize(mean, std)]) as ONE object that runs on the device (or, bit for bit the same, on the host)
def _get_transforms(train=True):
    return DeviceAugment((3, 32, 32), padding=4 if train else 0, flip=train,
                         mean=(0.4914, 0.4822, 0.4465), std=(0.2023, 0.1994, 0.2010), seed=32)
"""),
        code("""
if os.path.isdir(training_dir):
    raw_train = torchvision.datasets.CIFAR10(root=training_dir, train=True, download=False)
    raw_val = torchvision.datasets.CIFAR10(root=training_dir, train=False, download=False)
    classes = raw_train.classes
    train_set = U8ImageDataset.from_torchvision(raw_train, _get_transforms(True), train=True)
    val_set = U8ImageDataset.from_torchvision(raw_val, _get_transforms(False), train=False)
else:
    print('no CIFAR-10 on disk: using a CIFAR-shaped synthetic stand-in')
    classes = list(CIFAR10_CLASSES)
    train_set = U8ImageDataset(*synthetic_cifar(4096, seed=0), augment=_get_transforms(True), train=True)
    val_set = U8ImageDataset(*synthetic_cifar(1024, seed=1), augment=_get_transforms(False), train=False)
"""),
        code("val_set.inputs.view(-1, 32, 32, 3).shape"),
        code("datasets = (train_set, val_set)"),
        code("model = MLModel()"),
        code("""
config = {
    'seed': 32,
    'scheduler': None,
    'optimizer': 'sgd',
    'momentum': 0.9,
    'weight_decay': 0.0,
    'lr': 0.001,
    'criterion': 'cross_entropy',
    'metric': 'accuracy',
    'pred_function': 'softmax',
    'model_dir': 'model_output',
    'backend': 'smddp'
}
"""),
        code("trainer = Trainer(model, datasets=datasets, epochs=6, batch_size=32, progress=False, **config)"),
        code("""
# class histogram (the reference loops over the dataset; the labels are one tensor here)
counts = torch.bincount(train_set.targets, minlength=len(classes)).tolist()
class_count = {classes[i]: counts[i] for i in range(len(classes))}
class_count
"""),
        code("""
# one augmented training batch, as the model sees it (de-normalised for display)
for images, _ in trainer.train_loader:
    print('images.shape:', images.shape)
    mean = torch.tensor((0.4914, 0.4822, 0.4465)).view(1, 3, 1, 1)
    std = torch.tensor((0.2023, 0.1994, 0.2010)).view(1, 3, 1, 1)
    if images.dtype == torch.uint8:          # a CUDA run hands the loader's raw uint8 rows to the device kernel
        images = torch.stack([train_set.augment.apply_cpu(r, i, True) for i, r in enumerate(images)])
    grid = make_grid((images * std + mean).clamp(0, 1), nrow=16)
    show(torchvision.transforms.functional.to_pil_image(grid))
    break
"""),
        code("history = trainer.fit()\nfor k in ('train_loss', 'val_loss', 'train_metric', 'val_metric'):\n    print(k, [round(v, 4) for v in history[k]])"),
        code("trainer.save_history_('model_output')"),
        code("history = load_history('model_output')"),
        code("""
try:
    plot_history(history)
except ImportError as ex:
    print('matplotlib is not installed here (', ex, '); the curves as numbers:')
    for e, tl, vl, ta, va in zip(history['epochs'], history['train_loss'], history['val_loss'], history['train_metric'], history['val_metric']):
        print(f'epoch {e}: train loss {tl:.4f}  val loss {vl:.4f}  train acc {ta:.3f}  val acc {va:.3f}')
"""),
        code("""
model = MLModel()
model = load_model(model, 'model_output/model.pth')
test_loss, test_accuracy = trainer.test(model, trainer.val_loader)
"""),
        code("print(f'Model accuracy on test: {test_accuracy}')"),
    ]


def nb02():
    return [
        md("""
# 02 - Distributed training on one NVSwitch node

The reference's notebook 02 uploads CIFAR-10 to S3 and starts a SageMaker `PyTorch` estimator with
`distribution={'smdistributed': {'dataparallel': {'enabled': True}}}`; the estimator runs `main.py` once per GPU with the
hyper-parameters as command-line flags.  Here the same `main.py` with the same flags is started by the in-tree launcher
(`python -m ml_trainer_b200.parallel.launch`, one process per GPU, env:// contract, first failure stops the group).  On
GPUs the gradient exchange is the hand-written NVLink / NVSwitch kernels fused with the optimizer; without GPUs the same
command runs two ranks over gloo.
"""),
        code("""
import os, subprocess, sys
sys.path.insert(0, '..')
import torch
"""),
        code("""
hyperparameters = {
    'seed': 32,
    'optimizer': 'sgd',
    'momentum': 0.9,
    'lr': 0.001,
    'criterion': 'cross_entropy',
    'pred_function': 'softmax',
    'metric': 'accuracy',
    'custom_function': True,
    'backend': 'smddp',      # the reference's default string: the native NVLink backend on GPUs, gloo on CPUs
    'epochs': 2,
    'batch_size': 64,
}
"""),
        code("""
# instance_type / instance_count of the estimator become: how many local ranks
nproc = min(torch.cuda.device_count(), 8) if torch.cuda.is_available() else 2
data_flags = ['--data_dir', 'cifar10-dataset'] if os.path.isdir('../cifar10-dataset') else ['--synthetic', '--synthetic_train', '1024', '--synthetic_val', '256']
flags = [f'--{k}={v}' for k, v in hyperparameters.items()] + data_flags + ['--no_progress', '--model_dir', 'notebooks/model_output_dist']
cmd = [sys.executable, '-m', 'ml_trainer_b200.parallel.launch', '--nproc', str(nproc), 'main.py', *flags]
print(' '.join(cmd))
"""),
        code("""
env = dict(os.environ, MLT_LOG_LEVEL='INFO')
out = subprocess.run(cmd, cwd='..', env=env, capture_output=True, text=True)
print(out.stdout[-1500:])
print('\\n'.join(l for l in out.stderr.splitlines() if 'loss' in l or 'Training' in l or 'rank' in l)[-2500:])
out.returncode
"""),
        code("""
from src.utils.utils import load_history, plot_history
history = load_history('model_output_dist')
history
"""),
        code("""
try:
    plot_history(history)
except ImportError as ex:
    print('matplotlib is not installed here:', ex)
"""),
    ]


def nb03():
    return [
        md("""
# 03 - Testing a trained model
The reference's `03_ML_Testing.ipynb`, cell for cell: transforms -> CIFAR-10 test split -> `Loader` -> `load_model` ->
`Trainer(model)` -> `trainer.test`.  The checkpoint is the one notebook 01 wrote (`model_output/model.pth`); it loads through
the same `load_model` the reference ships (both the plain and the `module.`-prefixed key style).  Without `cifar10-dataset/` on
disk (no network here) the CIFAR-shaped synthetic validation set of notebook 01 stands in.
"""),
        code("""
import os, sys
sys.path.insert(0, '..')
import torchvision
from torchvision.utils import make_grid
from src.model import MLModel
from src.trainer import Trainer
from src.dataloader import Loader
from src.utils.utils import load_model
from ml_trainer_b200 import DeviceAugment, U8ImageDataset
from ml_trainer_b200.data import synthetic_cifar
"""),
        code("training_dir = 'cifar10-dataset'"),
        code("""
# Test-time transform: the reference's ToTensor + Normalize (its RandomCrop / flip belong to training only); one object that
# runs as a GPU kernel on uint8 batches, or bit for bit the same on the host
def _get_transforms():
    return DeviceAugment((3, 32, 32), padding=0, flip=False,
                         mean=(0.4914, 0.4822, 0.4465), std=(0.2023, 0.1994, 0.2010))
"""),
        code("""
if os.path.isdir(training_dir):
    raw_val = torchvision.datasets.CIFAR10(root=training_dir, train=False, download=False)
    val_set = U8ImageDataset.from_torchvision(raw_val, _get_transforms(), train=False)
else:
    print('no CIFAR-10 on disk: using the CIFAR-shaped synthetic stand-in of notebook 01')
    val_set = U8ImageDataset(*synthetic_cifar(1024, seed=1), augment=_get_transforms(), train=False)
"""),
        code("val_set.inputs.view(-1, 32, 32, 3).shape"),
        code("test_loader = Loader(val_set, batch_size=32, shuffle=True)"),
        code("""
model = MLModel()
model = load_model(model, 'model_output/model.pth')
"""),
        code("trainer = Trainer(model)"),
        code("test_loss, test_accuracy = trainer.test(model, test_loader)\nprint(f'test loss {test_loss:.4f}')"),
        code("print(f'Model accuracy on test: {test_accuracy}')"),
    ]


if __name__ == "__main__":
    which = sys.argv[1:] or ["01", "02", "03"]
    if "01" in which:
        write("01_ML_Training_local.ipynb", run_cells(nb01()))
    if "02" in which:
        write("02_ML_Training_distributed.ipynb", run_cells(nb02()))
    if "03" in which:
        write("03_ML_Testing.ipynb", run_cells(nb03()))
