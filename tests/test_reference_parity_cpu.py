"""End-to-end loss-curve parity with the REFERENCE Trainer itself ("End to end" tier; BASELINE.json config 1).

The unmodified reference (``baseline/_ref``, installed by ``__graft_entry__.build()``) trains MLModel on a seeded synthetic
tensor dataset in a subprocess (both packages are called ``src``, so they cannot share an interpreter); this repo's Trainer
then trains the same model on the same tensors with the same seed and flags.  Per-epoch train / validation loss and accuracy
and the final weights must agree: same initial weights (reference src/model.py via the same ``torch.manual_seed``), same shuffle
order (the reference's ``DataLoader(shuffle=True)`` draws, /root/reference/src/trainer.py:77-79), same SGD arithmetic."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")

REF_SCRIPT = r"""
import json, os, sys
ref, out, optimizer, scheduler = sys.argv[1], sys.argv[2], sys.argv[3], (sys.argv[4] if sys.argv[4] != "none" else None)
os.environ["TQDM_DISABLE"] = "1"
sys.path[:0] = [ref, os.path.join(ref, "_shims")]
import logging
logging.disable(logging.INFO)
import torch
from torch.utils.data import TensorDataset
from src.trainer import Trainer
from src.model import MLModel
g = torch.Generator().manual_seed(5)
xt, yt = torch.randn(256, 3, 32, 32, generator=g), torch.randint(0, 10, (256,), generator=g)
xv, yv = torch.randn(64, 3, 32, 32, generator=g), torch.randint(0, 10, (64,), generator=g)
torch.manual_seed(32)
model = MLModel()
t = Trainer(model, datasets=(TensorDataset(xt, yt), TensorDataset(xv, yv)), epochs=3, batch_size=32, seed=32, optimizer=optimizer,
            lr=0.01, momentum=0.9, criterion="cross_entropy", metric="accuracy", pred_function="softmax", model_dir=out,
            scheduler=scheduler)
os.makedirs(out, exist_ok=True)
t.fit()
json.dump({"train_loss": t.train_losses, "val_loss": t.val_losses, "train_metric": [float(v) for v in t.train_metrics],
           "val_metric": [float(v) for v in t.val_metrics]}, open(os.path.join(out, "ref_history.json"), "w"))
torch.save({k: v.cpu() for k, v in t.model.state_dict().items()}, os.path.join(out, "ref_weights.pt"))
"""


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference arm not installed (baseline/_ref)")
@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU statement of both arms (the GPU tier has its own oracles)")
@pytest.mark.parametrize("optimizer,scheduler", [("sgd", None), ("adamw", "StepLR"), ("adam", "CosineAnnealingWarmRestarts"),
                                                 ("adagrad", None)])
def test_loss_curves_match_the_reference_trainer(tmp_path, optimizer, scheduler):
    from torch.utils.data import TensorDataset

    from ml_trainer_b200 import MLModel, Trainer

    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "-c", REF_SCRIPT, REF, str(tmp_path / "ref"), optimizer, scheduler or "none"], env=env,
                       capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr[-2000:]
    ref_hist = json.load(open(tmp_path / "ref" / "ref_history.json"))
    ref_w = torch.load(tmp_path / "ref" / "ref_weights.pt")

    g = torch.Generator().manual_seed(5)
    xt, yt = torch.randn(256, 3, 32, 32, generator=g), torch.randint(0, 10, (256,), generator=g)
    xv, yv = torch.randn(64, 3, 32, 32, generator=g), torch.randint(0, 10, (64,), generator=g)
    torch.manual_seed(32)
    model = MLModel()
    t = Trainer(model, datasets=(TensorDataset(xt, yt), TensorDataset(xv, yv)), epochs=3, batch_size=32, seed=32, optimizer=optimizer,
                lr=0.01, momentum=0.9, criterion="cross_entropy", metric="accuracy", pred_function="softmax",
                model_dir=str(tmp_path / "ours"), progress=False, scheduler=scheduler)
    t.fit()
    for key, ours in (("train_loss", t.train_losses), ("val_loss", t.val_losses), ("train_metric", t.train_metrics),
                      ("val_metric", t.val_metrics)):
        assert len(ours) == len(ref_hist[key]) == 3
        for a, b in zip(ours, ref_hist[key]):
            assert abs(float(a) - float(b)) < 2e-4, (key, ours, ref_hist[key])
    for k, v in t.model.state_dict().items():
        assert torch.allclose(v.cpu(), ref_w[k], rtol=1e-4, atol=1e-5), k
