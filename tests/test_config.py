"""Config whitelist + factories (reference src/trainer.py:26-41,105-150,307-311)."""
import pytest
import torch

from ml_trainer_b200 import MLModel, Trainer, TrainerConfig
from ml_trainer_b200.config import REFERENCE_KEYS


def test_reference_defaults():
    c = TrainerConfig()
    assert c.reference_view() == dict(seed=32, scheduler=None, optimizer="sgd", momentum=0.9, weight_decay=0.0,
                                      lr=0.001, criterion="cross_entropy", metric="accuracy",
                                      pred_function="softmax", model_dir="model_output", backend="smddp")
    assert set(REFERENCE_KEYS) <= TrainerConfig.allowed()


def test_unknown_kwarg_raises_typeerror():
    with pytest.raises(TypeError) as e:
        Trainer(MLModel(), bogus=1)
    assert e.value.args == ("Keyword argument not understood:", "bogus")


def test_unknown_scheduler_keyerror():
    with pytest.raises(KeyError):
        Trainer(MLModel(), scheduler="Nope")


@pytest.mark.parametrize("name", ["sgd", "adam", "adagrad", "adamax", "adamw"])
def test_optimizer_factory(name):
    t = Trainer(MLModel(), optimizer=name, progress=False)
    assert t.optimizer.kind == name
    assert t.optimizer.numel >= 62006


def test_unknown_optimizer_is_an_error_not_none():
    with pytest.raises(ValueError):
        Trainer(MLModel(), optimizer="lion")


@pytest.mark.parametrize("name", ["cross_entropy", "neg-loss", "l1", "l2", "custom"])
def test_all_criteria_construct(name):
    # neg-loss / l2 / custom crash in the reference ctor (class objects / plain function get .to(device))
    t = Trainer(MLModel(), criterion=name, metric=None, progress=False)
    out = torch.randn(4, 10)
    if name in ("cross_entropy", "neg-loss"):
        tgt = torch.randint(0, 10, (4,))
        out = out.log_softmax(-1) if name == "neg-loss" else out
    else:
        tgt = torch.randn(4, 10)
    assert torch.isfinite(t.criterion(out, tgt))


@pytest.mark.parametrize("name", ["CosineAnnealingWarmRestarts", "ReduceLROnPlateau", "StepLR"])
def test_scheduler_table(name):
    t = Trainer(MLModel(), scheduler=name, progress=False)
    assert type(t.scheduler).__name__ == name


def test_test_only_trainer_has_no_loaders():
    t = Trainer(MLModel())
    assert t.train_loader is None and t.val_loader is None
    with pytest.raises(RuntimeError):
        t.fit()
