import torch

from ml_trainer_b200.utils import metrics as M
from ml_trainer_b200.utils.functions import custom_loss_function


def test_accuracy_matches_sklearn():
    from sklearn.metrics import accuracy_score

    g = torch.Generator().manual_seed(0)
    out = torch.randn(64, 10, generator=g)
    tgt = torch.randint(0, 10, (64,), generator=g)
    ref = accuracy_score(tgt.numpy(), torch.argmax(torch.softmax(out, -1), -1).numpy())
    assert abs(M.accuracy(out, tgt).item() - ref) < 1e-7


def test_mcrmse():
    out, tgt = torch.randn(32, 6), torch.randn(32, 6)
    ref = torch.mean(torch.sqrt(torch.mean((tgt - out) ** 2, dim=0)))
    assert torch.allclose(M.mcrmse(out, tgt), ref)


def test_pred_functions():
    assert isinstance(M.get_prediction_function("softmax"), torch.nn.Softmax)
    assert isinstance(M.get_prediction_function("logsoftmax"), torch.nn.LogSoftmax)
    assert M.get_prediction_function(None) is None


def test_custom_loss():
    a, b = torch.randn(5, 3), torch.randn(5, 3)
    assert torch.allclose(custom_loss_function(a, b), torch.nn.functional.mse_loss(a, b))
