"""Prediction functions and metrics (reference src/trainer.py:115-121,160-172).

The reference computes accuracy by shipping targets and predictions to the host and
calling ``sklearn.metrics.accuracy_score`` every step (two D2H copies + a stream
sync).  Here every metric stays on the device and returns a 0-d tensor; softmax /
log-softmax are monotone so argmax is taken on the raw logits (the prediction
function is still exposed for callers that want probabilities).
"""
from __future__ import annotations

import torch

PRED_FUNCTIONS = ("softmax", "logsoftmax", None)
METRICS = ("accuracy", "mcrmse", None)


def get_prediction_function(kind):
    if kind == "logsoftmax":
        return torch.nn.LogSoftmax(dim=-1)
    if kind == "softmax":
        return torch.nn.Softmax(dim=-1)
    return None


def get_predictions(outputs: torch.Tensor, pred_function=None) -> torch.Tensor:
    # argmax(softmax(x)) == argmax(log_softmax(x)) == argmax(x); skip the extra pass.
    return torch.argmax(outputs, dim=-1)


def accuracy(outputs: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    """Fraction of rows whose argmax equals the target, as a device scalar."""
    if outputs.is_cuda and outputs.dim() == 2 and targets.dtype == torch.int64:
        from ..ops import functional as F

        if F.native_available():
            return F.argmax_accuracy(outputs, targets)
    return (torch.argmax(outputs, dim=-1) == targets).float().mean()


def mcrmse(outputs: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    """Mean column-wise RMSE (reference src/trainer.py:160-163); one kernel on the GPU (csrc/loss.cu)."""
    if outputs.is_cuda and outputs.dim() == 2 and outputs.shape == targets.shape and 0 < outputs.shape[1] <= 1024 \
            and outputs.dtype in (torch.float32, torch.bfloat16) and targets.dtype in (torch.float32, torch.bfloat16):
        from ..ops import functional as F

        if F.native_available():
            acc = torch.zeros(3, dtype=torch.float32, device=outputs.device)
            F.regression_loss(outputs, targets, "l2", acc, True, need_grad=False)
            return acc[1]
    colwise_mse = torch.mean(torch.square(targets.float() - outputs.float()), dim=0)
    return torch.mean(torch.sqrt(colwise_mse), dim=0)


def evaluate(metric, outputs, targets):
    if metric == "accuracy":
        return accuracy(outputs, targets)
    if metric == "mcrmse":
        return mcrmse(outputs, targets)
    if metric is None:
        return None
    raise KeyError(f"unknown metric {metric!r}; expected one of {METRICS}")
