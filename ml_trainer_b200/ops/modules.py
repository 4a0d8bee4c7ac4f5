"""``nn.Module`` front ends for the op library.

They subclass the matching ``torch.nn`` layers so ``state_dict`` keys, initialisation and
constructor signatures are unchanged (checkpoints stay loadable by the reference,
SURVEY §3.5); only ``forward`` is re-routed to the fused kernels.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import functional as F


class Linear(nn.Linear):
    """``act(x W^T + b)`` with bias and activation fused into the GEMM epilogue."""

    def __init__(self, in_features, out_features, bias=True, act: Optional[str] = None, **kw):
        super().__init__(in_features, out_features, bias=bias, **kw)
        self.act = act

    def forward(self, x):
        return F.linear_act(x, self.weight, self.bias, self.act)

    def extra_repr(self):
        return super().extra_repr() + (f", act={self.act}" if self.act else "")


class RMSNorm(nn.Module):
    """Root-mean-square norm over the last dimension with a learned scale (``torch.nn.RMSNorm`` semantics, same
    ``weight`` parameter name); ``forward(x, residual=None)`` fuses an optional residual add."""

    def __init__(self, normalized_shape, eps: Optional[float] = None, elementwise_affine: bool = True, device=None, dtype=None):
        super().__init__()
        if isinstance(normalized_shape, int):
            normalized_shape = (normalized_shape,)
        if len(normalized_shape) != 1:
            raise ValueError("RMSNorm here normalises over the last dimension only")
        self.normalized_shape = tuple(normalized_shape)
        self.eps = eps
        self.elementwise_affine = elementwise_affine
        if elementwise_affine:
            self.weight = nn.Parameter(torch.ones(self.normalized_shape, device=device, dtype=dtype))
        else:
            self.register_buffer("weight", torch.ones(self.normalized_shape, device=device, dtype=dtype), persistent=False)

    def forward(self, x, residual=None):
        eps = self.eps if self.eps is not None else torch.finfo(x.dtype).eps
        return F.rms_norm(x, self.weight, eps, residual)

    def extra_repr(self):
        return f"{self.normalized_shape}, eps={self.eps}, elementwise_affine={self.elementwise_affine}"


class ConvReLUPool2d(nn.Conv2d):
    """Small-channel conv (stride 1, no padding) + bias [+ ReLU] [+ 2x2 max-pool] in one kernel."""

    def __init__(self, in_channels, out_channels, kernel_size, relu=True, pool=True, bias=True, **kw):
        super().__init__(in_channels, out_channels, kernel_size, bias=bias, **kw)
        if self.stride != (1, 1) or self.padding != (0, 0) or self.dilation != (1, 1) or self.groups != 1:
            raise ValueError("ConvReLUPool2d covers stride-1, unpadded, dense convolutions")
        self.relu, self.pool = relu, pool

    def forward(self, x):
        return F.conv_relu_pool(x, self.weight, self.bias, self.relu, self.pool)


class CrossEntropyLoss(nn.Module):
    """Mean softmax cross-entropy (reference criterion 'cross_entropy', src/trainer.py:141-142).

    ``fused_step`` is what the engine calls: loss, dlogits and the device-side accumulators in one
    kernel, bypassing an autograd node for the loss.
    """

    native_fused = True
    is_cross_entropy = True                 # the single-kernel model steps (MLModel) implement exactly this criterion
    metrics = (None, "accuracy")            # what the kernel can accumulate next to the loss

    def forward(self, logits, targets):
        return F.cross_entropy(logits, targets)

    def accepts(self, outputs, targets) -> bool:
        return outputs.dim() == 2 and targets.dtype == torch.int64 and targets.dim() == 1

    def fused_step(self, logits, targets, acc3=None, want_acc=False, need_grad=True):
        return F.softmax_xent(logits, targets, acc3, want_acc, need_grad)


class NLLLoss(nn.Module):
    """Mean negative log-likelihood on log-probabilities (reference criterion 'neg-loss', src/trainer.py:143-144)."""

    native_fused = True
    metrics = (None, "accuracy")

    def forward(self, logp, targets):
        if logp.dim() != 2 or targets.dtype != torch.int64:
            return torch.nn.functional.nll_loss(logp, targets)
        return F._FusedLossFn.apply(logp, targets, "nll")

    def accepts(self, outputs, targets) -> bool:
        return outputs.dim() == 2 and targets.dtype == torch.int64 and targets.dim() == 1

    def fused_step(self, logp, targets, acc3=None, want_acc=False, need_grad=True):
        return F.nll_loss(logp, targets, acc3, want_acc, need_grad)


class _RegressionLoss(nn.Module):
    """Mean L1 / L2 loss over all elements; the mcrmse metric rides in the same kernel."""

    native_fused = True
    metrics = (None, "mcrmse")
    kind = "l2"

    def forward(self, pred, target):
        if pred.shape != target.shape:
            d = pred - target
            return d.abs().mean() if self.kind == "l1" else (d * d).mean()
        return F._FusedLossFn.apply(pred, target, self.kind)

    def accepts(self, outputs, targets) -> bool:
        return targets.is_floating_point() and outputs.shape == targets.shape

    def fused_step(self, pred, target, acc3=None, want_metric=False, need_grad=True):
        return F.regression_loss(pred, target, self.kind, acc3, want_metric, need_grad)


class L1Loss(_RegressionLoss):
    """Reference criterion 'l1' (src/trainer.py:145-146)."""
    kind = "l1"


class MSELoss(_RegressionLoss):
    """Reference criterion 'l2' (src/trainer.py:147-148)."""
    kind = "l2"


class CustomLoss(_RegressionLoss):
    """Reference criterion 'custom' = custom_loss_function = mean((output - target)**2) (src/utils/functions.py:15-17)."""
    kind = "custom"
