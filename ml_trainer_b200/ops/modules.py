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

    def forward(self, logits, targets):
        return F.cross_entropy(logits, targets)

    def fused_step(self, logits, targets, acc3=None, want_acc=False, need_grad=True):
        return F.softmax_xent(logits, targets, acc3, want_acc, need_grad)
