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


# ---- drop-in layers for user models (convert_model) -----------------------------------------------------------------------
# A user's CNN computes in logical NCHW.  These layers keep that contract but hold the data channels-last: a channels_last
# NCHW tensor IS a contiguous NHWC tensor, so the NHWC kernels read and write it without a copy and the logical shape the rest
# of the user's forward() sees never changes.  They take the native path for bf16 CUDA tensors and defer to torch otherwise.
def _as_nhwc(x: torch.Tensor) -> torch.Tensor:
    return x.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)


class Conv2d(nn.Conv2d):
    """``nn.Conv2d`` on the implicit-GEMM tcgen05 path (dense, undilated, symmetric padding, channels multiple of 8)."""

    def _native(self, x) -> bool:
        k, s, p = self.kernel_size, self.stride, self.padding
        return (x.is_cuda and x.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16 and F.native_available()
                and self.groups == 1 and self.dilation == (1, 1) and k[0] == k[1] and s[0] == s[1] and isinstance(p, tuple)
                and p[0] == p[1] and self.padding_mode == "zeros" and self.in_channels % 8 == 0 and self.out_channels % 8 == 0)

    def forward(self, x):
        if not self._native(x):
            return super().forward(x)
        from . import conv as C

        y = C.conv2d_nhwc(_as_nhwc(x), self.weight, self.stride[0], self.padding[0])
        if self.bias is not None:
            y = y + self.bias
        return y.permute(0, 3, 1, 2)


class BatchNorm2d(nn.BatchNorm2d):
    """``nn.BatchNorm2d`` through the NHWC statistics / apply kernels (training and evaluation mode)."""

    def forward(self, x):
        if not (x.is_cuda and x.dtype == torch.bfloat16 and F.native_available() and self.affine and self.track_running_stats
                and self.num_features % 8 == 0):
            return super().forward(x)
        from . import conv as C

        return C.batch_norm_nhwc(_as_nhwc(x), self).permute(0, 3, 1, 2)


class LayerNorm(nn.LayerNorm):
    """``nn.LayerNorm`` over the last dimension through the fused kernels (bf16 activations, width a multiple of 8 <= 1024)."""

    def forward(self, x, residual=None):
        if (len(self.normalized_shape) == 1 and self.elementwise_affine and self.bias is not None and x.is_cuda
                and x.dtype == torch.bfloat16 and residual is None):
            return F.layer_norm(x, self.weight, self.bias, self.eps)
        if residual is not None:
            return F.layer_norm(x, self.weight, self.bias, self.eps, residual)
        return super().forward(x)
