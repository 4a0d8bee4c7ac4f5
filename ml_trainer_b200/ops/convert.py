"""Swap ``torch.nn`` layers of a user model for the op library, in place.

The reference lets the user bring any ``nn.Module`` (reference main.py:30).  ``convert_model`` keeps that contract: layers it
knows are re-classed onto the fused implementations (same parameters, same ``state_dict`` keys), everything else is left on
stock PyTorch and still trains through the generic gradient hooks.

* ``nn.Linear``      -> :class:`ops.modules.Linear`       (CUDA-core kernel for tiny shapes, tcgen05 GEMM otherwise)
* ``nn.Conv2d``      -> :class:`ops.modules.Conv2d`       (implicit-GEMM convolution on channels-last data)
* ``nn.BatchNorm2d`` -> :class:`ops.modules.BatchNorm2d`  (NHWC statistics / apply kernels)
* ``nn.LayerNorm``   -> :class:`ops.modules.LayerNorm`
* ``nn.RMSNorm``     -> :class:`ops.modules.RMSNorm`
* a ``torchvision`` bottleneck ResNet (resnet50 / 101 / 152 and the wide variants) is recognised as a whole and re-classed onto
  :mod:`models.resnet` - same module tree and parameter names - so it runs NHWC end to end with BatchNorm + ReLU + residual
  fused, exactly like the library's own ``resnet50``.
"""
from __future__ import annotations

import torch
from torch import nn

from . import modules as M


def _reclass(mod: nn.Module, cls, **attrs):
    mod.__class__ = cls
    for k, v in attrs.items():
        setattr(mod, k, v)
    return mod


def _convert_torchvision_resnet(model: nn.Module) -> bool:
    """Re-class an unmodified ``torchvision.models.ResNet`` built from ``Bottleneck`` blocks onto the library's ResNet."""
    try:
        from torchvision.models import resnet as tv
    except Exception:  # torchvision absent: nothing to recognise
        return False
    if type(model) is not tv.ResNet:
        return False
    blocks = [m for layer in (model.layer1, model.layer2, model.layer3, model.layer4) for m in layer]
    if not blocks or any(type(b) is not tv.Bottleneck for b in blocks):
        return False
    for b in blocks:          # dense, undilated 3x3 (torchvision v1.5 puts the stride there), stock BatchNorm
        c2 = b.conv2
        if c2.groups != 1 or c2.dilation != (1, 1) or type(b.bn1) is not nn.BatchNorm2d or b.conv1.bias is not None:
            return False
    from ..models import resnet as ours

    for b in blocks:
        _reclass(b, ours.Bottleneck, stride=b.conv2.stride[0])
    _reclass(model, ours.ResNet)
    for m in model.modules():
        if isinstance(m, nn.Conv2d):
            m.weight._mlt_channels_last = True          # the optimizer arena keeps it as [Cout, KH, KW, Cin]
    return True


def convert_model(model: nn.Module) -> nn.Module:
    if getattr(model, "_mlt_native", False):      # library models route their own forward
        return model
    if _convert_torchvision_resnet(model):
        return model
    rms = getattr(nn, "RMSNorm", None)
    for mod in model.modules():
        t = type(mod)
        if t is nn.Linear:
            _reclass(mod, M.Linear, act=None)
        elif t is nn.Conv2d:
            _reclass(mod, M.Conv2d)
            if mod.groups == 1 and mod.in_channels % 8 == 0 and mod.out_channels % 8 == 0:
                mod.weight._mlt_channels_last = True
        elif t is nn.BatchNorm2d:
            _reclass(mod, M.BatchNorm2d)
        elif t is nn.LayerNorm:
            _reclass(mod, M.LayerNorm)
        elif rms is not None and t is rms and mod.elementwise_affine:
            _reclass(mod, M.RMSNorm)
    return model
