"""Swap ``torch.nn`` layers of a user model for the op library, in place.

The reference lets the user bring any ``nn.Module`` (reference main.py:30).  ``convert_model``
keeps that contract: layers it knows are re-classed onto the fused implementations
(same parameters, same ``state_dict`` keys), everything else is left on stock PyTorch and
still trains through the generic gradient hooks.
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


def convert_model(model: nn.Module) -> nn.Module:
    if getattr(model, "_mlt_native", False):      # library models route their own forward
        return model
    for mod in model.modules():
        if type(mod) is nn.Linear:
            _reclass(mod, M.Linear, act=None)
    return model
