"""Optimizer update rules on flat fp32 slices.

These are the plain-PyTorch statements of the five rules the reference can select
(reference src/trainer.py:123-138: sgd / adam / adagrad / adamax / adamw, all with
torch defaults beyond lr / momentum / weight_decay).  They serve three purposes:
the CPU execution path, the oracle the CUDA kernels in ``csrc/optim_rules.cuh`` are
tested against, and executable documentation of the exact arithmetic
(bias-correction placement, eps placement, coupled vs decoupled weight decay).

Every rule has the signature ``rule(p, g, s1, s2, h)`` where ``p`` is the fp32
master slice, ``g`` the (already averaged) fp32 gradient slice, ``s1``/``s2`` the
state slices and ``h`` a :class:`Hyper`.  All updates are in place.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

RULE_IDS = {"sgd": 0, "adam": 1, "adamw": 2, "adagrad": 3, "adamax": 4}
# number of fp32 state slots each rule needs
RULE_STATES = {"sgd": 1, "adam": 2, "adamw": 2, "adagrad": 1, "adamax": 2}
DEFAULT_EPS = {"sgd": 0.0, "adam": 1e-8, "adamw": 1e-8, "adagrad": 1e-10, "adamax": 1e-8}
# torch.optim defaults: AdamW decays by 1e-2 unless told otherwise, but the reference
# always passes weight_decay explicitly (default 0.0), so no special-casing is needed.

HYPER_LEN = 8  # lr, wd, b1 (or momentum), b2, eps, bias_corr1, bias_corr2_sqrt, grad_scale


@dataclass
class Hyper:
    lr: float
    wd: float = 0.0
    b1: float = 0.9      # momentum for sgd
    b2: float = 0.999
    eps: float = 1e-8
    step: int = 1        # 1-based step count of the update being applied
    grad_scale: float = 1.0

    def pack(self) -> list:
        bc1 = 1.0 - self.b1 ** self.step
        bc2 = 1.0 - self.b2 ** self.step
        return [self.lr, self.wd, self.b1, self.b2, self.eps, bc1, math.sqrt(bc2), self.grad_scale]


def sgd(p, g, s1, s2, h: Hyper):
    if h.wd != 0.0:
        g = g.add(p, alpha=h.wd)
    if h.b1 != 0.0:
        # torch seeds the buffer with the first gradient; with a zero-initialised
        # buffer  b = mu*0 + g  gives the same value, so no first-step branch.
        s1.mul_(h.b1).add_(g)
        g = s1
    p.add_(g, alpha=-h.lr)


def _adam_core(p, g, m, v, h: Hyper):
    m.lerp_(g, 1.0 - h.b1)
    v.mul_(h.b2).addcmul_(g, g, value=1.0 - h.b2)
    bc1 = 1.0 - h.b1 ** h.step
    bc2_sqrt = math.sqrt(1.0 - h.b2 ** h.step)
    denom = (v.sqrt() / bc2_sqrt).add_(h.eps)
    p.addcdiv_(m, denom, value=-h.lr / bc1)


def adam(p, g, s1, s2, h: Hyper):
    if h.wd != 0.0:
        g = g.add(p, alpha=h.wd)          # L2 penalty folded into the gradient
    _adam_core(p, g, s1, s2, h)


def adamw(p, g, s1, s2, h: Hyper):
    if h.wd != 0.0:
        p.mul_(1.0 - h.lr * h.wd)          # decoupled decay
    _adam_core(p, g, s1, s2, h)


def adagrad(p, g, s1, s2, h: Hyper):
    if h.wd != 0.0:
        g = g.add(p, alpha=h.wd)
    s1.addcmul_(g, g, value=1.0)
    p.addcdiv_(g, s1.sqrt().add_(h.eps), value=-h.lr)   # lr_decay = 0


def adamax(p, g, s1, s2, h: Hyper):
    if h.wd != 0.0:
        g = g.add(p, alpha=h.wd)
    s1.lerp_(g, 1.0 - h.b1)
    torch.maximum(s2.mul_(h.b2), g.abs().add_(h.eps), out=s2)
    bc1 = 1.0 - h.b1 ** h.step
    p.addcdiv_(s1, s2, value=-h.lr / bc1)


RULES = {"sgd": sgd, "adam": adam, "adamw": adamw, "adagrad": adagrad, "adamax": adamax}


def apply_rule(kind: str, p, g, s1, s2, h: Hyper):
    """Apply ``kind`` in place.  ``g`` is scaled by ``h.grad_scale`` first."""
    g = g.float()
    if h.grad_scale != 1.0:
        g = g * h.grad_scale
    RULES[kind](p, g, s1, s2, h)
