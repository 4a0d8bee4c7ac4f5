"""FusedOptimizer (CPU path) against torch.optim for every rule the reference can select."""
import copy

import pytest
import torch

from ml_trainer_b200 import MLModel
from ml_trainer_b200.optim import FusedOptimizer

TORCH = {
    "sgd": lambda p, wd: torch.optim.SGD(p, lr=0.05, momentum=0.9, weight_decay=wd),
    "adam": lambda p, wd: torch.optim.Adam(p, lr=0.01, weight_decay=wd),
    "adamw": lambda p, wd: torch.optim.AdamW(p, lr=0.01, weight_decay=wd),
    "adagrad": lambda p, wd: torch.optim.Adagrad(p, lr=0.05, weight_decay=wd),
    "adamax": lambda p, wd: torch.optim.Adamax(p, lr=0.01, weight_decay=wd),
}


@pytest.mark.parametrize("kind", list(TORCH))
@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_rule_matches_torch(kind, wd):
    torch.manual_seed(0)
    ref_model = MLModel()
    model = copy.deepcopy(ref_model)
    ref_opt = TORCH[kind](ref_model.parameters(), wd)
    lr = ref_opt.param_groups[0]["lr"]
    opt = FusedOptimizer(model.parameters(), kind=kind, lr=lr, momentum=0.9, weight_decay=wd)
    x = torch.randn(16, 3, 32, 32)
    y = torch.randint(0, 10, (16,))
    for _ in range(5):
        for m, o in ((ref_model, ref_opt), (model, opt)):
            o.zero_grad()
            torch.nn.functional.cross_entropy(m(x), y).backward()
            o.step()
    for (n, a), (_, b) in zip(ref_model.named_parameters(), model.named_parameters()):
        assert torch.allclose(a, b, rtol=2e-5, atol=2e-6), f"{kind} wd={wd}: {n} diverged by {(a - b).abs().max()}"


def test_arena_layout_is_reverse_order_and_aligned():
    m = MLModel()
    opt = FusedOptimizer(m.parameters(), kind="sgd")
    plist = list(m.parameters())
    assert opt.params_in_order[0] is plist[-1]
    assert all(o % 128 == 0 for o in opt.offsets)
    for p in plist:
        assert p.data_ptr() >= opt.master.data_ptr()
        assert p._mlt_grad_sink.shape == p.shape


def test_bf16_shadow_keeps_fp32_master():
    m = MLModel()
    ref = {n: p.detach().clone() for n, p in m.named_parameters()}
    opt = FusedOptimizer(m.parameters(), kind="adamw", compute_dtype=torch.bfloat16)
    for n, p in m.named_parameters():
        assert p.dtype == torch.bfloat16
    sd = opt.master_state_dict(m)
    for n in ref:
        assert sd[n].dtype == torch.float32 and torch.equal(sd[n], ref[n])
