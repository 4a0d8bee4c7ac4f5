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


def test_hyper_rows_match_per_step_pack():
    """The vectorised per-epoch hyper table == Hyper.pack() evaluated step by step (what the native executors upload)."""
    import torch

    from ml_trainer_b200.optim.fused import FusedOptimizer

    for kind in ("sgd", "adam", "adamw", "adamax", "adagrad"):
        p = [torch.nn.Parameter(torch.randn(300))]
        o = FusedOptimizer(p, kind, lr=0.02, weight_decay=0.01)
        o.step_count = 7
        lrs = [0.02 * (0.9 ** i) for i in range(5)]
        rows = o.hyper_rows(5, 0.25, lrs)
        want = []
        for i in range(5):
            o.param_groups[0]["lr"] = lrs[i]
            want.append(o.hyper_for_next_step(0.25).pack())
            o.step_count += 1
        assert torch.allclose(rows, torch.tensor(want, dtype=torch.float32), rtol=1e-6, atol=0), kind


def test_packed_parameters_are_contiguous_views():
    """param._mlt_pack: members sit back to back in the arena and one view covers them (BERT's Q/K/V projections)."""
    import torch

    from ml_trainer_b200.optim.fused import FusedOptimizer

    torch.manual_seed(0)
    lin = [torch.nn.Linear(128, 128) for _ in range(3)]
    other = torch.nn.Linear(128, 64)
    for i, l in enumerate(lin):
        l.weight._mlt_pack = (("w", 1), i, 3)
        l.bias._mlt_pack = (("b", 1), i, 3)
    params = [p for m in (lin[0], other, lin[1], lin[2]) for p in m.parameters()]
    before = [p.detach().clone() for p in params]
    opt = FusedOptimizer(params, "sgd", lr=0.1)
    for p, q in zip(params, before):
        assert torch.equal(p.detach(), q)                      # values survive the re-homing
    pk = lin[0].weight._mlt_packed
    assert pk.data.shape == (384, 128) and pk.grad.shape == (384, 128)
    for i, l in enumerate(lin):
        assert pk.data[128 * i:128 * (i + 1)].data_ptr() == l.weight.data_ptr()
        assert torch.equal(pk.data[128 * i:128 * (i + 1)], l.weight.detach())
        assert pk.grad[128 * i:128 * (i + 1)].data_ptr() == l.weight._mlt_grad_sink.data_ptr()
    bk = lin[0].bias._mlt_packed
    assert bk.data.shape == (384,) and bk.data[128:256].data_ptr() == lin[1].bias.data_ptr()
    # an incomplete pack (one member missing from the optimizer) is ignored
    a, b = torch.nn.Linear(128, 128), torch.nn.Linear(128, 128)
    a.weight._mlt_pack = (("w", 2), 0, 3)
    b.weight._mlt_pack = (("w", 2), 1, 3)
    FusedOptimizer(list(a.parameters()) + list(b.parameters()), "sgd", lr=0.1)
    assert not hasattr(a.weight, "_mlt_packed")
