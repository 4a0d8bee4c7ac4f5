"""The single-kernel MLModel train step (csrc/lenet_step.cu) against the layer-kernel path and stock PyTorch."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _trainer(tmp_path, fused, optimizer="sgd", metric="accuracy", wd=0.0):
    from ml_trainer_b200 import Loader, MLModel, SyntheticDataset, Trainer

    torch.manual_seed(32)
    tr = SyntheticDataset((3, 32, 32), 10, length=160, seed=3)
    va = SyntheticDataset((3, 32, 32), 10, length=48, seed=4)
    t = Trainer(MLModel(), datasets=(tr, va), epochs=2, batch_size=32, model_dir=str(tmp_path), progress=False,
                optimizer=optimizer, lr=0.02, weight_decay=wd, metric=metric, graph=False, fused_step=fused, seed=32)
    t.train_loader = Loader(tr, batch_size=32, shuffle=False, pin_memory=True)
    t.val_loader = Loader(va, batch_size=32, shuffle=False, pin_memory=True)     # epoch loss = mean of batch means: fix the order
    return t


@pytest.mark.parametrize("optimizer,wd", [("sgd", 0.0), ("sgd", 0.01), ("adamw", 0.01), ("adam", 0.0), ("adagrad", 0.0), ("adamax", 0.0)])
def test_fused_step_matches_layer_kernels(tmp_path, optimizer, wd):
    a = _trainer(tmp_path / "a", True, optimizer, wd=wd)
    b = _trainer(tmp_path / "b", False, optimizer, wd=wd)
    assert a.engine.fused is not None and b.engine.fused is None
    a.fit()
    b.fit()
    assert a.train_losses == pytest.approx(b.train_losses, rel=2e-4)
    assert a.val_losses == pytest.approx(b.val_losses, rel=2e-4)
    assert a.train_metrics == pytest.approx(b.train_metrics, abs=1e-6)
    assert a.val_metrics == pytest.approx(b.val_metrics, abs=1e-6)
    atol = 2e-5 if optimizer == "sgd" else 0.02 * 10 * 0.3
    for (n, p), (_, q) in zip(a.model.named_parameters(), b.model.named_parameters()):
        assert torch.allclose(p, q, rtol=1e-3, atol=atol), (n, (p - q).abs().max())


def test_fused_step_one_step_exact_against_torch(tmp_path):
    """One SGD step, compared element-wise with stock PyTorch autograd (fp32, TF32 off)."""
    from ml_trainer_b200 import MLModel

    t = _trainer(tmp_path, True, "sgd")
    ref = copy.deepcopy(t.model).cpu()
    for p, q in zip(ref.parameters(), t.model.parameters()):
        p.data = q.detach().cpu().clone()
    x = torch.randn(32, 3, 32, 32)
    y = torch.randint(0, 10, (32,))
    opt = torch.optim.SGD(ref.parameters(), lr=0.02, momentum=0.9)
    loss = torch.nn.functional.cross_entropy(ref(x), y)
    loss.backward()
    opt.step()
    t.engine.reset_accumulators()
    t.engine.train_step(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    got_loss, got_acc, n = t.engine.read_accumulators()
    assert n == 1 and abs(got_loss - loss.item()) < 1e-5
    for (name, p), q in zip(ref.named_parameters(), t.model.parameters()):
        assert torch.allclose(p, q.cpu(), rtol=1e-4, atol=2e-6), (name, (p - q.cpu()).abs().max())


@pytest.mark.parametrize("B", [1, 7, 32, 200])
def test_fused_step_batch_sizes(tmp_path, B):
    t = _trainer(tmp_path, True)
    u = _trainer(tmp_path / "u", False)
    x = torch.randn(B, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (B,), device="cuda")
    for eng in (t.engine, u.engine):
        eng.reset_accumulators()
        eng.train_step(x, y)
        eng.train_step(x, y)
    torch.cuda.synchronize()
    assert t.engine.read_accumulators() == pytest.approx(u.engine.read_accumulators(), rel=1e-4)
    for p, q in zip(t.model.parameters(), u.model.parameters()):
        assert torch.allclose(p, q, rtol=1e-3, atol=1e-5)


def test_native_executor_matches_python_loop(tmp_path):
    """Whole epochs in the C++ executor (gather thread + copy stream + single-kernel steps) == the Python loop."""
    a = _trainer(tmp_path / "a", True)
    b = _trainer(tmp_path / "b", True)
    b.cfg.native_loop = False
    assert a._native_ok(a.train_loader, a.model) and not b._native_ok(b.train_loader, b.model)
    a.fit()
    b.fit()
    assert a.train_losses == pytest.approx(b.train_losses, rel=1e-6)
    assert a.val_losses == pytest.approx(b.val_losses, rel=1e-6)
    assert a.train_metrics == pytest.approx(b.train_metrics, abs=1e-7)
    for p, q in zip(a.model.parameters(), b.model.parameters()):
        assert torch.equal(p, q)
    assert a.last_native_stats["h2d_bytes"] > 0 and len(a.last_step_losses) == len(a.train_loader)


def test_native_executor_honours_lr_schedule(tmp_path):
    from ml_trainer_b200 import Loader, MLModel, SyntheticDataset, Trainer

    outs = []
    for native in (True, False):
        torch.manual_seed(32)
        tr = SyntheticDataset((3, 32, 32), 10, length=160, seed=3)
        t = Trainer(MLModel(), datasets=(tr, tr), epochs=2, batch_size=32, model_dir=str(tmp_path / str(native)),
                    progress=False, lr=0.05, scheduler="CosineAnnealingWarmRestarts", native_loop=native, seed=32)
        t.train_loader = Loader(tr, batch_size=32, shuffle=False, pin_memory=True)
        t.val_loader = Loader(tr, batch_size=32, shuffle=False, pin_memory=True)
        t.fit()
        outs.append(t)
    assert outs[0].optimizer.param_groups[0]["lr"] == pytest.approx(outs[1].optimizer.param_groups[0]["lr"])
    for p, q in zip(outs[0].model.parameters(), outs[1].model.parameters()):
        assert torch.equal(p, q)


@pytest.mark.parametrize("optimizer", ["sgd", "adamw"])
def test_multi_step_launch_equals_single_steps(tmp_path, optimizer):
    """K steps in ONE launch over a device pool == K single-step launches (same kernel body, in-kernel barriers between steps)."""
    K, P, B = 7, 3, 32
    torch.manual_seed(5)
    xp = torch.randn(P, B, 3, 32, 32, device="cuda")
    yp = torch.randint(0, 10, (P, B), device="cuda")
    ts = []
    for _ in range(2):
        t = _trainer(tmp_path / str(len(ts)), True, optimizer)
        ts.append(t)
    a, b = ts
    for (p, q) in zip(a.model.parameters(), b.model.parameters()):
        q.data.copy_(p.data)
    b.optimizer.master.copy_(a.optimizer.master)
    fa, fb = a.engine.fused, b.engine.fused
    # per-step hyper rows (AdamW bias corrections change every step)
    rows = []
    la = torch.zeros(K, device="cuda")
    for i in range(K):
        h = a.optimizer.begin_step(1.0)
        rows.append(torch.tensor(h.pack(), dtype=torch.float32))
        fa(xp[i % P], yp[i % P], True, None, False, la[i:i + 1])
    table = torch.stack(rows).cuda()
    lb = torch.zeros(K, device="cuda")
    acc = torch.zeros(3, device="cuda")
    fb.run_pool(xp, yp, K, True, hyper_table=table, losses=lb, acc3=acc, want_acc=True)
    torch.cuda.synchronize()
    assert torch.equal(la, lb), (la, lb)
    assert torch.equal(a.optimizer.master, b.optimizer.master)
    assert acc[2].item() == K and abs(acc[0].item() - lb.sum().item()) < 1e-4


@pytest.mark.parametrize("persistent", ["1", "0"])
def test_native_executor_persistent_and_per_step_modes(tmp_path, monkeypatch, persistent):
    """Executor kinds 2 (one persistent launch per run) and 0 (one launch per step), ragged last batch, cosine LR per step."""
    from ml_trainer_b200 import Loader, MLModel, SyntheticDataset, Trainer

    monkeypatch.setenv("MLT_LENET_PERSISTENT", persistent)
    outs = []
    for native in (True, False):
        torch.manual_seed(32)
        tr = SyntheticDataset((3, 32, 32), 10, length=150, seed=3)        # 4 x 32 + 22
        va = SyntheticDataset((3, 32, 32), 10, length=70, seed=4)
        t = Trainer(MLModel(), datasets=(tr, va), epochs=3, batch_size=32, model_dir=str(tmp_path / f"{persistent}{native}"), progress=False,
                    optimizer="adamw", lr=0.01, metric="accuracy", scheduler="CosineAnnealingWarmRestarts", native_loop=native, seed=32)
        t.train_loader = Loader(tr, batch_size=32, shuffle=False, pin_memory=True)
        t.val_loader = Loader(va, batch_size=32, shuffle=False, pin_memory=True)
        t.fit()
        outs.append(t)
    a, b = outs
    assert a._nloop.last_kind == (2 if persistent == "1" else 0)
    assert a.train_losses == pytest.approx(b.train_losses, rel=1e-5)
    assert a.val_losses == pytest.approx(b.val_losses, rel=1e-5)
    assert a.val_metrics == pytest.approx(b.val_metrics, abs=1e-6)
    for p, q in zip(a.model.parameters(), b.model.parameters()):
        assert torch.allclose(p, q, rtol=1e-4, atol=1e-6)
