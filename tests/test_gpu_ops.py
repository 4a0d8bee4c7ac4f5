"""Each sm_100a kernel against a plain PyTorch fp32 statement of the same op (SURVEY §4 GPU kernel tier)."""
import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda", 0)


def test_native_library_is_loaded():
    from ml_trainer_b200.ops import _lib

    assert _lib.native_available()
    assert _lib.lib().mlt_version() >= 100


@pytest.mark.parametrize("B,Cin,H,Cout,K", [(32, 3, 32, 6, 5), (32, 6, 14, 16, 5), (4, 3, 32, 6, 5), (5, 2, 11, 3, 3),
                                            (1, 1, 6, 1, 5), (7, 4, 9, 5, 2)])
@pytest.mark.parametrize("relu,pool", [(True, True), (False, False), (True, False), (False, True)])
def test_conv_relu_pool_fwd_bwd(B, Cin, H, Cout, K, relu, pool):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    dev = _dev()
    x = torch.randn(B, Cin, H, H, device=dev, requires_grad=True)
    w = torch.randn(Cout, Cin, K, K, device=dev, requires_grad=True) * 0.2
    w = w.detach().requires_grad_()
    b = torch.randn(Cout, device=dev, requires_grad=True)
    y = F.conv_relu_pool(x, w, b, relu=relu, pool=pool)
    ref = TF.conv2d(x, w, b)
    if relu:
        ref = TF.relu(ref)
    if pool:
        ref = TF.max_pool2d(ref, 2, 2)
    assert y.shape == ref.shape
    assert torch.allclose(y, ref, rtol=1e-4, atol=1e-4), (y - ref).abs().max()
    g = torch.randn_like(ref)
    gx, gw, gb = torch.autograd.grad(y, (x, w, b), g)
    rx, rw, rb = torch.autograd.grad(ref, (x, w, b), g)
    assert torch.allclose(gx, rx, rtol=1e-4, atol=1e-4), (gx - rx).abs().max()
    assert torch.allclose(gw, rw, rtol=1e-3, atol=1e-3), (gw - rw).abs().max()
    assert torch.allclose(gb, rb, rtol=1e-3, atol=1e-3), (gb - rb).abs().max()


def test_conv_bwd_scratch_self_resets():
    """The wgrad ticket/partial scratch must be reusable launch after launch without re-zeroing."""
    from ml_trainer_b200.ops import functional as F

    dev = _dev()
    torch.manual_seed(1)
    w = (torch.randn(6, 3, 5, 5, device=dev) * 0.2).requires_grad_()
    b = torch.zeros(6, device=dev, requires_grad=True)
    for i in range(4):
        x = torch.randn(32, 3, 32, 32, device=dev)
        y = F.conv_relu_pool(x, w, b)
        ref = TF.max_pool2d(TF.relu(TF.conv2d(x, w, b)), 2, 2)
        g = torch.randn_like(ref)
        gw, = torch.autograd.grad(y, w, g)
        rw, = torch.autograd.grad(ref, w, g)
        assert torch.allclose(gw, rw, rtol=1e-3, atol=1e-3), (i, (gw - rw).abs().max())


@pytest.mark.parametrize("B,K,N", [(32, 400, 120), (32, 120, 84), (32, 84, 10), (4, 400, 120), (3, 17, 5), (130, 33, 7)])
@pytest.mark.parametrize("act", [None, "relu", "tanh"])
def test_linear_small_fwd_bwd(B, K, N, act):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    dev = _dev()
    x = torch.randn(B, K, device=dev, requires_grad=True)
    w = (torch.randn(N, K, device=dev) * 0.1).requires_grad_()
    b = torch.randn(N, device=dev, requires_grad=True)
    y = F.linear_act(x, w, b, act)
    ref = F._torch_act(TF.linear(x, w, b), act)
    assert torch.allclose(y, ref, rtol=1e-4, atol=1e-4), (y - ref).abs().max()
    g = torch.randn_like(ref)
    got = torch.autograd.grad(y, (x, w, b), g)
    exp = torch.autograd.grad(ref, (x, w, b), g)
    for a, e, name in zip(got, exp, "xwb"):
        assert torch.allclose(a, e, rtol=1e-3, atol=1e-3), (name, (a - e).abs().max())


@pytest.mark.parametrize("B,C", [(32, 10), (4, 10), (256, 1000), (32, 2), (1000, 33), (5000, 10)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_softmax_xent(B, C, dtype):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    dev = _dev()
    logits = (torch.randn(B, C, device=dev) * 3).to(dtype)
    tgt = torch.randint(0, C, (B,), device=dev)
    acc = torch.zeros(3, device=dev)
    for rep in range(2):                                  # twice: the workspace must self-reset
        loss, dl = F.softmax_xent(logits, tgt, acc, want_acc=True)
    lf = logits.float().requires_grad_()
    ref = TF.cross_entropy(lf, tgt)
    rdl, = torch.autograd.grad(ref, lf)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert abs(loss.item() - ref.item()) < 1e-4 * max(1.0, abs(ref.item()))
    assert torch.allclose(dl.float(), rdl, rtol=tol, atol=tol / 10 if dtype == torch.float32 else 1e-3)
    ref_acc = (lf.argmax(-1) == tgt).float().mean().item()
    a = acc.tolist()
    assert abs(a[0] - 2 * ref.item()) < 2e-4 * max(1.0, abs(ref.item())) and abs(a[1] - 2 * ref_acc) < 1e-5 and a[2] == 2.0
    assert abs(F.argmax_accuracy(logits, tgt).item() - ref_acc) < 1e-6
    # autograd front end
    l2 = logits.float().requires_grad_()
    out = F.cross_entropy(l2, tgt)
    out.backward()
    assert torch.allclose(l2.grad, rdl, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("kind", ["sgd", "adam", "adamw", "adagrad", "adamax"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_fused_update_matches_rules(kind, dtype, wd):
    """CUDA rule kernels vs the PyTorch oracle in optim/rules.py (itself tested against torch.optim on CPU)."""
    from ml_trainer_b200.optim import FusedOptimizer

    torch.manual_seed(0)
    shapes = [(300, 7), (129,), (64, 64), (5,)]
    def make(device):
        torch.manual_seed(0)
        ps = [torch.nn.Parameter(torch.randn(s).to(device)) for s in shapes]      # same values on both devices
        return ps, FusedOptimizer(ps, kind=kind, lr=0.01, momentum=0.9, weight_decay=wd, compute_dtype=dtype)
    pc, oc = make("cpu")
    pg, og = make(_dev())
    for step in range(4):
        torch.manual_seed(100 + step)
        g = torch.randn(oc.numel).to(dtype)
        oc.grad.copy_(g)
        og.grad.copy_(g.to(_dev()))
        oc.begin_step(); oc.step_range(0, oc.numel)
        og.begin_step(); og.step_range(0, og.numel)
    assert torch.allclose(og.master.cpu(), oc.master, rtol=1e-5, atol=1e-6), (og.master.cpu() - oc.master).abs().max()
    assert torch.allclose(og.s1.cpu(), oc.s1, rtol=1e-5, atol=1e-6)
    if dtype == torch.bfloat16:
        assert torch.equal(og.model.cpu(), og.master.cpu().to(torch.bfloat16))
        assert pg[0].dtype == torch.bfloat16


def test_fused_update_partial_ranges():
    from ml_trainer_b200.optim import FusedOptimizer

    dev = _dev()
    ps = [torch.nn.Parameter(torch.randn(1000, device=dev)) for _ in range(3)]
    opt = FusedOptimizer(ps, kind="sgd", lr=0.1, momentum=0.0)
    opt.grad.fill_(1.0)
    before = opt.master.clone()
    opt.begin_step()
    lo, hi = opt.span(1, 2)
    opt.step_range(lo, hi)
    torch.cuda.synchronize()
    delta = opt.master - before
    assert torch.allclose(delta[lo:hi], torch.full_like(delta[lo:hi], -0.1))
    assert delta[:lo].abs().max() == 0 and delta[hi:].abs().max() == 0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_dropout_philox_mask(dtype):
    """Keep rate, scaling, backward regenerates the forward's mask, masks change with the device-side step counter."""
    from ml_trainer_b200.ops import functional as F

    dev = torch.device("cuda", 0)
    p = 0.25
    x = torch.ones(1 << 20, device=dev, dtype=dtype).requires_grad_()
    F.dropout_begin_step(dev)
    y = F.dropout(x, p, True)
    keep = (y != 0).float().mean().item()
    assert abs(keep - (1 - p)) < 5e-3, keep
    nz = y[y != 0].float()
    assert torch.allclose(nz, torch.full_like(nz, 1 / (1 - p)), rtol=1e-2)
    g = torch.full_like(y, 2.0)
    y.backward(g)
    assert torch.equal((x.grad != 0), (y != 0))            # same mask in backward
    assert torch.allclose(x.grad[x.grad != 0].float(), torch.full_like(nz, 2 / (1 - p)), rtol=1e-2)
    F.dropout_begin_step(dev)
    y2 = F.dropout(x.detach(), p, True)
    diff = ((y2 != 0) != (y != 0)).float().mean().item()
    assert 0.25 < diff < 0.5, diff                         # independent masks: 2 p (1 - p) = 0.375
    assert F.dropout(x, p, False) is x


# ---- the reference's other criteria and metric (src/trainer.py:140-150, 160-163) ---------------------------------------
@pytest.mark.parametrize("kind", ["l1", "l2", "custom"])
@pytest.mark.parametrize("shape,dtype", [((32, 6), torch.float32), ((257, 10), torch.float32), ((4096, 33), torch.bfloat16),
                                         ((7,), torch.float32)])
def test_regression_loss_fwd_bwd_mcrmse(kind, shape, dtype):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    pred = torch.randn(*shape, device="cuda").to(dtype)
    target = torch.randn(*shape, device="cuda")
    want_m = len(shape) == 2
    acc = torch.tensor([1.0, 2.0, 3.0], device="cuda")
    for _ in range(2):                                   # twice: the workspace must come back zeroed
        loss, dp = F.regression_loss(pred, target, kind, acc, want_m, need_grad=True)
    p32 = pred.float().requires_grad_(True)
    d = p32 - target
    ref = d.abs().mean() if kind == "l1" else (d * d).mean()
    ref.backward()
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert abs(loss.item() - ref.item()) < 1e-5 * max(1.0, abs(ref.item()))
    assert torch.allclose(dp.float(), p32.grad, rtol=tol, atol=tol * float(p32.grad.abs().max()))
    m_ref = torch.mean(torch.sqrt(torch.mean((pred.float() - target) ** 2, dim=0))).item() if want_m else 0.0
    exp = [1.0 + 2 * ref.item(), 2.0 + 2 * m_ref, 5.0]
    assert acc.tolist() == pytest.approx(exp, rel=2e-5)


@pytest.mark.parametrize("B,C,dtype", [(32, 10, torch.float32), (513, 1000, torch.float32), (64, 2, torch.bfloat16)])
def test_nll_loss_fwd_bwd(B, C, dtype):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(1)
    logp = torch.log_softmax(torch.randn(B, C, device="cuda"), -1).to(dtype)
    y = torch.randint(0, C, (B,), device="cuda")
    acc = torch.zeros(3, device="cuda")
    loss, dl = F.nll_loss(logp, y, acc, True, need_grad=True)
    l32 = logp.float().requires_grad_(True)
    ref = torch.nn.functional.nll_loss(l32, y)
    ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-5 * max(1.0, abs(ref.item()))
    assert torch.allclose(dl.float(), l32.grad, rtol=1e-2 if dtype == torch.bfloat16 else 1e-6, atol=1e-7)
    a_ref = (logp.float().argmax(-1) == y).float().mean().item()
    assert acc.tolist() == pytest.approx([ref.item(), a_ref, 1.0], rel=1e-5)


@pytest.mark.parametrize("criterion,metric", [("l1", "mcrmse"), ("l2", "mcrmse"), ("custom", None), ("neg-loss", "accuracy")])
def test_trainer_criteria_run_native(tmp_path, criterion, metric):
    """Every criterion of the reference's table trains through the fused loss kernels (launch counters prove it)."""
    from ml_trainer_b200 import Trainer
    from ml_trainer_b200.ops import _lib
    from torch.utils.data import TensorDataset

    torch.manual_seed(0)
    x = torch.randn(128, 16)
    if criterion == "neg-loss":
        y = torch.randint(0, 6, (128,))
        model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.ReLU(), torch.nn.Linear(32, 6), torch.nn.LogSoftmax(-1))
    else:
        y = torch.randn(128, 6)
        model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.ReLU(), torch.nn.Linear(32, 6))
    ds = TensorDataset(x, y)
    _lib.launches.reset()
    tr = Trainer(model, datasets=(ds, ds), epochs=3, batch_size=32, model_dir=str(tmp_path), progress=False, criterion=criterion,
                 metric=metric, pred_function=None, lr=0.05, graph=False)
    h = tr.fit()
    names = dict(_lib.launches.by_name)
    assert names.get("nll_loss" if criterion == "neg-loss" else "regression_loss", 0) >= 3 * 4, names
    assert h["train_loss"][-1] < h["train_loss"][0]
    if metric:
        assert len(h["train_metric"]) == 3 and all(v == v for v in h["train_metric"])


# ---- device-side image augmentation (csrc/augment.cu; reference pipeline src/utils/functions.py:5-12) ----------------------
@pytest.mark.parametrize("layout", ["hwc", "chw"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_augment_u8_kernel_matches_host_statement(layout, dtype):
    from ml_trainer_b200 import DeviceAugment

    g = torch.Generator().manual_seed(3)
    B = 37
    shape = (B, 32, 32, 3) if layout == "hwc" else (B, 3, 32, 32)
    data = torch.randint(0, 256, shape, dtype=torch.uint8, generator=g)
    aug = DeviceAugment((3, 32, 32), padding=4, seed=2024, layout=layout)
    out = aug.apply_cuda(data.cuda().reshape(B, -1), pos0=1000, train=True, dtype=dtype)
    want = torch.stack([aug.apply_cpu(data[b], 1000 + b, True) for b in range(B)])
    tol = 1e-6 if dtype == torch.float32 else 2e-2
    assert out.shape == (B, 3, 32, 32) and out.dtype == dtype
    assert torch.allclose(out.float().cpu(), want, atol=tol, rtol=tol)
    ev = aug.apply_cuda(data.cuda().reshape(B, -1), pos0=0, train=False, dtype=dtype)
    want_ev = torch.stack([aug.apply_cpu(data[b], None, False) for b in range(B)])
    assert torch.allclose(ev.float().cpu(), want_ev, atol=tol, rtol=tol)
    # the stream position advances by itself when none is given
    aug.position = 5
    aug.apply_cuda(data.cuda().reshape(B, -1))
    assert aug.position == 5 + B


def test_native_epoch_on_uint8_dataset_matches_python_loop(tmp_path):
    """The C++ executor with the in-loop augmentation kernel trains exactly like the Python loop that augments batch by batch
    (same stream positions -> same crops and flips), and uploads 4x fewer bytes than the fp32 path."""
    from ml_trainer_b200 import DeviceAugment, MLModel, Trainer, U8ImageDataset

    g = torch.Generator().manual_seed(11)
    data = torch.randint(0, 256, (256, 32, 32, 3), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (256,), generator=g)

    def run(native):
        torch.manual_seed(32)
        ds = U8ImageDataset(data, labels, augment=DeviceAugment((3, 32, 32), padding=4, seed=9))
        va = U8ImageDataset(data[:64], labels[:64], augment=DeviceAugment((3, 32, 32), padding=0, flip=False), train=False)
        tr = Trainer(MLModel(), datasets=(ds, va), epochs=2, batch_size=32, model_dir=str(tmp_path / str(native)), progress=False,
                     lr=0.01, native_loop=native, seed=32)
        from ml_trainer_b200 import Loader

        tr.train_loader = Loader(ds, batch_size=32, shuffle=False, pin_memory=True)
        tr.fit()
        return tr

    a, b = run(True), run(False)
    assert a.last_native_stats["h2d_bytes"] // 8 == 32 * (3072 + 8) + 32       # uint8 rows + int64 labels + the hyper row
    assert a.train_losses == pytest.approx(b.train_losses, rel=1e-5)
    assert a.val_losses == pytest.approx(b.val_losses, rel=1e-5)
    for p, q in zip(a.model.parameters(), b.model.parameters()):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6)


# ---- RMSNorm (csrc/rmsnorm.cu) -------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,H", [(64, 768), (1000, 1024), (33, 8), (257, 4096), (16, 8192), (5, 1032)])
@pytest.mark.parametrize("pdtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("with_res", [False, True])
def test_rmsnorm_fwd_bwd(T, H, pdtype, with_res):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(T + H)
    x = torch.randn(T, H, device="cuda").bfloat16().requires_grad_(True)
    res = torch.randn(T, H, device="cuda").bfloat16().requires_grad_(True) if with_res else None
    gamma = (1.0 + 0.1 * torch.randn(H, device="cuda")).to(pdtype).requires_grad_(True)
    for _ in range(2):                                   # twice: the scratch accumulator must come back zeroed
        for t in (x, res, gamma):
            if t is not None:
                t.grad = None
        y = F.rms_norm(x, gamma, 1e-6, res)
        dy = torch.randn_like(y)
        y.backward(dy)
    xf = x.detach().float().requires_grad_(True)
    rf = res.detach().float().requires_grad_(True) if with_res else None
    gf = gamma.detach().float().requires_grad_(True)
    h = xf + rf if with_res else xf
    if with_res:
        h = h + (h.bfloat16().float() - h).detach()      # the kernel rounds the pre-norm sum once (what backward reads)
    yf = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + 1e-6) * gf
    yf.backward(dy.float())
    assert torch.allclose(y.float(), yf, rtol=2e-2, atol=2e-2)
    assert torch.allclose(x.grad.float(), xf.grad, rtol=3e-2, atol=3e-2)
    if with_res:
        assert torch.allclose(res.grad.float(), rf.grad, rtol=3e-2, atol=3e-2)
    scale = float(gf.grad.abs().max())
    assert torch.allclose(gamma.grad.float(), gf.grad, rtol=3e-2, atol=(2e-2 if pdtype == torch.bfloat16 else 5e-3) * scale + 1e-3)
