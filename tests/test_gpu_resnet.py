"""NHWC conv (im2col + tcgen05 GEMM), BatchNorm, pooling kernels and the native ResNet path against fp32 PyTorch."""
import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _close(a, b, tol, what):
    err = (a.float() - b.float()).abs().max().item()
    ref = max(1.0, b.float().abs().max().item())
    assert err <= tol * ref, f"{what}: max abs err {err} (ref scale {ref})"


@pytest.mark.parametrize("B,H,Cin,Cout,K,stride,pad", [(2, 16, 16, 32, 1, 1, 0), (2, 16, 16, 32, 3, 1, 1), (3, 15, 8, 24, 3, 2, 1),
                                                      (2, 32, 8, 64, 7, 2, 3), (2, 14, 64, 128, 1, 2, 0), (4, 56, 64, 64, 3, 1, 1),
                                                      (3, 28, 128, 128, 3, 2, 1), (2, 14, 256, 256, 3, 1, 1), (5, 9, 64, 320, 3, 1, 1),
                                                      (2, 7, 512, 512, 3, 1, 1)])
def test_conv2d_nhwc_fwd_bwd(B, H, Cin, Cout, K, stride, pad):
    from ml_trainer_b200.ops import conv as C

    torch.manual_seed(0)
    x = torch.randn(B, H, H, Cin, device=DEV).to(torch.bfloat16).requires_grad_()
    w = (torch.randn(Cout, Cin, K, K, device=DEV) * 0.1).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    y = C.conv2d_nhwc(x, w, stride, pad)
    xr = x.detach().float().permute(0, 3, 1, 2).requires_grad_()
    wr = w.detach().float().requires_grad_()
    ref = TF.conv2d(xr, wr, None, stride, pad)
    _close(y.permute(0, 3, 1, 2), ref, 2e-2, "conv fwd")
    dy = torch.randn_like(ref).to(torch.bfloat16)
    gx, gw = torch.autograd.grad(y, (x, w), dy.permute(0, 2, 3, 1).contiguous())
    rx, rw = torch.autograd.grad(ref, (xr, wr), dy.float())
    _close(gx.permute(0, 3, 1, 2), rx, 3e-2, "dgrad")
    _close(gw, rw, 3e-2, "wgrad")


@pytest.mark.parametrize("M_shape,C", [((4, 8, 8), 64), ((2, 28, 28), 256), ((1, 7, 7), 2048), ((3, 5, 5), 8)])
@pytest.mark.parametrize("relu,with_res", [(False, False), (True, False), (True, True)])
def test_batchnorm_nhwc(M_shape, C, relu, with_res):
    from ml_trainer_b200.ops import conv as Cv

    torch.manual_seed(0)
    B, H, W = M_shape
    x = (torch.randn(B, H, W, C, device=DEV) * 2 + 0.5).to(torch.bfloat16).requires_grad_()
    r = torch.randn(B, H, W, C, device=DEV).to(torch.bfloat16).requires_grad_() if with_res else None
    bn = torch.nn.BatchNorm2d(C).to(DEV)
    with torch.no_grad():
        bn.weight.copy_(1 + 0.1 * torch.randn(C)); bn.bias.copy_(0.1 * torch.randn(C))
    ref_bn = torch.nn.BatchNorm2d(C).to(DEV)
    ref_bn.load_state_dict(bn.state_dict())
    y = Cv.batch_norm_nhwc(x, bn, relu=relu, residual=r)
    xr = x.detach().float().permute(0, 3, 1, 2).requires_grad_()
    rr = r.detach().float().permute(0, 3, 1, 2).requires_grad_() if with_res else None
    o = ref_bn(xr)
    if with_res:
        o = o + rr
    if relu:
        o = TF.relu(o)
    _close(y.permute(0, 3, 1, 2), o, 2e-2, "bn fwd")
    _close(bn.running_mean, ref_bn.running_mean, 1e-2, "running_mean")
    _close(bn.running_var, ref_bn.running_var, 2e-2, "running_var")
    dy = torch.randn_like(o).to(torch.bfloat16)
    ins = (x, r, bn.weight, bn.bias) if with_res else (x, bn.weight, bn.bias)
    rins = (xr, rr, ref_bn.weight, ref_bn.bias) if with_res else (xr, ref_bn.weight, ref_bn.bias)
    got = torch.autograd.grad(y, ins, dy.permute(0, 2, 3, 1).contiguous())
    exp = torch.autograd.grad(o, rins, dy.float())
    for i, (a, e) in enumerate(zip(got, exp)):
        a = a.permute(0, 3, 1, 2) if a.dim() == 4 else a
        _close(a, e, 4e-2, f"bn grad {i}")


def test_pools_nhwc():
    from ml_trainer_b200.ops import conv as Cv

    torch.manual_seed(0)
    x = torch.randn(2, 17, 17, 16, device=DEV).to(torch.bfloat16).requires_grad_()
    y = Cv.max_pool_nhwc(x, 3, 2, 1)
    xr = x.detach().float().permute(0, 3, 1, 2).requires_grad_()
    ref = TF.max_pool2d(xr, 3, 2, 1)
    _close(y.permute(0, 3, 1, 2), ref, 1e-6, "maxpool")
    dy = torch.randn_like(ref).to(torch.bfloat16)
    gx, = torch.autograd.grad(y, x, dy.permute(0, 2, 3, 1).contiguous())
    rx, = torch.autograd.grad(ref, xr, dy.float())
    _close(gx.permute(0, 3, 1, 2), rx, 1e-2, "maxpool bwd")
    z = Cv.global_avg_pool_nhwc(x)
    _close(z, x.detach().float().mean((1, 2)), 1e-2, "avgpool")
    gz, = torch.autograd.grad(z, x, torch.ones_like(z))
    _close(gz, torch.full_like(gz, 1.0 / 289), 1e-2, "avgpool bwd")


def test_resnet_tiny_native_matches_fp32_reference():
    """Native bf16 NHWC path vs fp32: its gradient error must be in the same range as stock PyTorch's bf16 (cuDNN) path."""
    import copy

    from ml_trainer_b200 import build_model

    torch.manual_seed(0)
    ref = build_model("resnet_tiny")
    x = torch.randn(32, 3, 64, 64)
    y = torch.randint(0, 16, (32,))

    def run(model, native):
        m = copy.deepcopy(model).to(DEV)
        if native:
            for p in m.parameters():                      # what the Trainer does: bf16 weights, fp32 BN statistics
                p.data = p.data.to(torch.bfloat16)
        else:
            m = m.to(torch.bfloat16)                      # stock path: everything bf16 (cuDNN wants matching stats dtype)
        xin = x.to(DEV)
        out = m(xin) if native else m._forward_torch(xin.to(torch.bfloat16)).float()
        TF.cross_entropy(out, y.to(DEV)).backward()
        return out.detach().cpu(), {n: p.grad.float().cpu() for n, p in m.named_parameters()}

    lr = ref(x)
    TF.cross_entropy(lr, y).backward()
    rg = {n: p.grad for n, p in ref.named_parameters()}
    lo, ng = run(ref, True)
    lt, tg = run(ref, False)
    _close(lo, lr.detach(), 6e-2, "logits")
    rel = lambda g: sum(((g[n] - rg[n]).norm() / (rg[n].norm() + 1e-6)).item() for n in rg) / len(rg)
    e_native, e_torch = rel(ng), rel(tg)
    print(f"mean relative gradient error vs fp32: native {e_native:.3f}, torch bf16 {e_torch:.3f}")
    assert e_native < 2.0 * e_torch + 0.02, (e_native, e_torch)


def test_trainer_resnet_tiny_bf16(tmp_path):
    from ml_trainer_b200 import SyntheticDataset, Trainer, build_model
    from ml_trainer_b200.ops import _lib

    _lib.launches.reset()
    ds = SyntheticDataset((3, 32, 32), 16, length=64, seed=1)
    t = Trainer(build_model("resnet_tiny"), datasets=(ds, ds), epochs=4, batch_size=16, optimizer="sgd", lr=0.05, dtype="bf16",
                model_dir=str(tmp_path), progress=False, metric="accuracy")
    t.fit()
    assert t.train_losses[-1] < t.train_losses[0]
    names = set(_lib.launches.by_name)
    assert {"gemm_bf16_tcgen05", "im2col_nhwc", "col2im_nhwc", "bn_fwd", "bn_bwd", "maxpool_fwd", "avgpool_fwd"} <= names, names


def test_unmodified_torchvision_resnet_trains_on_the_library(tmp_path):
    """The reference's contract is "bring your own nn.Module" (reference main.py:30): an unmodified torchvision bottleneck ResNet
    handed to the Trainer is recognised, re-classed onto the library's NHWC / tcgen05 path (same parameter names) and trains like
    the library's own model with the same weights."""
    import torchvision

    from ml_trainer_b200 import SyntheticDataset, Trainer
    from ml_trainer_b200.models.resnet import ResNet
    from ml_trainer_b200.ops import _lib

    torch.manual_seed(0)
    tv = torchvision.models.ResNet(torchvision.models.resnet.Bottleneck, [1, 1, 1, 1], num_classes=16)
    ours = ResNet((1, 1, 1, 1), num_classes=16)
    ours.load_state_dict(tv.state_dict())
    data = SyntheticDataset((3, 64, 64), 16, length=16, seed=3)       # ONE step: a random-init ResNet amplifies rounding noise quickly
    res = []
    for model in (tv, ours):
        _lib.launches.reset()
        tr = Trainer(model, datasets=(data, data), epochs=1, batch_size=16, model_dir=str(tmp_path), progress=False, optimizer="sgd",
                     lr=0.01, momentum=0.0, metric="accuracy", dtype="bf16", graph=False, seed=32)
        from ml_trainer_b200 import Loader

        tr.train_loader = Loader(data, batch_size=16, shuffle=False, pin_memory=True)
        tr.fit()
        names = dict(_lib.launches.by_name)
        assert names.get("conv_fwd_implicit_gemm", 0) > 0 and names.get("bn_apply", 0) > 0, names
        res.append((tr.train_losses[0], {k: v.float().clone() for k, v in tr.model.state_dict().items()}))
    assert type(tv).__module__.endswith("models.resnet") and type(tv).__name__ == "ResNet" and getattr(tv, "_mlt_native", False)
    assert list(res[0][1]) == list(res[1][1])                      # state_dict keys are torchvision's
    assert abs(res[0][0] - res[1][0]) < 2e-2 * max(1.0, abs(res[1][0]))
    for k in res[0][1]:
        if "num_batches_tracked" in k:
            continue
        a, b = res[0][1][k], res[1][1][k]
        # same kernels, but split-K / BatchNorm reductions fold with atomics: compare in norm, not element by element
        assert float((a - b).norm()) <= 0.1 * float(b.norm()) + 1e-3, (k, float((a - b).norm()), float(b.norm()))


def test_convert_model_generic_cnn_matches_fp32_torch(tmp_path):
    """A hand-written CNN (Conv2d / BatchNorm2d / LayerNorm / Linear in logical NCHW) converted layer by layer: forward and
    parameter gradients against the same network in fp32 torch."""
    import copy

    from ml_trainer_b200.ops import _lib
    from ml_trainer_b200.ops.convert import convert_model

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.c1 = torch.nn.Conv2d(8, 64, 3, padding=1, bias=False)
            self.b1 = torch.nn.BatchNorm2d(64)
            self.c2 = torch.nn.Conv2d(64, 64, 3, stride=2, padding=1, bias=True)
            self.b2 = torch.nn.BatchNorm2d(64)
            self.ln = torch.nn.LayerNorm(64)
            self.fc = torch.nn.Linear(64, 10)

        def forward(self, x):
            x = torch.relu(self.b1(self.c1(x)))
            x = torch.relu(self.b2(self.c2(x)))
            x = x.mean((2, 3))
            return self.fc(self.ln(x))

    torch.manual_seed(1)
    ref = Net().cuda()
    nat = convert_model(copy.deepcopy(ref)).to(torch.bfloat16)
    x = torch.randn(16, 8, 16, 16, device="cuda")
    y = torch.randint(0, 10, (16,), device="cuda")
    _lib.launches.reset()
    lo = nat(x.to(torch.bfloat16))
    lr = ref(x)
    torch.nn.functional.cross_entropy(lo.float(), y).backward()
    torch.nn.functional.cross_entropy(lr, y).backward()
    names = dict(_lib.launches.by_name)
    assert any(k.startswith("conv") or k.startswith("im2col") for k in names) and any(k.startswith("bn_") for k in names), names
    assert torch.allclose(lo.float(), lr, rtol=5e-2, atol=5e-2), float((lo.float() - lr).abs().max())
    rp = dict(ref.named_parameters())
    for n, p in nat.named_parameters():
        if n == "c2.bias":
            continue            # a bias in front of a BatchNorm has a gradient of exactly zero (the mean is subtracted): only rounding noise
        g, e = p.grad.float(), rp[n].grad
        rel = float((g - e).norm() / (e.norm() + 1e-6))
        assert rel < 0.1, (n, rel)
