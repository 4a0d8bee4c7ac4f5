"""tcgen05 GEMM (csrc/gemm_tcgen05.cu) against fp32 torch.matmul, all operand-major combinations and tile widths."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _operands(M, N, K, a_mn, b_mn, seed=0):
    g = torch.Generator().manual_seed(seed)
    a = (torch.randn((K, M) if a_mn else (M, K), generator=g) * 0.5).to(torch.bfloat16).to(DEV)
    b = (torch.randn((K, N) if b_mn else (N, K), generator=g) * 0.5).to(torch.bfloat16).to(DEV)
    return a, b


def _check(d, ref, K, what):
    err = (d.float() - ref).abs().max().item()
    tol = 2e-2 * max(1.0, ref.abs().max().item()) if d.dtype == torch.bfloat16 else 1e-3 * (K ** 0.5)
    assert err <= tol, f"{what}: max abs err {err} > {tol}"


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("block_n", [64, 128, 256])
def test_gemm_majors_and_tiles(a_mn, b_mn, block_n):
    from ml_trainer_b200.ops.gemm import gemm, gemm_reference

    M, N, K = 256, 512, 320
    a, b = _operands(M, N, K, a_mn, b_mn)
    d = gemm(a, b, a_mn, b_mn, out_dtype=torch.float32, block_n=block_n)
    torch.cuda.synchronize()
    _check(d, gemm_reference(a, b, a_mn, b_mn), K, f"a_mn={a_mn} b_mn={b_mn} BN={block_n}")


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (100, 72, 40), (1000, 200, 136), (4096, 768, 768), (512, 3072, 768),
                                   (333, 2304, 776), (16384, 768, 3072)])
def test_gemm_shapes_with_tails(M, N, K):
    from ml_trainer_b200.ops.gemm import gemm, gemm_reference

    a, b = _operands(M, N, K, False, False, seed=1)
    d = gemm(a, b, out_dtype=torch.bfloat16)
    torch.cuda.synchronize()
    _check(d, gemm_reference(a, b), K, f"{M}x{N}x{K}")


@pytest.mark.parametrize("act", [None, "relu", "gelu", "tanh"])
@pytest.mark.parametrize("bias_dtype", [torch.bfloat16, torch.float32])
def test_gemm_epilogues(act, bias_dtype):
    from ml_trainer_b200.ops.gemm import gemm, gemm_reference

    M, N, K = 384, 256, 192
    a, b = _operands(M, N, K, False, False, seed=2)
    bias = torch.randn(N, device=DEV).to(bias_dtype)
    aux = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    d = gemm(a, b, bias=bias, act=act, aux_mode=1, aux=aux, out_dtype=torch.float32)
    torch.cuda.synchronize()
    _check(d, gemm_reference(a, b, bias=bias, act=act), K, f"act={act}")
    _check(aux, gemm_reference(a, b, bias=bias), K, "pre-activation aux")


def test_gemm_act_grad_epilogue():
    from ml_trainer_b200.ops.gemm import gemm, gemm_reference

    M, N, K = 256, 128, 128
    a, b = _operands(M, N, K, False, False, seed=3)
    z = torch.randn(M, N, device=DEV).to(torch.bfloat16)
    d = gemm(a, b, act="gelu", aux_mode=2, aux=z, out_dtype=torch.float32)
    zf = z.float().requires_grad_()
    torch.nn.functional.gelu(zf).sum().backward()
    _check(d, gemm_reference(a, b) * zf.grad, K, "gelu-grad epilogue")


@pytest.mark.parametrize("act", [None, "relu", "gelu"])
def test_linear_tc_fwd_bwd(act):
    from ml_trainer_b200.ops import functional as F
    from ml_trainer_b200.ops.gemm import linear_act_tc

    torch.manual_seed(0)
    B, K, N = 512, 256, 384
    x = torch.randn(B, K, device=DEV).to(torch.bfloat16).requires_grad_()
    w = (torch.randn(N, K, device=DEV) * 0.05).to(torch.bfloat16).requires_grad_()
    b = torch.randn(N, device=DEV).to(torch.bfloat16).requires_grad_()
    y = linear_act_tc(x, w, b, act)
    xr, wr, br = (t.detach().float().requires_grad_() for t in (x, w, b))
    ref = F._torch_act(torch.nn.functional.linear(xr, wr, br), act)
    _check(y, ref.detach(), K, "linear fwd")
    g = torch.randn_like(ref)
    gx, gw, gb = torch.autograd.grad(y, (x, w, b), g.to(torch.bfloat16))
    rx, rw, rb = torch.autograd.grad(ref, (xr, wr, br), g.to(torch.bfloat16).float())
    _check(gx, rx, N, "dX")
    _check(gw, rw, B, "dW")
    _check(gb, rb, B, "db")


def test_gemm_throughput_report():
    """Not an assertion on speed: prints achieved TFLOP/s next to torch.matmul (cuBLAS) for the BERT-base shapes."""
    from ml_trainer_b200.ops.gemm import gemm

    rows = []
    for (M, N, K) in [(16384, 768, 768), (16384, 2304, 768), (16384, 3072, 768), (16384, 768, 3072), (8192, 8192, 8192)]:
        a, b = _operands(M, N, K, False, False)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
        for fn, name in ((lambda: gemm(a, b, out=out), "tcgen05"), (lambda: torch.matmul(a, b.t(), out=out), "cublas")):
            for _ in range(3):
                fn()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(10):
                fn()
            e.record()
            torch.cuda.synchronize()
            rows.append((M, N, K, name, 2.0 * M * N * K * 10 / (s.elapsed_time(e) * 1e-3) / 1e12))
    for r in rows:
        print("GEMM %6d x %5d x %5d  %-8s %8.1f TFLOP/s" % r)


@pytest.mark.parametrize("M,N,K", [(768, 768, 16384), (768, 3072, 8192), (128, 256, 4096), (2304, 768, 1000)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_gemm_split_k_weight_gradient_shapes(M, N, K, out_dtype):
    """dW = dY^T X: both operands MN-major, few output tiles, long K -> the split-K path."""
    from ml_trainer_b200.ops.gemm import gemm, gemm_reference

    a, b = _operands(M, N, K, True, True, seed=5)
    d = gemm(a, b, True, True, out_dtype=out_dtype)
    d2 = gemm(a, b, True, True, out_dtype=out_dtype, split_k=False)
    torch.cuda.synchronize()
    ref = gemm_reference(a, b, True, True)
    _check(d, ref, K, f"split-K {M}x{N}x{K}")
    _check(d2, ref, K, f"no split {M}x{N}x{K}")


@pytest.mark.parametrize("M,N", [(16384, 768), (16384, 3072), (1000, 136), (33, 7), (4096, 2304)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_colsum(M, N, dtype):
    from ml_trainer_b200.ops import functional as F

    x = torch.randn(M, N, device=DEV).to(dtype)
    for _ in range(2):
        out = F.colsum(x)
    ref = x.float().sum(0)
    tol = 2e-2 * max(1.0, ref.abs().max().item()) if dtype == torch.bfloat16 else 1e-3 * M ** 0.5
    assert (out.float() - ref).abs().max().item() <= tol


@pytest.mark.parametrize("T,H,I,act", [(512, 128, 512, "gelu"), (1000, 256, 1024, "gelu"), (300, 64, 264, "relu")])
def test_ffn_block_matches_torch(T, H, I, act):
    """Fused MLP block: activation derivative stored by the first GEMM, applied in the second layer's dgrad epilogue."""
    from ml_trainer_b200.ops.gemm import ffn_tc

    torch.manual_seed(3)
    x = (torch.randn(T, H, device=DEV) * 0.5).to(torch.bfloat16).requires_grad_()
    w1 = (torch.randn(I, H, device=DEV) * H ** -0.5).to(torch.bfloat16).requires_grad_()
    b1 = (torch.randn(I, device=DEV) * 0.1).to(torch.bfloat16).requires_grad_()
    w2 = (torch.randn(H, I, device=DEV) * I ** -0.5).to(torch.bfloat16).requires_grad_()
    b2 = (torch.randn(H, device=DEV) * 0.1).to(torch.bfloat16).requires_grad_()
    go = torch.randn(T, H, device=DEV).to(torch.bfloat16)
    y = ffn_tc(x, w1, b1, w2, b2, act)
    y.backward(go)
    torch.cuda.synchronize()
    f32 = [t.detach().float().requires_grad_() for t in (x, w1, b1, w2, b2)]
    fn = torch.nn.functional.gelu if act == "gelu" else torch.relu
    yr = torch.nn.functional.linear(fn(torch.nn.functional.linear(f32[0], f32[1], f32[2])), f32[3], f32[4])
    yr.backward(go.float())
    assert (y.float() - yr).abs().max().item() < 3e-2 * max(1.0, yr.abs().max().item())
    for name, got, want in zip(("dx", "dw1", "db1", "dw2", "db2"), (x, w1, b1, w2, b2), f32):
        rel = (got.grad.float() - want.grad).norm() / (want.grad.norm() + 1e-6)
        assert rel < 2e-2, (name, rel.item())
    with torch.no_grad():
        y2 = ffn_tc(x, w1, b1, w2, b2, act)
    assert (y2.float() - yr).abs().max().item() < 3e-2 * max(1.0, yr.abs().max().item())


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(512, 256, 128), (1024, 768, 320), (712, 1000, 192)])
def test_gemm_cta_pair(a_mn, b_mn, M, N, K):
    """cta_group::2 kernel (256 x 256 tile on a CTA pair), forced with block_n=512."""
    from ml_trainer_b200.ops.gemm import gemm, gemm_reference

    a, b = _operands(M, N, K, a_mn, b_mn, seed=11)
    bias = torch.randn(N, device=DEV).to(torch.bfloat16)
    d = gemm(a, b, a_mn, b_mn, block_n=512, bias=bias, split_k=False)
    torch.cuda.synchronize()
    _check(d, gemm_reference(a, b, a_mn, b_mn, bias=bias), K, f"pair {M}x{N}x{K} a_mn={a_mn} b_mn={b_mn}")


@pytest.mark.parametrize("M,K,N,act", [(512, 1024, 768, None), (256, 4096, 1000 // 8 * 8, "relu"), (1024, 512, 2048, "gelu")])
def test_fp32_linear_runs_on_the_tensor_cores_via_three_term_split(M, K, N, act):
    """An fp32 nn.Linear too large for the CUDA-core kernel: forward, data and weight gradients through the bf16x3 split GEMM,
    against fp64 torch; the error must be far below what a single bf16 (or TF32) product gives."""
    from ml_trainer_b200.ops import _lib
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    x = torch.randn(M, K, device="cuda", requires_grad=True)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).requires_grad_()
    b = torch.randn(N, device="cuda", requires_grad=True)
    _lib.launches.reset()
    y = F.linear_act(x, w, b, act)
    go = torch.randn_like(y)
    y.backward(go)
    names = dict(_lib.launches.by_name)
    assert names.get("gemm_bf16_tcgen05", 0) == 3 and names.get("split3_bf16", 0) == 6, names
    xd, wd, bd = (t.detach().double().requires_grad_() for t in (x, w, b))
    z = xd @ wd.t() + bd
    # ReLU: pre-activations within rounding distance of zero may land on the other side; the oracle uses the kernel's own
    # on/off decisions there (one flipped element would change a whole row of dx by a few per cent)
    relu_like_ours = lambda t: t * (y.detach() > 0).double()
    yd = {None: lambda t: t, "relu": relu_like_ours, "gelu": torch.nn.functional.gelu}[act](z)
    yd.backward(go.double())
    for name, got, want in (("y", y, yd), ("dx", x.grad, xd.grad), ("dw", w.grad, wd.grad), ("db", b.grad, bd.grad)):
        err = float((got.double() - want).abs().max() / (want.abs().max() + 1e-12))
        assert err < 5e-5, (name, err)          # one bf16 product: ~4e-3, TF32: ~5e-4
