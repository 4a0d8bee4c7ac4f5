"""Transformer kernels (LayerNorm+residual, attention via batched tcgen05 GEMMs + row softmax, embeddings) and
the native BERT path against fp32 PyTorch."""
import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _close(a, b, tol, what):
    err = (a.float() - b.float()).abs().max().item()
    ref = max(1.0, b.float().abs().max().item())
    assert err <= tol * ref, f"{what}: max abs err {err} (ref scale {ref})"


@pytest.mark.parametrize("T,H", [(64, 768), (1000, 128), (37, 1024), (4096, 768)])
@pytest.mark.parametrize("pdtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("with_res", [False, True])
def test_layernorm_fwd_bwd(T, H, pdtype, with_res):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    x = torch.randn(T, H, device=DEV).to(torch.bfloat16).requires_grad_()
    r = torch.randn(T, H, device=DEV).to(torch.bfloat16).requires_grad_() if with_res else None
    g = (1 + 0.1 * torch.randn(H, device=DEV)).to(pdtype).requires_grad_()
    b = (0.1 * torch.randn(H, device=DEV)).to(pdtype).requires_grad_()
    y = F.layer_norm(x, g, b, 1e-12, residual=r)
    xr = x.detach().float().requires_grad_()
    rr = r.detach().float().requires_grad_() if with_res else None
    gr, br = g.detach().float().requires_grad_(), b.detach().float().requires_grad_()
    h = xr + rr if with_res else xr
    if with_res:
        h = h + (h.to(torch.bfloat16).float() - h).detach()         # the kernel rounds the sum to bf16 once
    ref = TF.layer_norm(h, (H,), gr, br, 1e-12)
    _close(y, ref, 2e-2, "y")
    dy = torch.randn_like(ref).to(torch.bfloat16)
    ins = (x, r, g, b) if with_res else (x, g, b)
    rins = (xr, rr, gr, br) if with_res else (xr, gr, br)
    got = torch.autograd.grad(y, ins, dy)
    exp = torch.autograd.grad(ref, rins, dy.float())
    names = ("dx", "dres", "dgamma", "dbeta") if with_res else ("dx", "dgamma", "dbeta")
    for n, a, e in zip(names, got, exp):
        _close(a, e, 3e-2 if n.startswith("dx") or n == "dres" else 2e-2, n)


@pytest.mark.parametrize("nb1,nb2,M,N,K", [(2, 3, 128, 128, 64), (12, 2, 512, 512, 64), (4, 2, 512, 64, 512)])
def test_batched_gemm_strided_heads(nb1, nb2, M, N, K):
    from ml_trainer_b200.ops.gemm import gemm_batched

    torch.manual_seed(0)
    a = (torch.randn(nb2, nb1, M, K, device=DEV) * 0.5).to(torch.bfloat16)
    b = (torch.randn(nb2, nb1, N, K, device=DEV) * 0.5).to(torch.bfloat16)
    out = torch.empty(nb2, nb1, M, N, dtype=torch.float32, device=DEV)
    gemm_batched(a, K, M * K, nb1 * M * K, False, b, K, N * K, nb1 * N * K, False, out, N, M * N, nb1 * M * N, M, N, K, nb1, nb2)
    torch.cuda.synchronize()
    _close(out, a.float() @ b.float().transpose(-1, -2), 1e-3 * K ** 0.5, "batched")


@pytest.mark.parametrize("B,S,heads,d", [(2, 128, 2, 64), (2, 512, 12, 64), (3, 64, 4, 32)])
def test_attention_fwd_bwd(B, S, heads, d):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    q, k, v = ((torch.randn(B, S, heads * d, device=DEV) * 0.5).to(torch.bfloat16).requires_grad_() for _ in range(3))
    o = F.attention(q, k, v, heads)
    qr, kr, vr = (t.detach().float().requires_grad_() for t in (q, k, v))
    sp = lambda t: t.view(B, S, heads, d).transpose(1, 2)
    ref = TF.scaled_dot_product_attention(sp(qr), sp(kr), sp(vr)).transpose(1, 2).reshape(B, S, heads * d)
    _close(o, ref, 2e-2, "attention out")
    do = torch.randn_like(ref).to(torch.bfloat16)
    got = torch.autograd.grad(o, (q, k, v), do)
    exp = torch.autograd.grad(ref, (qr, kr, vr), do.float())
    for n, a, e in zip(("dq", "dk", "dv"), got, exp):
        _close(a, e, 4e-2, n)


def test_embeddings_fwd_bwd():
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    V, P, H, B, S = 500, 64, 128, 4, 32
    word = torch.randn(V, H, device=DEV).to(torch.bfloat16).requires_grad_()
    pos = torch.randn(P, H, device=DEV).to(torch.bfloat16).requires_grad_()
    typ = torch.randn(2, H, device=DEV).to(torch.bfloat16).requires_grad_()
    ids = torch.randint(0, V, (B, S), device=DEV)
    h = F.bert_embeddings(ids, word, pos, typ)
    wr, pr, tr = (t.detach().float().requires_grad_() for t in (word, pos, typ))
    ref = wr[ids] + pr[:S].unsqueeze(0) + tr[0]
    _close(h, ref, 1e-2, "embedding sum")
    dh = torch.randn_like(ref).to(torch.bfloat16)
    got = torch.autograd.grad(h, (word, pos, typ), dh)
    exp = torch.autograd.grad(ref, (wr, pr, tr), dh.float())
    for n, a, e in zip(("dword", "dpos", "dtype"), got, exp):
        _close(a, e, 3e-2, n)


def test_bert_tiny_native_matches_fp32_reference():
    """bf16 native forward/backward of a small BERT against the same weights run in fp32 PyTorch on the CPU."""
    from ml_trainer_b200 import build_model

    torch.manual_seed(0)
    ref = build_model("bert_tiny")
    nat = build_model("bert_tiny")
    nat.load_state_dict(ref.state_dict())
    nat = nat.to(DEV).to(torch.bfloat16)
    ids = torch.randint(0, 1000, (4, 64))
    y = torch.tensor([0, 1, 1, 0])
    lo = nat(ids.to(DEV))
    lr = ref(ids)
    _close(lo.cpu(), lr, 3e-2, "logits")
    TF.cross_entropy(lo, y.to(DEV)).backward()
    TF.cross_entropy(lr, y).backward()
    rp = dict(ref.named_parameters())
    worst = 0.0
    for n, p in nat.named_parameters():
        if n.endswith("key.bias"):
            continue            # softmax is invariant to a constant added to every key: this gradient is exactly 0 up to rounding
        g, e = p.grad.float().cpu(), rp[n].grad
        rel = (g - e).norm() / (e.norm() + 1e-6)
        worst = max(worst, rel.item())
        assert rel < 0.08, (n, rel.item())
    print("worst relative gradient error", worst)


def test_trainer_bert_tiny_bf16(tmp_path):
    from ml_trainer_b200 import SyntheticDataset, Trainer, build_model
    from ml_trainer_b200.ops import _lib

    _lib.launches.reset()
    tr = SyntheticDataset((64,), 2, length=64, kind="tokens", vocab=1000, seed=1)
    t = Trainer(build_model("bert_tiny"), datasets=(tr, tr), epochs=3, batch_size=16, optimizer="adamw", lr=1e-3,
                dtype="bf16", model_dir=str(tmp_path), progress=False, metric="accuracy")
    t.fit()
    assert t.train_losses[-1] < t.train_losses[0]
    names = set(_lib.launches.by_name)
    # head dim 64, S = 64: the fused tcgen05 attention kernels (S may be any multiple of 64 up to 512)
    assert {"gemm_bf16_tcgen05", "attention_fwd_tcgen05", "attention_bwd_tcgen05", "layernorm_fwd", "layernorm_bwd",
            "embedding_fwd", "embedding_bwd"} <= names, names
    sd = torch.load(tmp_path / "model.pth")
    assert all(v.dtype == torch.float32 for v in sd.values() if v.is_floating_point())


def test_native_graph_executor_matches_python_loop(tmp_path):
    """Generic models: C++ executor replaying per-slot CUDA graphs == the Python loop over the same batches."""
    from ml_trainer_b200 import Loader, SyntheticDataset, Trainer, build_model

    outs = []
    for native in (True, False):
        torch.manual_seed(0)
        tr = SyntheticDataset((64,), 2, length=16 * 13 + 5, kind="tokens", vocab=1000, seed=1)
        t = Trainer(build_model("bert_tiny"), datasets=(tr, tr), epochs=2, batch_size=16, optimizer="adamw", lr=1e-3,
                    dtype="bf16", model_dir=str(tmp_path / str(native)), progress=False, metric="accuracy",
                    native_loop=native, seed=0)
        t.train_loader = Loader(tr, batch_size=16, shuffle=False, pin_memory=True)
        t.val_loader = Loader(tr, batch_size=16, shuffle=False, pin_memory=True)
        assert t._native_graph_ok(t.train_loader, t.model) == native
        t.fit()
        outs.append(t)
    a, b = outs
    assert a.engine.graph_replays > 0
    assert a.train_losses == pytest.approx(b.train_losses, rel=2e-3)
    # embedding gradients accumulate with atomics (order varies run to run) and AdamW turns that noise into +-lr steps
    for p, q in zip(a.model.parameters(), b.model.parameters()):
        assert torch.allclose(p.float(), q.float(), rtol=2e-2, atol=1e-3 * 28 * 2)
    assert a.last_native_stats["h2d_bytes"] > 0


@pytest.mark.parametrize("B,S,heads", [(1, 128, 1), (2, 256, 3), (2, 512, 12), (3, 384, 2), (2, 64, 2), (2, 192, 3), (1, 448, 1)])
def test_fused_attention_forward(B, S, heads):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    d = 64
    q, k, v = ((torch.randn(B, S, heads * d, device=DEV) * 0.7).to(torch.bfloat16) for _ in range(3))
    out, lse = F.attention_fused_fwd(q, k, v, heads, want_lse=True)
    torch.cuda.synchronize()
    sp = lambda t: t.float().view(B, S, heads, d).transpose(1, 2)
    ref = TF.scaled_dot_product_attention(sp(q), sp(k), sp(v)).transpose(1, 2).reshape(B, S, heads * d)
    _close(out, ref, 2e-2, "fused attention out")
    ref_lse = torch.logsumexp(sp(q) @ sp(k).transpose(-1, -2) / d ** 0.5, dim=-1)
    _close(lse, ref_lse, 2e-3, "lse")
    with torch.no_grad():
        _close(F.attention(q, k, v, heads), ref, 2e-2, "dispatch under no_grad")


@pytest.mark.parametrize("B,S,heads", [(1, 128, 1), (2, 256, 3), (2, 512, 12), (1, 384, 2), (2, 64, 2), (2, 192, 3), (1, 320, 2)])
def test_fused_attention_backward(B, S, heads):
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(1)
    d = 64
    q, k, v = ((torch.randn(B, S, heads * d, device=DEV) * 0.7).to(torch.bfloat16).requires_grad_() for _ in range(3))
    go = (torch.randn(B, S, heads * d, device=DEV) * 0.5).to(torch.bfloat16)
    out = F.attention(q, k, v, heads)
    out.backward(go)
    torch.cuda.synchronize()
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    sp = lambda t: t.view(B, S, heads, d).transpose(1, 2)
    ref = TF.scaled_dot_product_attention(sp(qf), sp(kf), sp(vf)).transpose(1, 2).reshape(B, S, heads * d)
    ref.backward(go.float())
    _close(out, ref, 2e-2, "fused attention out")
    for name, got, want in (("dq", q.grad, qf.grad), ("dk", k.grad, kf.grad), ("dv", v.grad, vf.grad)):
        err = (got.float() - want).abs().max().item() / (want.abs().max().item() + 1e-12)
        assert err < 3e-2, f"{name}: relative max error {err}"


@pytest.mark.parametrize("B,S,heads", [(3, 128, 2), (2, 512, 12), (4, 192, 3), (2, 64, 1)])
@pytest.mark.parametrize("packed", [False, True])
def test_fused_attention_key_padding_mask(B, S, heads, packed):
    """Key-padding mask as a per-key additive bias inside the fused kernels, forward and backward, against fp32
    scaled_dot_product_attention with a boolean mask.  Padded QUERY rows are excluded from the comparison of
    the outputs (they attend uniformly to garbage in both implementations but nothing reads them)."""
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(S + heads)
    d = 64
    HD = heads * d
    lengths = torch.randint(S // 4, S + 1, (B,))
    lengths[0] = S                                           # one full-length sample
    amask = (torch.arange(S)[None, :] < lengths[:, None]).to(DEV)          # [B, S] True = token
    kb = F.prepare_key_bias(amask.long())
    if packed:
        qkv = (torch.randn(B, S, 3 * HD, device=DEV) * 0.7).to(torch.bfloat16).requires_grad_()
        out = F.attention_packed(qkv, heads, kb)
        q, k, v = qkv[..., :HD], qkv[..., HD:2 * HD], qkv[..., 2 * HD:]
    else:
        q, k, v = ((torch.randn(B, S, HD, device=DEV) * 0.7).to(torch.bfloat16).requires_grad_() for _ in range(3))
        out = F.attention(q, k, v, heads, kb)
    go = (torch.randn(B, S, HD, device=DEV) * 0.5).to(torch.bfloat16) * amask[:, :, None]      # no gradient enters at padded rows
    out.backward(go)
    torch.cuda.synchronize()
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    sp = lambda t: t.reshape(B, S, heads, d).transpose(1, 2)
    ref = TF.scaled_dot_product_attention(sp(qf), sp(kf), sp(vf), attn_mask=amask[:, None, None, :]).transpose(1, 2).reshape(B, S, HD)
    ref.backward(go.float())
    valid = amask[:, :, None].float()
    _close(out.float() * valid, ref * valid, 2e-2, "masked attention out")
    grads = (qkv.grad[..., :HD], qkv.grad[..., HD:2 * HD], qkv.grad[..., 2 * HD:]) if packed else (q.grad, k.grad, v.grad)
    for name, got, want in zip(("dq", "dk", "dv"), grads, (qf.grad, kf.grad, vf.grad)):
        err = (got.float() - want).abs().max().item() / (want.abs().max().item() + 1e-12)
        assert err < 3e-2, f"{name}: relative max error {err}"
    # padded keys receive no gradient at all
    assert float((grads[1].float() * (1 - valid)).abs().max()) < 1e-3
    assert float((grads[2].float() * (1 - valid)).abs().max()) < 1e-3


@pytest.mark.parametrize("B,S,heads", [(2, 128, 2), (1, 512, 3), (3, 192, 1), (2, 64, 2)])
@pytest.mark.parametrize("masked", [False, True])
@pytest.mark.parametrize("packed", [False, True])
def test_fused_attention_dropout(B, S, heads, masked, packed):
    """Dropout on the attention probabilities inside the fused kernels: forward and the three gradients against an fp32
    statement that applies the SAME keep mask (``attention_keep_mask`` restates the kernels' counter hash on the host), so the
    comparison is exact up to bf16 rounding - and it shows that all three kernels regenerate one and the same mask."""
    from ml_trainer_b200.ops import _lib
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(S * 3 + heads)
    d, p_drop = 64, 0.1
    HD = heads * d
    dev = torch.device(DEV)
    F.dropout_seed(0x1234ABCD5678)
    F.dropout_begin_step(dev)
    F.dropout_begin_step(dev)
    kb, amask = None, torch.ones(B, S, dtype=torch.bool, device=DEV)
    if masked:
        lengths = torch.randint(S // 4, S + 1, (B,))
        lengths[0] = S
        amask = (torch.arange(S)[None, :] < lengths[:, None]).to(DEV)
        kb = F.prepare_key_bias(amask.long())
    _lib.launches.reset()
    if packed:
        qkv = (torch.randn(B, S, 3 * HD, device=DEV) * 0.7).to(torch.bfloat16).requires_grad_()
        out = F.attention_packed(qkv, heads, kb, p_drop, True)
        q, k, v = qkv[..., :HD], qkv[..., HD:2 * HD], qkv[..., 2 * HD:]
    else:
        q, k, v = ((torch.randn(B, S, HD, device=DEV) * 0.7).to(torch.bfloat16).requires_grad_() for _ in range(3))
        out = F.attention(q, k, v, heads, kb, p_drop, True)
    call_id = F._drop["call"]
    go = (torch.randn(B, S, HD, device=DEV) * 0.5).to(torch.bfloat16) * amask[:, :, None]
    out.backward(go)
    torch.cuda.synchronize()
    names = dict(_lib.launches.by_name)
    assert names.get("attention_fwd_tcgen05") == 1 and names.get("attention_bwd_tcgen05") == 3 and "dropout" not in names, names

    keep, scale = F.attention_keep_mask(B, heads, S, p_drop, call_id, device=DEV)
    assert abs(keep.float().mean().item() - 230 / 256) < 0.01
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    sp = lambda t: t.reshape(B, S, heads, d).transpose(1, 2)
    scores = sp(qf) @ sp(kf).transpose(-1, -2) / d ** 0.5
    scores = scores.masked_fill(~amask[:, None, None, :], float("-inf"))
    ref = ((torch.softmax(scores, -1) * keep * scale) @ sp(vf)).transpose(1, 2).reshape(B, S, HD)
    ref.backward(go.float())
    valid = amask[:, :, None].float()
    _close(out.float() * valid, ref * valid, 2e-2, "attention out with dropout")
    grads = (qkv.grad[..., :HD], qkv.grad[..., HD:2 * HD], qkv.grad[..., 2 * HD:]) if packed else (q.grad, k.grad, v.grad)
    for name, got, want in zip(("dq", "dk", "dv"), grads, (qf.grad, kf.grad, vf.grad)):
        err = (got.float() - want).abs().max().item() / (want.abs().max().item() + 1e-12)
        assert err < 3e-2, f"{name}: relative max error {err}"
    # another step draws another mask; the same (step, call site) draws the same one
    with torch.no_grad():
        F._drop["call"] = call_id - 1
        again = F.attention(q, k, v, heads, kb, p_drop, True)
        F.dropout_begin_step(dev)
        F._drop["call"] = call_id - 1
        other = F.attention(q, k, v, heads, kb, p_drop, True)
    assert torch.equal(again, out.detach())
    assert not torch.equal(other, out.detach())


def test_attention_dropout_rate_is_unbiased():
    """Mean of the dropped output over many call sites approaches the undropped output (survivors are scaled by 256 / thr)."""
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(5)
    B, S, heads, HD = 2, 128, 2, 128
    q, k, v = ((torch.randn(B, S, HD, device=DEV) * 0.5).to(torch.bfloat16) for _ in range(3))
    F.dropout_begin_step(torch.device(DEV))
    base = F.attention(q, k, v, heads).float()
    acc = torch.zeros_like(base)
    n = 200
    for _ in range(n):
        acc += F.attention(q, k, v, heads, None, 0.25, True).float()
    err = (acc / n - base).abs().max().item() / base.abs().max().item()
    assert err < 0.08, err


def test_bert_masked_batch_uses_fused_attention():
    """A padded batch (HuggingFace attention_mask) stays on the fused attention kernels and matches the fp32 CPU statement."""
    from ml_trainer_b200 import build_model
    from ml_trainer_b200.ops import _lib

    torch.manual_seed(0)
    ref = build_model("bert_tiny")
    nat = build_model("bert_tiny")
    nat.load_state_dict(ref.state_dict())
    nat = nat.to(DEV).to(torch.bfloat16)
    ids = torch.randint(0, 1000, (4, 64))
    mask = torch.ones(4, 64, dtype=torch.long)
    mask[1, 40:] = 0
    mask[3, 9:] = 0
    _lib.launches.reset()
    lo = nat(ids.to(DEV), attention_mask=mask.to(DEV))
    lr = ref(ids, attention_mask=mask)
    _close(lo.cpu(), lr, 3e-2, "masked logits")
    names = dict(_lib.launches.by_name)
    assert names.get("attention_fwd_tcgen05", 0) == 2 and "softmax_rows" not in names, names


def test_packed_qkv_matches_separate_projections():
    """Q/K/V packed by the FusedOptimizer (one GEMM per direction + packed flash attention) == three separate Linears."""
    from ml_trainer_b200 import build_model
    from ml_trainer_b200.optim.fused import FusedOptimizer

    torch.manual_seed(0)
    base = build_model("bert_tiny")
    ids = torch.randint(0, 1000, (4, 128), device=DEV)
    y = torch.tensor([0, 1, 1, 0], device=DEV)
    grads, logits = [], []
    for packed in (True, False):
        m = build_model("bert_tiny")
        m.load_state_dict(base.state_dict())
        m = m.to(DEV)
        if not packed:
            for p in m.parameters():
                if hasattr(p, "_mlt_pack"):
                    del p._mlt_pack
        opt = FusedOptimizer(m.parameters(), "sgd", lr=0.1, compute_dtype=torch.bfloat16)
        att = m.bert.encoder.layer[0].attention.self
        assert (getattr(att.query.weight, "_mlt_packed", None) is not None) == packed
        if packed:
            pk = att.query.weight._mlt_packed
            assert pk.data.shape == (3 * 128, 128)
            assert pk.data[128:256].data_ptr() == att.key.weight.data_ptr()
            assert pk.grad[256:].data_ptr() == att.value.weight._mlt_grad_sink.data_ptr()
        opt.grad.zero_()
        lo = m(ids)
        TF.cross_entropy(lo, y).backward()
        torch.cuda.synchronize()
        logits.append(lo.detach().float())
        grads.append({n: opt.grad_view(p._mlt_index).float().clone() for n, p in m.named_parameters()})
    _close(logits[0], logits[1], 2e-2, "logits")
    for n in grads[0]:
        a, b = grads[0][n], grads[1][n]
        rel = (a - b).norm() / (b.norm() + 1e-6)
        assert rel < 0.05, (n, rel.item())


def test_trainer_bert_tiny_with_dropout(tmp_path):
    """Hidden + attention-probability dropout through the Trainer: trains, and graph replays draw fresh masks."""
    from ml_trainer_b200 import SyntheticDataset, Trainer, build_model

    tr = SyntheticDataset((64,), 2, length=96, kind="tokens", vocab=1000, seed=1)
    m = build_model("bert_tiny", hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1)
    assert m._mlt_uses_dropout
    t = Trainer(m, datasets=(tr, tr), epochs=4, batch_size=16, optimizer="adamw", lr=1e-3, dtype="bf16", model_dir=str(tmp_path),
                progress=False, metric="accuracy")
    t.fit()
    assert t.train_losses[-1] < t.train_losses[0]
    assert t.engine.graph_replays > 0
    # same batch twice in train mode: outputs differ because the step counter moved; in eval mode they are identical
    ids = torch.randint(0, 1000, (4, 64), device=DEV)
    from ml_trainer_b200.ops import functional as F

    t.model.train()
    with torch.no_grad():
        F.dropout_begin_step(torch.device(DEV)); a = t.model(ids)
        F.dropout_begin_step(torch.device(DEV)); b = t.model(ids)
    assert not torch.equal(a, b)
    t.model.eval()
    with torch.no_grad():
        assert torch.equal(t.model(ids), t.model(ids))
