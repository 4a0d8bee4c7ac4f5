"""BERT definition against HuggingFace's (CPU, fp32): same parameter names, same logits, same gradients."""
import pytest
import torch


def _hf(cfg_kw):
    transformers = pytest.importorskip("transformers")
    cfg = transformers.BertConfig(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, **cfg_kw)
    return transformers.BertForSequenceClassification(cfg)


def test_parameter_names_and_count_match_hf_base():
    from ml_trainer_b200.models.bert import BertForSequenceClassification, BertConfig

    with torch.device("meta"):
        ours = BertForSequenceClassification(BertConfig())
    names = {n: tuple(p.shape) for n, p in ours.named_parameters()}
    assert sum(torch.Size(s).numel() for s in names.values()) == 109_483_778
    with torch.device("meta"):
        hf = _hf(dict(num_labels=2))
    hf_names = {n: tuple(p.shape) for n, p in hf.named_parameters()}
    assert names == hf_names


def test_tiny_bert_matches_hf_forward_and_backward():
    from ml_trainer_b200 import build_model

    torch.manual_seed(0)
    kw = dict(vocab_size=1000, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
              max_position_embeddings=128, num_labels=2)
    ours = build_model("bert_tiny")
    hf = _hf(kw)
    missing, unexpected = hf.load_state_dict(ours.state_dict(), strict=False)
    assert not unexpected and all("position_ids" in m or "token_type_ids" in m for m in missing), (missing, unexpected)
    ids = torch.randint(0, 1000, (3, 64))
    y = torch.tensor([0, 1, 1])
    lo = ours(ids)
    lh = hf(input_ids=ids).logits
    assert torch.allclose(lo, lh, atol=2e-5), (lo - lh).abs().max()
    torch.nn.functional.cross_entropy(lo, y).backward()
    torch.nn.functional.cross_entropy(lh, y).backward()
    hp = dict(hf.named_parameters())
    for n, p in ours.named_parameters():
        assert torch.allclose(p.grad, hp[n].grad, atol=2e-5, rtol=1e-3), (n, (p.grad - hp[n].grad).abs().max())


def test_trainer_runs_bert_tiny_on_cpu(tmp_path):
    from ml_trainer_b200 import SyntheticDataset, Trainer, build_model

    tr = SyntheticDataset((32,), 2, length=32, kind="tokens", vocab=1000, seed=1)
    t = Trainer(build_model("bert_tiny"), datasets=(tr, tr), epochs=2, batch_size=8, optimizer="adamw", lr=1e-3,
                model_dir=str(tmp_path), progress=False)
    t.fit()
    assert t.train_losses[1] < t.train_losses[0]


def test_attention_mask_matches_hf():
    """HuggingFace-style key-padding mask (1 = attend, 0 = padding): same logits as HF on padded batches."""
    from ml_trainer_b200 import build_model

    torch.manual_seed(1)
    kw = dict(vocab_size=1000, hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
              max_position_embeddings=128, num_labels=2)
    ours = build_model("bert_tiny").eval()
    hf = _hf(kw).eval()
    hf.load_state_dict(ours.state_dict(), strict=False)
    ids = torch.randint(0, 1000, (3, 48))
    mask = torch.ones(3, 48, dtype=torch.long)
    mask[0, 30:] = 0
    mask[2, 10:] = 0
    with torch.no_grad():
        lo = ours(ids, attention_mask=mask)
        lh = hf(input_ids=ids, attention_mask=mask).logits
        assert torch.allclose(lo, lh, atol=2e-5), (lo - lh).abs().max()
        assert not torch.equal(lo, ours(ids))                      # the mask matters


def test_attention_keep_mask_statement():
    """The host statement of the in-kernel attention dropout mask (ops/functional.py::attention_keep_mask): keep rate equals
    thr / 256, the mask is a pure function of (seed, step, call site), and changing any of the three changes it."""
    from ml_trainer_b200.ops import functional as F

    keep, scale = F.attention_keep_mask(2, 3, 128, 0.1, call_id=4, step=7, seed=1234)
    assert keep.shape == (2, 3, 128, 128) and keep.dtype == torch.bool
    thr = round(0.9 * 256)
    assert abs(scale - 256.0 / thr) < 1e-9
    assert abs(keep.float().mean().item() - thr / 256) < 0.01
    # rows and columns are not constant (a 2 x 2 block shares one hash word, its four bytes differ)
    assert 0.5 < keep[0, 0, 0].float().mean() < 1.0 and 0.5 < keep[0, 0, :, 0].float().mean() < 1.0
    again, _ = F.attention_keep_mask(2, 3, 128, 0.1, call_id=4, step=7, seed=1234)
    assert torch.equal(keep, again)
    for kw in (dict(call_id=5, step=7, seed=1234), dict(call_id=4, step=8, seed=1234), dict(call_id=4, step=7, seed=1235)):
        other, _ = F.attention_keep_mask(2, 3, 128, 0.1, **kw)
        assert not torch.equal(keep, other)
    # heads and samples draw different masks
    assert not torch.equal(keep[0, 0], keep[0, 1]) and not torch.equal(keep[0, 0], keep[1, 0])


def test_attention_probability_dropout_on_cpu():
    """CPU statement of attention with dropout on the probabilities: eval is deterministic, train differs and keeps the scale."""
    from ml_trainer_b200.ops import functional as F

    torch.manual_seed(0)
    q, k, v = (torch.randn(2, 16, 64) for _ in range(3))
    base = F.attention(q, k, v, 2)
    assert torch.allclose(F.attention(q, k, v, 2, None, 0.5, False), base)
    outs = torch.stack([F.attention(q, k, v, 2, None, 0.5, True) for _ in range(400)])
    assert not torch.allclose(outs[0], base)
    assert (outs.mean(0) - base).abs().max() < 0.25 * base.abs().max()
