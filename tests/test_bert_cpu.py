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
