"""ResNet-50 definition against torchvision's (CPU, fp32): same names, same outputs, same gradients."""
import pytest
import torch


def test_resnet50_names_and_count_match_torchvision():
    tv = pytest.importorskip("torchvision")
    from ml_trainer_b200 import build_model

    with torch.device("meta"):
        ours = build_model("resnet50")
        ref = tv.models.resnet50(weights=None)
    a = {k: tuple(v.shape) for k, v in ours.state_dict().items()}
    b = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    assert a == b
    assert sum(p.numel() for p in ours.parameters()) == 25_557_032


def test_resnet50_matches_torchvision_forward_backward():
    tv = pytest.importorskip("torchvision")
    from ml_trainer_b200 import build_model

    torch.manual_seed(0)
    ours = build_model("resnet50")
    ref = tv.models.resnet50(weights=None)
    ref.load_state_dict(ours.state_dict())
    x = torch.randn(2, 3, 64, 64)
    y = torch.tensor([3, 7])
    lo, lr = ours(x), ref(x)
    assert torch.allclose(lo, lr, atol=1e-4), (lo - lr).abs().max()
    torch.nn.functional.cross_entropy(lo, y).backward()
    torch.nn.functional.cross_entropy(lr, y).backward()
    rp = dict(ref.named_parameters())
    for n, p in ours.named_parameters():
        assert torch.allclose(p.grad, rp[n].grad, atol=1e-4, rtol=1e-3), n


def test_channels_last_arena_layout_roundtrip(tmp_path):
    """Conv weights sit physically as [Cout,KH,KW,Cin] in the arena but checkpoints keep the logical shape."""
    from ml_trainer_b200 import SyntheticDataset, Trainer, build_model, load_model

    torch.manual_seed(0)
    m = build_model("resnet_tiny")
    ref = {k: v.clone() for k, v in m.state_dict().items()}
    ds = SyntheticDataset((3, 32, 32), 16, length=8, seed=0)
    t = Trainer(m, datasets=(ds, ds), epochs=1, batch_size=4, model_dir=str(tmp_path), progress=False, lr=0.0, momentum=0.0)
    w = m.layer1[0].conv2.weight
    assert w.shape == (16, 16, 3, 3) and w.permute(0, 2, 3, 1).is_contiguous()
    assert torch.equal(w, ref["layer1.0.conv2.weight"])
    t.save_model(str(tmp_path))
    m2 = load_model(build_model("resnet_tiny"), str(tmp_path / "model.pth"))
    for k, v in m2.state_dict().items():
        if "num_batches" not in k and "running" not in k:
            assert torch.equal(v, ref[k]), k
