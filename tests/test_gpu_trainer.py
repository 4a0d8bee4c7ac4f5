"""Trainer on a B200: native kernels, eager vs CUDA-graph vs the CPU statement of the same training run."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _sets():
    from ml_trainer_b200 import SyntheticDataset

    return (SyntheticDataset((3, 32, 32), 10, length=256, seed=3), SyntheticDataset((3, 32, 32), 10, length=64, seed=4))


def _run(tmp_path, device_is_cuda, graph, optimizer="sgd", epochs=2, metric="accuracy", fused_step=False):
    import ml_trainer_b200.trainer as T
    from ml_trainer_b200 import MLModel, Trainer

    torch.manual_seed(32)
    model = MLModel()
    if not device_is_cuda:
        orig = torch.cuda.is_available
        torch.cuda.is_available = lambda: False
    try:
        tr = Trainer(model, datasets=_sets(), epochs=epochs, batch_size=32, model_dir=str(tmp_path), progress=False,
                     optimizer=optimizer, lr=0.01, metric=metric, graph=graph, seed=32, fused_step=fused_step)
        # identical batch order on every run
        from ml_trainer_b200 import Loader
        tr.train_loader = Loader(tr.train_loader.dataset, batch_size=32, shuffle=False, pin_memory=device_is_cuda)
        tr.fit()
    finally:
        if not device_is_cuda:
            torch.cuda.is_available = orig
    return tr


@pytest.mark.parametrize("optimizer", ["sgd", "adamw"])
def test_gpu_training_matches_cpu(tmp_path, optimizer):
    cpu = _run(tmp_path / "cpu", False, False, optimizer)
    gpu = _run(tmp_path / "gpu", True, False, optimizer)
    assert gpu.device.type == "cuda" and cpu.device.type == "cpu"
    for a, b in zip(cpu.train_losses, gpu.train_losses):
        assert abs(a - b) < 2e-3 * max(1.0, abs(a)), (cpu.train_losses, gpu.train_losses)
    # Adam-family updates are m/sqrt(v): where a gradient is ~0 its sign is noise, so compare those loosely
    atol = 2e-4 if optimizer == "sgd" else 0.01 * 16 * 0.25
    for (n, p), (_, q) in zip(cpu.model.named_parameters(), gpu.model.named_parameters()):
        assert torch.allclose(p, q.cpu(), rtol=2e-3, atol=atol), (n, (p - q.cpu()).abs().max())


def test_graph_replay_matches_eager(tmp_path):
    eager = _run(tmp_path / "e", True, False)
    graph = _run(tmp_path / "g", True, True)
    assert graph.engine.graph_replays > 0 and graph.engine.launches_per_step > 0
    assert eager.train_losses == pytest.approx(graph.train_losses, rel=1e-5)
    assert eager.train_metrics == pytest.approx(graph.train_metrics, rel=1e-6)
    for p, q in zip(eager.model.parameters(), graph.model.parameters()):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6)


def test_native_kernels_ran_and_loss_falls(tmp_path):
    from ml_trainer_b200.ops import _lib

    _lib.launches.reset()
    tr = _run(tmp_path, True, False, epochs=4)
    names = set(_lib.launches.by_name)
    assert {"conv2d_direct_fwd", "conv2d_direct_bwd", "linear_small_fwd", "linear_small_bwd", "softmax_xent",
            "fused_update"} <= names, names
    assert tr.train_losses[-1] < tr.train_losses[0]


def test_default_path_is_the_single_kernel_step(tmp_path):
    from ml_trainer_b200.ops import _lib

    _lib.launches.reset()
    tr = _run(tmp_path, True, "auto", epochs=3, fused_step=True)
    assert tr.engine.fused is not None
    assert set(_lib.launches.by_name) == {"lenet_step"}, _lib.launches.by_name
    assert tr.train_losses[-1] < tr.train_losses[0]


def test_checkpoint_from_gpu_is_cpu_fp32(tmp_path):
    tr = _run(tmp_path, True, True, epochs=1)
    sd = torch.load(tmp_path / "model.pth")
    assert all(v.device.type == "cpu" and v.dtype == torch.float32 for v in sd.values())
    for k, v in tr.model.state_dict().items():
        assert torch.equal(sd[k], v.cpu())
    assert next(tr.model.parameters()).is_cuda          # the live model never left the GPU


def test_smoke_entry():
    import __graft_entry__ as g

    g.smoke()
