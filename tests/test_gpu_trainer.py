"""Trainer on a B200: native kernels, eager vs CUDA-graph vs the CPU statement of the same training run."""
import copy

import numpy as np
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


class _Twice(torch.nn.Module):
    """One Linear applied twice per forward pass (and a second, single-use one): its gradient is the SUM of two uses."""

    def __init__(self, d, out):
        super().__init__()
        self.shared = torch.nn.Linear(d, d)
        self.head = torch.nn.Linear(d, out)

    def forward(self, x):
        x = x.flatten(1)
        h = torch.tanh(self.shared(x))
        h = torch.tanh(self.shared(h))
        return self.head(h)


@pytest.mark.parametrize("d,dtype", [(48, "fp32"), (512, "bf16")])       # CUDA-core linear path / tcgen05 GEMM path
def test_parameter_used_twice_accumulates(tmp_path, d, dtype):
    """native backward kernels WRITE weight gradients into the arena, so a reused parameter must accumulate over
    its uses and report "ready" once, after the last.  Oracle: torch autograd on an fp32 copy."""
    from ml_trainer_b200 import SyntheticDataset, Trainer

    torch.manual_seed(5)
    ref = _Twice(d, 10)
    model = copy.deepcopy(ref)
    B = 2048 if dtype == "bf16" else 32
    data = SyntheticDataset((d,), 10, length=B, seed=9)
    tr = Trainer(model, datasets=(data, data), epochs=1, batch_size=B, model_dir=str(tmp_path), progress=False, optimizer="sgd",
                 lr=0.1, momentum=0.0, metric=None, graph=False, dtype=dtype, seed=32)
    x, y = next(iter(tr.train_loader))
    tr.engine.train_step(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    ref = ref.cuda()
    xr = x.cuda().float()
    torch.nn.functional.cross_entropy(ref(xr), y.cuda()).backward()
    g_ref = ref.shared.weight.grad
    i = tr.optimizer.index_of[id(tr.model.shared.weight)]
    g = tr.optimizer.grad_view(i).float()
    tol = dict(rtol=2e-3, atol=2e-5) if dtype == "fp32" else dict(rtol=6e-2, atol=2e-2 * float(g_ref.abs().max()) + 1e-6)
    assert torch.allclose(g, g_ref, **tol), float((g - g_ref).abs().max())
    # and a single use would have produced a visibly different gradient (the test can tell the two apart)
    gb = tr.optimizer.grad_view(tr.optimizer.index_of[id(tr.model.shared.bias)]).float()
    assert torch.allclose(gb, ref.shared.bias.grad, **tol)
    w_new = tr.optimizer.master_view(i)
    with torch.no_grad():
        w_exp = ref.shared.weight - 0.1 * g_ref
    assert torch.allclose(w_new, w_exp, rtol=1e-2 if dtype == "bf16" else 1e-4, atol=1e-3 if dtype == "bf16" else 1e-5)


class _Syncs(torch.nn.Module):
    """A user module that cannot be captured in a CUDA graph: it reads a value back to the host in forward."""

    def __init__(self, d, classes):
        super().__init__()
        self.a = torch.nn.Linear(d, 64)
        self.b = torch.nn.Linear(64, classes)
        self.seen = 0.0

    def forward(self, x):
        h = torch.relu(self.a(x.flatten(1)))
        self.seen = float(h.abs().max().item())          # device -> host sync: illegal while a stream is capturing
        return self.b(h)


class _RefusesCapture(_Syncs):
    """Fails only while a stream is capturing (stands for any module whose forward breaks inside a capture)."""

    def forward(self, x):
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("this module cannot be captured")
        self.seen = 1.0
        return self.b(torch.relu(self.a(x.flatten(1))))


@pytest.mark.parametrize("model_cls,precheck", [(_Syncs, True), (_Syncs, False), (_RefusesCapture, True)])
def test_uncapturable_module_falls_back_to_eager_steps(tmp_path, monkeypatch, model_cls, precheck):
    """graph='auto' must not break 'bring your own nn.Module'.  A module that syncs with the host is spotted
    during the eager warm-up steps (precheck); without that check, or with a module that fails only inside the capture, the
    failed capture ends in eager steps too - and torch's current stream and device generator are usable afterwards (a capture
    that dies in capture_end leaves both in capture mode)."""
    import ml_trainer_b200.engine as E
    from ml_trainer_b200 import SyntheticDataset, Trainer

    if not precheck:
        monkeypatch.setattr(E.StepEngine, "_eager_step_watching_for_syncs", lambda self, x, y: self._train_body(x, y))
    torch.manual_seed(3)
    data = SyntheticDataset((32,), 4, length=32 * 8, seed=2)
    tr = Trainer(model_cls(32, 4), datasets=(data, data), epochs=3, batch_size=32, model_dir=str(tmp_path), progress=False,
                 optimizer="sgd", lr=0.05, metric="accuracy", seed=1)
    assert tr.engine.use_graph
    stream_before = torch.cuda.current_stream()
    tr.fit()
    assert not tr.engine.use_graph and tr.engine.graph_replays == 0
    assert all(np.isfinite(v) for v in tr.train_losses) and tr.train_losses[-1] < tr.train_losses[0]
    assert tr.model.seen > 0
    assert torch.cuda.current_stream() == stream_before
    assert torch.isfinite(torch.randn(8, device="cuda")).all()          # the device generator is not stuck in capture mode


def test_capturable_user_module_keeps_cuda_graphs(tmp_path):
    """The sync detector of the warm-up steps must not fire on the library's own ops: a plain user MLP ends up replaying graphs."""
    from ml_trainer_b200 import SyntheticDataset, Trainer

    class Plain(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.b = torch.nn.Linear(32, 64), torch.nn.Linear(64, 4)

        def forward(self, x):
            return self.b(torch.relu(self.a(x.flatten(1))))

    torch.manual_seed(3)
    data = SyntheticDataset((32,), 4, length=32 * 8, seed=2)
    tr = Trainer(Plain(), datasets=(data, data), epochs=2, batch_size=32, model_dir=str(tmp_path), progress=False,
                 optimizer="adamw", lr=0.01, metric="accuracy", seed=1)
    tr.fit()
    assert tr.engine.use_graph and tr.engine.graph_replays > 0
    assert tr.train_losses[-1] < tr.train_losses[0]
