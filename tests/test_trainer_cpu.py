"""End-to-end Trainer behaviour on CPU (BASELINE.json config 1) and checkpoint compatibility."""
import importlib.util
import os
import pickle
import subprocess
import sys
import types

import pytest
import torch

from ml_trainer_b200 import Loader, MLModel, SyntheticDataset, Trainer, load_history, load_model

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def _fit(tmp_path, tiny_sets, **kw):
    cfg = dict(model_dir=str(tmp_path), progress=False)
    cfg.update(kw)
    t = Trainer(MLModel(), datasets=tiny_sets, epochs=2, batch_size=32, save_history=True, **cfg)
    t.fit()
    return t


def test_fit_history_schema_and_files(tmp_path, tiny_sets):
    t = _fit(tmp_path, tiny_sets)
    assert list(t.history) == ["epochs", "train_loss", "val_loss", "train_metric", "val_metric", "metric_type"]
    assert t.history["epochs"] == [1, 2] and t.history["metric_type"] == "accuracy"
    assert len(t.train_losses) == len(t.val_losses) == len(t.train_metrics) == len(t.val_metrics) == 2
    assert os.path.exists(tmp_path / "model.pth") and os.path.exists(tmp_path / "history.pkl")
    with open(tmp_path / "history.pkl", "rb") as fp:
        assert pickle.load(fp) == t.history
    assert load_history(str(tmp_path)) == t.history


def test_metric_none_returns_loss_only(tmp_path, tiny_sets):
    t = _fit(tmp_path, tiny_sets, metric=None)
    assert t.train_metrics == [] and t.history["metric_type"] is None
    out = t.test(t.model, t.val_loader)
    assert isinstance(out, float)


def test_checkpoint_is_plain_fp32_state_dict(tmp_path, tiny_sets):
    t = _fit(tmp_path, tiny_sets)
    sd = torch.load(tmp_path / "model.pth")
    ref_keys = list(MLModel().state_dict().keys())
    assert list(sd.keys()) == ref_keys
    assert all(v.dtype == torch.float32 and v.device.type == "cpu" for v in sd.values())
    m = load_model(MLModel(), str(tmp_path / "model.pth"))
    loss, acc = t.test(m, t.val_loader)
    assert abs(loss - t.val_losses[-1]) < 1e-5 and abs(acc - t.val_metrics[-1]) < 1e-6


def _reference_utils():
    """Import the reference's own utils.py (needs a matplotlib stub on this box)."""
    if not os.path.exists(os.path.join(REF, "src/utils/utils.py")):
        pytest.skip("reference tree not mounted")
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib  # noqa: F401
        except Exception:
            mpl = types.ModuleType("matplotlib")
            mpl.pyplot = types.ModuleType("matplotlib.pyplot")
            sys.modules["matplotlib"] = mpl
            sys.modules["matplotlib.pyplot"] = mpl.pyplot
    spec = importlib.util.spec_from_file_location("ref_utils", os.path.join(REF, "src/utils/utils.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("parallel_keys", [False, True])
def test_reference_load_model_accepts_our_checkpoints(tmp_path, tiny_sets, parallel_keys):
    ref_utils = _reference_utils()
    t = _fit(tmp_path, tiny_sets)
    if parallel_keys:                       # what parallel mode writes: module.-prefixed keys
        t.is_parallel = True
        t.save_model(str(tmp_path))
        assert all(k.startswith("module.") for k in torch.load(tmp_path / "model.pth"))
    spec = importlib.util.spec_from_file_location("ref_model", os.path.join(REF, "src/model.py"))
    ref_model_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_model_mod)
    m = ref_utils.load_model(ref_model_mod.MLModel(), str(tmp_path / "model.pth"))
    x = torch.randn(4, 3, 32, 32)
    with torch.no_grad():
        assert torch.allclose(m(x), t.model(x), atol=1e-5)
    assert ref_utils.load_history(str(tmp_path)) == t.history


def test_same_seed_same_init_as_reference_model():
    spec = importlib.util.spec_from_file_location("ref_model", os.path.join(REF, "src/model.py"))
    if not os.path.exists(os.path.join(REF, "src/model.py")):
        pytest.skip("reference tree not mounted")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.manual_seed(32)
    a = mod.MLModel()
    torch.manual_seed(32)
    b = MLModel()
    for (ka, va), (kb, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)
    assert sum(p.numel() for p in b.parameters()) == 62006


def test_loss_curve_matches_stock_pytorch(tmp_path):
    """Same data, seed and SGD settings through plain torch == through the Trainer (fp32, CPU)."""
    tr = SyntheticDataset((3, 32, 32), 10, length=96, seed=5)
    va = SyntheticDataset((3, 32, 32), 10, length=32, seed=6)
    torch.manual_seed(32)
    model = MLModel()
    t = Trainer(model, datasets=(tr, va), epochs=1, batch_size=32, model_dir=str(tmp_path), progress=False,
                metric=None, seed=32)
    # oracle: replay the exact batches the loader will produce
    torch.manual_seed(32)
    ref = MLModel()
    opt = torch.optim.SGD(ref.parameters(), lr=0.001, momentum=0.9)
    g = torch.Generator()
    g.manual_seed(1234)
    t.train_loader = Loader(tr, batch_size=32, shuffle=False)
    losses = []
    for x, y in Loader(tr, batch_size=32, shuffle=False):
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(ref(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    t._train_one_epoch(1)
    assert abs(t.train_losses[0] - sum(losses) / len(losses)) < 1e-5
    for a, b in zip(ref.parameters(), t.model.parameters()):
        assert torch.allclose(a, b, atol=1e-6)


def test_resume_restores_optimizer_and_epoch(tmp_path, tiny_sets):
    t = _fit(tmp_path, tiny_sets, optimizer="adam")
    t2 = Trainer(MLModel(), datasets=tiny_sets, epochs=3, batch_size=32, model_dir=str(tmp_path), progress=False,
                 optimizer="adam", resume=str(tmp_path))
    assert t2.start_epoch == 3 and t2.optimizer.step_count == t.optimizer.step_count
    assert torch.equal(t2.optimizer.master, t.optimizer.master) and torch.equal(t2.optimizer.s1, t.optimizer.s1)
    t2.fit()
    assert len(t2.train_losses) == 3


@pytest.mark.parametrize("sched", ["CosineAnnealingWarmRestarts", "StepLR", "ReduceLROnPlateau"])
def test_schedulers_drive_the_fused_optimizer(tmp_path, tiny_sets, sched):
    t = Trainer(MLModel(), datasets=tiny_sets, epochs=3, batch_size=32, model_dir=str(tmp_path), progress=False,
                scheduler=sched, lr=0.01)
    t.fit()
    lr = t.optimizer.param_groups[0]["lr"]
    if sched == "StepLR":
        assert abs(lr - 0.01 * 0.1) < 1e-9          # step_size=2, gamma=0.1, 3 epochs
    if sched == "CosineAnnealingWarmRestarts":
        assert lr < 0.01


@pytest.mark.parametrize("crit,metric", [("l1", "mcrmse"), ("l2", "mcrmse"), ("custom", None)])
def test_regression_criteria_train(tmp_path, crit, metric):
    class Reg(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.f = torch.nn.Linear(8, 3)

        def forward(self, x):
            return self.f(x)

    tr = SyntheticDataset((8,), length=64, seed=1, target_shape=(3,))
    va = SyntheticDataset((8,), length=32, seed=2, target_shape=(3,))
    t = Trainer(Reg(), datasets=(tr, va), epochs=2, batch_size=16, model_dir=str(tmp_path), progress=False,
                criterion=crit, metric=metric, lr=0.05)
    t.fit()
    assert t.train_losses[1] < t.train_losses[0]


def test_main_py_runs_locally_with_synthetic_data(tmp_path):
    env = dict(os.environ, MLT_LOG_LEVEL="WARNING")
    env.pop("RANK", None)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "main.py"), "--synthetic", "--synthetic_train", "64",
                        "--synthetic_val", "32", "--epochs", "1", "--batch_size", "16", "--model_dir", str(tmp_path),
                        "--metric", "accuracy", "--no_progress", "--local"],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert os.path.exists(tmp_path / "model.pth") and os.path.exists(tmp_path / "history.pkl")


def test_test_py_runs(tmp_path):
    env = dict(os.environ, MLT_LOG_LEVEL="WARNING", MLT_MODEL_DIR=str(tmp_path), MLT_EPOCHS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "test.py")], capture_output=True, text=True, env=env,
                       cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert os.path.exists(tmp_path / "model.pth")


def test_tracing_phases_are_noops_on_cpu(tmp_path):
    """trace=True must not change results and must not need a GPU (NVTX / CUDA events are skipped on CPU)."""
    from ml_trainer_b200 import SyntheticDataset, Trainer, build_model
    from ml_trainer_b200.utils import tracing

    ds = SyntheticDataset((3, 32, 32), 10, length=64, seed=3)
    losses = []
    for trace in (False, True):
        torch.manual_seed(11)
        t = Trainer(build_model("mlmodel"), datasets=(ds, ds), epochs=1, batch_size=16, optimizer="sgd", lr=0.01, progress=False,
                    model_dir=str(tmp_path / str(trace)), seed=5, trace=trace)
        t.fit()
        losses.append(t.train_losses[-1])
        tracing.enable(False, None)
    assert losses[0] == losses[1]
    with tracing.phase("anything"):
        pass
    assert tracing.PhaseTimer().summary() == {}


def test_epoch_indices_matches_batch_sampler():
    """The flattened epoch order handed to the native executor == what iterating the loader's batch sampler yields."""
    from ml_trainer_b200 import Loader, SyntheticDataset
    from ml_trainer_b200.runtime import epoch_indices

    ds = SyntheticDataset((3, 32, 32), 10, length=77, seed=3)
    for shuffle, drop_last in ((False, False), (True, False), (True, True)):
        g = torch.Generator().manual_seed(5)
        ld = Loader(ds, batch_size=16, shuffle=shuffle, drop_last=drop_last, generator=g)
        g.manual_seed(5)
        want = [list(b) for b in ld.batch_sampler]
        g.manual_seed(5)
        flat, sizes = epoch_indices(ld)
        assert sizes == [len(b) for b in want]
        assert flat.tolist() == [i for b in want for i in b]


def test_sampler_order_reproduces_stock_samplers():
    from torch.utils.data import RandomSampler, SequentialSampler
    from torch.utils.data.distributed import DistributedSampler

    from ml_trainer_b200 import SyntheticDataset
    from ml_trainer_b200.runtime import _sampler_order

    ds = SyntheticDataset((3, 32, 32), 10, length=53, seed=3)
    assert _sampler_order(SequentialSampler(ds)).tolist() == list(SequentialSampler(ds))
    g = torch.Generator().manual_seed(9)
    rs = RandomSampler(ds, generator=g)
    want = list(rs)
    g.manual_seed(9)
    assert _sampler_order(rs).tolist() == want
    torch.manual_seed(123)
    want = list(RandomSampler(ds))                    # no generator: seeded from the global RNG
    torch.manual_seed(123)
    assert _sampler_order(RandomSampler(ds)).tolist() == want
    for shuffle in (True, False):
        for drop_last in (True, False):
            for rank in (0, 2):
                d = DistributedSampler(ds, num_replicas=4, rank=rank, shuffle=shuffle, drop_last=drop_last, seed=5)
                d.set_epoch(3)
                assert _sampler_order(d).tolist() == list(d), (shuffle, drop_last, rank)


def test_device_augment_host_statement_matches_torch_ops():
    """uint8 pipeline: the host statement of crop / flip / normalise equals pad + slice + flip + (x/255 - mean)/std, the draw of a
    stream position is what the C library states (same Philox), and a Trainer trains from a U8ImageDataset on the CPU."""
    import ctypes

    import torch

    from ml_trainer_b200 import DeviceAugment, Loader, MLModel, Trainer, U8ImageDataset
    from ml_trainer_b200.ops import _lib
    from ml_trainer_b200.utils.functions import CIFAR_MEAN, CIFAR_STD

    aug = DeviceAugment((3, 32, 32), padding=4, seed=77)
    L = _lib.lib_or_none()
    if L is not None:
        out = (ctypes.c_int * 3)()
        for pos in (0, 1, 5, 123456, 2 ** 35 + 9):
            assert L.mlt_augment_params(ctypes.byref(aug.cfg), ctypes.c_ulonglong(pos), out) == 0
            assert tuple(out) == aug.params(pos)
    draws = [aug.params(i) for i in range(4000)]
    assert {d[0] for d in draws} == set(range(9)) and {d[1] for d in draws} == set(range(9)) and {d[2] for d in draws} == {0, 1}
    assert 0.45 < sum(d[2] for d in draws) / len(draws) < 0.55
    g = torch.Generator().manual_seed(0)
    img = torch.randint(0, 256, (32, 32, 3), dtype=torch.uint8, generator=g)
    for pos in range(6):
        top, left, flip = aug.params(pos)
        x = img.permute(2, 0, 1).float() / 255.0
        x = torch.nn.functional.pad(x, (4, 4, 4, 4))[:, top:top + 32, left:left + 32]
        if flip:
            x = x.flip(-1)
        want = (x - torch.tensor(CIFAR_MEAN).view(3, 1, 1)) / torch.tensor(CIFAR_STD).view(3, 1, 1)
        assert torch.allclose(aug.apply_cpu(img, pos), want, atol=1e-6)
    ev = aug.apply_cpu(img, None, train=False)
    assert torch.allclose(ev, (img.permute(2, 0, 1).float() / 255.0 - torch.tensor(CIFAR_MEAN).view(3, 1, 1)) / torch.tensor(CIFAR_STD).view(3, 1, 1), atol=1e-6)
    data = torch.randint(0, 256, (96, 32, 32, 3), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (96,), generator=g)
    ds = U8ImageDataset(data, labels, augment=aug)
    va = U8ImageDataset(data[:32], labels[:32], augment=aug, train=False)
    ld = Loader(ds, batch_size=32, shuffle=False)
    assert ld._fast and ld._augment is aug
    xb, yb = next(iter(ld))
    assert xb.shape == (32, 3, 32, 32) and xb.dtype == torch.float32 and torch.equal(yb, labels[:32])
    import tempfile

    h = Trainer(MLModel(), datasets=(ds, va), epochs=1, batch_size=32, progress=False, model_dir=tempfile.mkdtemp()).fit()
    assert len(h["train_loss"]) == 1 and h["train_loss"][0] == h["train_loss"][0]
