"""world_size=2 over gloo on CPU: our reducer + fused optimizer == stock DDP + torch.optim."""
import os
import sys

import pytest
import torch

from ml_trainer_b200.parallel.launch import launch
from ml_trainer_b200.parallel.reducer import plan_buckets

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("optimizer", ["sgd", "adamw"])
def test_ws2_gloo_matches_ddp(tmp_path, optimizer):
    env = {"MLT_LOG_LEVEL": "WARNING", "CUDA_VISIBLE_DEVICES": ""}
    rc = launch([sys.executable, os.path.join(HERE, "dist_worker.py"), str(tmp_path), optimizer, "gloo"], 2, env_extra=env)
    assert rc == 0
    res = torch.load(tmp_path / "result.pt", weights_only=False)
    for k in res["ours"]:
        a, b = res["ours"][k], res["ddp"][k]
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-6), f"{k}: max diff {(a - b).abs().max()}"
    # cross-rank reduced history: identical on both ranks (the reference leaves per-rank values unreduced)
    h0, h1 = res["hist"]
    assert h0 == h1 and len(h0["train_loss"]) == 2
    assert res["host_sum"] == pytest.approx([3.0, 5.0, 30.0])          # reduce_scalars_host: 1 + 2, 2.5 + 2.5, 10 + 20
    # rank 0 wrote module.-prefixed keys, exactly one history file
    sd = torch.load(tmp_path / "model.pth")
    assert all(k.startswith("module.") for k in sd)
    assert os.path.exists(tmp_path / "history.pkl")


def test_launcher_propagates_failure(tmp_path):
    script = tmp_path / "boom.py"
    script.write_text("import os, sys, time\n"
                      "if os.environ['RANK'] == '1':\n    sys.exit(3)\n"
                      "time.sleep(60)\n")
    rc = launch([sys.executable, str(script)], 2, grace_s=2.0)
    assert rc == 3


def test_bucket_plan_shapes():
    mib = 2 ** 20
    sizes = [1 * mib] * 10
    assert plan_buckets(sizes, 4 * mib, 1 * mib) == [(0, 1), (1, 5), (5, 9), (9, 10)]
    assert plan_buckets([100 * mib], 25 * mib, 1 * mib) == [(0, 1)]
    assert plan_buckets([10, 10, 10], 25 * mib) == [(0, 3)]
