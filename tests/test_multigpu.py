"""Real multi-GPU runs of the native NVLink backend (needs >= 2 GPUs; `gpurun --gpus 2 -- python -m pytest tests/test_multigpu.py`).

What is checked on real NVLink / NVSwitch (SURVEY §4 "GPU collective" tier; reference semantics src/trainer.py:152-158: SUM, then
divide by the world size):
  * the stand-alone one-shot / two-shot / NVLS all-reduce + update kernels against dist.all_reduce + the Python rule, fp32 and
    bf16 gradients, 1 KB .. 256 MB, bucket lengths that do not divide by world x 8, replicas bit-identical;
  * Trainer(world=2) through the Reducer (fused_step off, the launch counters prove which kernel ran) == stock DDP + torch.optim;
  * the single-kernel step with its in-kernel exchange == stock DDP + torch.optim;
  * bert_tiny / resnet_tiny in bf16 with several small buckets == the same Trainer over the library all-reduce;
  * checkpoints written under the sharded update kernels hold every rank's slices, and resume continues bit-identically;
  * replicas that start from different weights are equal to rank 0 after construction;
  * 10^5 back-to-back small collectives leave the flags clean;
  * a rank that dies silently stops its peer with a diagnostic within the time-out instead of training on stale data.
"""
import json
import os
import sys
import time

import pytest
import torch

from ml_trainer_b200.parallel.launch import launch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
HERE = os.path.dirname(os.path.abspath(__file__))
WORLD = int(os.environ.get("MLT_TEST_WORLD", "2"))


def run(tmp_path, *args, env=None, world=WORLD, expect=None):
    """Launch the worker on ``world`` ranks.  The rendezvous port is picked by probing, so another process can take it before
    the store binds (EADDRINUSE at start-up): when the expected result file is missing the launch is repeated once."""
    e = {"MLT_LOG_LEVEL": "WARNING"}
    e.update(env or {})
    rc = 1
    for _ in range(2):
        rc = launch([sys.executable, os.path.join(HERE, "p2p_worker.py"), str(tmp_path), *map(str, args)], world, env_extra=e)
        if expect is None or os.path.exists(os.path.join(str(tmp_path), expect)):
            break
    return rc


@pytest.mark.parametrize("optimizer,algo,fused", [("sgd", "oneshot", 0), ("adamw", "twoshot", 0), ("sgd", "nvls", 0),
                                                  ("adamw", "nvls", 0), ("sgd", "auto", 1)])
def test_ws2_p2p_matches_ddp(tmp_path, optimizer, algo, fused):
    env = {"MLT_LOG_LEVEL": "WARNING", "MLT_COMM_ALGO": algo, "MLT_FUSED_STEP": str(fused)}
    rc = launch([sys.executable, os.path.join(HERE, "dist_worker.py"), str(tmp_path), optimizer, "p2p"], 2, env_extra=env)
    assert rc == 0
    res = torch.load(tmp_path / "result.pt", weights_only=False)
    atol = 1e-5 if optimizer == "sgd" else 0.05        # Adam-family: sign noise where the gradient is ~0
    for k in res["ours"]:
        a, b = res["ours"][k], res["ddp"][k]
        assert torch.allclose(a, b, rtol=2e-3, atol=atol), f"{k}: max diff {(a - b).abs().max()}"
    h0, h1 = res["hist"]
    assert h0 == h1
    assert res["host_sum"] == pytest.approx([3.0, 5.0, 30.0])         # epoch-end reduction of host values + error check in one trip
    names = res["launches"]
    if fused:
        assert names.get("lenet_step", 0) > 0 and not any(k.startswith("p2p_allreduce_update") for k in names), names
    else:       # the Reducer path: the kernel that was asked for is the kernel that ran
        assert names.get(f"p2p_allreduce_update_{algo}", 0) > 0, names
        assert "lenet_step" not in names, names


def test_allreduce_update_kernels_vs_library(tmp_path):
    rc = run(tmp_path, "kernels", expect="kernels.json")
    rep = json.load(open(tmp_path / "kernels.json"))
    assert rc == 0 and not rep["failures"], json.dumps(rep["failures"][:4])
    algos = {r["algo"] for r in rep["report"]}
    assert {"oneshot", "twoshot"} <= algos
    assert rep["cases"] >= 40
    assert all(r["replicas_identical"] for r in rep["report"])


def test_single_kernel_step_keeps_replicas_identical(tmp_path):
    """MLModel's one-kernel step (in-kernel gradient exchange) back to back - through the C++ executor and as one multi-step
    launch, as bench.py runs it: bit-identical replicas, a loss that stays put or falls, clean flags."""
    rc = run(tmp_path, "lenet", 200, expect="lenet.json")
    rep = json.load(open(tmp_path / "lenet.json"))
    assert rc == 0 and rep["ok"], json.dumps(rep)
    r0 = rep["per_rank"][0]
    # random labels: 200 steps barely move the loss (that it falls is checked on learnable data in test_gpu_trainer.py); here it
    # must simply not run away, on top of the bit-identical replicas and clean flags above
    assert r0["loss_last"] < r0["loss_first"] + 0.05


def test_stress_small_collectives(tmp_path):
    rc = run(tmp_path, "stress", expect="stress.json")
    rep = json.load(open(tmp_path / "stress.json"))
    assert rc == 0 and not rep["bad"] and rep["error"] is None and rep["replicas_identical"], rep


@pytest.mark.parametrize("model,algo", [("bert_tiny", "twoshot"), ("bert_tiny", "nvls"), ("resnet_tiny", "nvls")])
def test_bf16_models_native_vs_library_collectives(tmp_path, model, algo):
    rc = run(tmp_path, "train", model, algo, "bf16", 0.05, expect="train.pt")
    assert rc == 0
    res = torch.load(tmp_path / "train.pt", weights_only=False)
    assert res["buckets"] >= 3, res["buckets"]
    assert res["launches"].get(f"p2p_allreduce_update_{algo}", 0) >= 3, res["launches"]
    assert res["replicas_identical"]
    a, b = res["ours"], res["ref"]
    # bf16 gradients: the library sums them in bf16, the native kernels in fp32; two epochs of AdamW at lr 1e-3
    assert torch.allclose(a, b, rtol=0.0, atol=2e-2), float((a - b).abs().max())
    assert float((a - b).abs().mean()) < 2e-3
    for x, y in zip(res["hist"][0] + res["hist"][1], res["hist_ref"][0] + res["hist_ref"][1]):
        assert abs(x - y) < 0.05 * max(1.0, abs(y))


@pytest.mark.parametrize("algo", ["twoshot", "nvls"])
def test_checkpoint_under_sharded_updates(tmp_path, algo):
    rc = run(tmp_path, "ckpt", algo, "bf16", expect="ckpt.json")
    rep = json.load(open(tmp_path / "ckpt.json"))
    assert rc == 0 and rep["ok"], json.dumps(rep)


@pytest.mark.parametrize("model", ["mlmodel", "resnet_tiny"])
def test_initial_state_follows_rank0(tmp_path, model):
    rc = run(tmp_path, "init", model, expect="init.json")
    rep = json.load(open(tmp_path / "init.json"))
    assert rc == 0 and rep["ok"], rep


@pytest.mark.parametrize("fused", [1, 0])
def test_dead_peer_stops_training_with_diagnostic(tmp_path, fused):
    t0 = time.time()
    rc = run(tmp_path, "fault", fused, env={"MLT_P2P_TIMEOUT_S": "3"})
    took = time.time() - t0
    rep = json.load(open(tmp_path / "fault_rank0.json"))
    assert rc != 0, "rank 0 must exit non-zero when its peer is gone"
    assert "timed out waiting for peer 1" in rep["message"], rep
    assert rep["weights_finite"]
    assert took < 60 and rep["seconds"] < 30, (took, rep)
