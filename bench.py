#!/usr/bin/env python
"""Headline benchmark: training samples/sec of the reference's model/config, device-timed, max over ranks.

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference] [--model mlmodel|resnet50|bert_base]

Contract (one JSON line on rank 0; see DESIGN.md "bench.py"):
  value   whole-job samples/s from CUDA-event timing of K training steps with inputs already staged on the
          device (kernel-only view); L2 is flushed before every timed step and the flush is excluded.
  e2e     the same metric through the public Trainer API, every step including the H2D copy of that step's
          inputs from pinned host memory and a D2H read of the step's loss.
  --impl reference runs the UNMODIFIED reference Trainer from baseline/_ref (stock DDP + NCCL + cuDNN/cuBLAS)
          on the same config and prints the same line with "impl": "reference".

Default config = the reference's own model and published setting: MLModel, batch 32 per GPU,
SGD lr 1e-3 momentum 0.9, fp32, cross-entropy.  Weak scaling: the reference's batch_size is GLOBAL and is divided
by the world size (src/trainer.py:63), so both arms are handed batch_size = 32 * N and every rank keeps 32 samples.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PUBLISHED = {"mlmodel": 966.0}      # best published train throughput, samples/s (CPU; 01_ML_Training_local.ipynb:359)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="mlmodel", choices=["mlmodel", "resnet50", "bert_base"])
    ap.add_argument("--no-pool", dest="no_pool", action="store_true",
                    help="single-kernel models: time one launch per step with an L2 flush before each, instead of one multi-step launch")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch size (0 = the model's default); global = batch * N")
    ap.add_argument("--metric", default=None, help="None (main.py default) or accuracy")
    ap.add_argument("--optimizer", default=None)
    ap.add_argument("--backend", default=None)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--seq-len", type=int, default=512)
    return ap.parse_args()


MODEL_CFG = {
    # model: (global batch, dtype, optimizer, lr, input kind)
    "mlmodel": dict(batch=32, dtype="fp32", optimizer="sgd", lr=1e-3),
    "resnet50": dict(batch=256, dtype="bf16", optimizer="sgd", lr=1e-3),
    "bert_base": dict(batch=32, dtype="bf16", optimizer="adamw", lr=2e-5),
}


# --------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi while the timed region runs
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def reduce_max(value, device):
    import torch
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        on_cpu = "gloo" in str(dist.get_backend())          # our arm bootstraps over gloo; the reference has only nccl
        t = torch.tensor([value], dtype=torch.float64, device="cpu" if on_cpu else device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return float(value)


def barrier_sync(device):
    import torch
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
    if device.type == "cuda":
        torch.cuda.synchronize(device)


def make_datasets(model, global_batch, steps, seq_len, seed=32):
    sys.path.insert(0, ROOT)
    from ml_trainer_b200.data.loader import SyntheticDataset

    n = global_batch * steps
    if model == "bert_base":
        return (SyntheticDataset((seq_len,), 2, length=n, seed=seed, kind="tokens", pool=max(global_batch * 4, 256)),
                SyntheticDataset((seq_len,), 2, length=global_batch, seed=seed + 1, kind="tokens"))
    if model == "resnet50":
        return (SyntheticDataset((3, 224, 224), 1000, length=n, seed=seed, pool=max(global_batch * 2, 256)),
                SyntheticDataset((3, 224, 224), 1000, length=global_batch, seed=seed + 1))
    return (SyntheticDataset((3, 32, 32), 10, length=n, seed=seed, pool=min(n, 8192)),
            SyntheticDataset((3, 32, 32), 10, length=global_batch, seed=seed + 1))


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def run_ours(a):
    import torch

    rank, local_rank, world = env_rank()
    assert world == a.gpus or (a.gpus == 1 and world == 1), f"--gpus {a.gpus} but WORLD_SIZE={world}"
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    sys.path.insert(0, ROOT)
    from ml_trainer_b200 import Trainer, build_model
    from ml_trainer_b200.ops import _lib

    cfg = MODEL_CFG[a.model]
    gb = (a.batch or cfg["batch"]) * world       # weak scaling: per-GPU batch fixed, the Trainer gets the GLOBAL batch
    K, W = a.steps, max(a.warmup, 3)
    warm, val = make_datasets(a.model, gb, W + 4, a.seq_len, seed=7)
    timed, _ = make_datasets(a.model, gb, K, a.seq_len)
    model = build_model(a.model)
    tr = Trainer(model, datasets=(warm, val), epochs=1, batch_size=gb, is_parallel=world > 1,
                 optimizer=a.optimizer or cfg["optimizer"], lr=cfg["lr"], momentum=0.9, metric=a.metric,
                 dtype=cfg["dtype"], backend=a.backend or "p2p", graph=not a.no_graph, progress=False,
                 model_dir=os.path.join("/tmp", f"mlt_bench_{os.getpid()}"), seed=32)
    per_rank = tr.batch_size
    eng = tr.engine
    # warm-up through the public epoch loop: module load, graph capture, native executor set-up, comm bootstrap
    tr.train_epoch(1)
    barrier_sync(dev)

    # ---- (1) device-timed steps, inputs already on the device --------------------------------------
    tr.model.train()
    feed = iter(tr._feed(tr.train_loader))
    slots = [next(feed) for _ in range(2)]
    for i in range(4):
        eng.train_step(*slots[i % 2])
    # The flush must outlast the host's enqueue of the next step, otherwise the event pair would time host latency.
    flush = None if a.no_flush else torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    barrier_sync(dev)
    _lib.launches.reset()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    replays0 = eng.graph_replays
    pool_mode = eng.fused is not None and hasattr(eng.fused, "run_pool") and not a.no_pool
    if pool_mode:
        # Single-kernel models: the K timed steps are ONE launch of the step kernel walking a device-resident pool of P distinct
        # batches (P x 393 KB = 201 MB > the 126 MB L2, every step reads a batch it has not seen); weights stay where
        # steady-state training keeps them.  L2 is flushed once before the launch.
        P = 512
        xp = torch.randn(P, per_rank, 3, 32, 32, device=dev)
        yp = torch.randint(0, 10, (P, per_rank), device=dev)
        tr.optimizer.begin_step(tr.reducer.grad_scale)
        eng.fused.run_pool(xp, yp, max(W, 8), True, acc3=eng.acc, want_acc=a.metric == "accuracy")
        if flush is not None:
            flush.fill_(1)
        barrier_sync(dev)
        _lib.launches.reset()
        s0, e0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        eng.fused.run_pool(xp, yp, K, True, acc3=eng.acc, want_acc=a.metric == "accuracy")
        e0.record()
        barrier_sync(dev)
        dev_ms = reduce_max(s0.elapsed_time(e0), dev)
        del xp, yp
    else:
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for i in range(K):
            x, y = slots[i % 2]
            if flush is not None:
                flush.fill_(i & 0xFF)                # evict L2 (1 GiB >> 126 MB); not timed
            ev[i][0].record()
            eng.train_step(x, y)
            ev[i][1].record()
        barrier_sync(dev)
        dev_ms = reduce_max(sum(s.elapsed_time(e) for s, e in ev), dev)
    gpu_launches = _lib.launches.n + eng.launches_per_step * (eng.graph_replays - replays0 if eng.use_graph and eng.fused is None else 0)
    by_kernel = dict(_lib.launches.by_name)
    del flush

    # ---- (2) end to end through the Trainer's public epoch loop ------------------------------------
    e2e = run_e2e_ours(tr, timed, K, dev, world, rank)
    clk = clocks.stop() if rank == 0 else None
    tr.close()
    if rank != 0:
        return
    sps = gb * K / (dev_ms / 1e3)
    out = {
        "metric": "train samples/sec (device-timed, max over ranks)", "value": round(sps, 2), "unit": "samples/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(dev_ms / K, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": round(sps / PUBLISHED[a.model], 3) if a.model in PUBLISHED else None,
        "dtype": cfg["dtype"], "data": "synthetic (random inputs/labels of the named shape, random-init weights)",
        "impl": "ours",
        "config": {"model": a.model, "global_batch": gb, "per_rank_batch": per_rank,
                   "seq_len": a.seq_len if a.model == "bert_base" else None,
                   "parallelism": f"dp{world}", "optimizer": a.optimizer or cfg["optimizer"], "metric": a.metric,
                   "step": "single fused kernel" if eng.fused is not None else ("cuda graph" if eng.use_graph else "eager"),
                   "comm": tr.comm.name,
                   "l2": ("inputs larger than L2: the K steps are one launch of the step kernel over a device pool of 512 distinct batches "
                          "(201 MB > 126 MB L2), L2 flushed once before it") if pool_mode else
                         ("no flush" if a.no_flush else "1 GiB fill before every timed step, excluded from timing")},
        "e2e": e2e, "gpu_launches": int(gpu_launches), "launches_by_kernel": by_kernel,
        "clocks": clk,
    }
    print(json.dumps(out), flush=True)


def run_e2e_ours(tr, timed, K, dev, world, rank):
    """K steps of Trainer.train_epoch: host dataset -> pinned staging -> H2D -> step -> D2H of the loss, every step."""
    import torch
    from torch.utils.data import distributed as tud

    from ml_trainer_b200 import Loader

    sampler = tud.DistributedSampler(timed, num_replicas=world, rank=rank) if world > 1 else None
    tr.train_sampler = sampler
    tr.train_loader = Loader(timed, batch_size=tr.batch_size, shuffle=sampler is None, sampler=sampler,
                             pin_memory=True, num_workers=0)
    barrier_sync(dev)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    s.record()
    tr.train_epoch(2)
    e.record()
    barrier_sync(dev)
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms = reduce_max(max(s.elapsed_time(e), 0.0), dev)
    wall_ms = reduce_max(wall_ms, dev)
    steps = len(tr.train_loader)
    st = getattr(tr, "last_native_stats", {}) or {}
    losses = tr.last_step_losses
    gb = tr.global_batch_size
    return {"value": round(gb * steps / (ms / 1e3), 2), "unit": "samples/s", "ms_per_step": round(ms / steps, 5),
            "wall_ms_per_step": round(wall_ms / steps, 5), "steps": steps,
            "h2d_bytes_per_step": int(st.get("h2d_bytes", 0) / max(steps, 1)),
            "d2h_bytes_per_step": int(st.get("d2h_bytes", 0) / max(steps, 1)),
            "api": "Trainer.train_epoch (what fit() runs): in-memory dataset -> pinned staging -> copy-stream H2D -> step",
            "executor": "native C++ epoch executor" if (tr._native_ok(tr.train_loader, tr.model) or
                                                        tr._native_graph_ok(tr.train_loader, tr.model)) else "python loop + PinnedPrefetcher",
            "final_loss": float(losses[steps - 1]) if losses is not None else None}


# --------------------------------------------------------------------------------------------------
# reference arm: the unmodified reference Trainer (stock DDP/NCCL/cuDNN/cuBLAS), same metric and config
# --------------------------------------------------------------------------------------------------
def run_reference(a):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "src", "trainer.py")):
        try:
            sys.path.insert(0, os.path.join(ROOT, "baseline"))
            import install_reference

            install_reference.install()
        except Exception as ex:  # noqa: BLE001
            print(json.dumps({"impl": "reference", "unavailable": f"reference not installed: {ex}"[:300]}))
            return
    import torch

    rank, local_rank, world = env_rank()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "reference", "unavailable": "no CUDA device"}))
        return
    cfg = MODEL_CFG[a.model]
    gb = (a.batch or cfg["batch"]) * world       # weak scaling: per-GPU batch fixed, the Trainer gets the GLOBAL batch
    K, W = a.steps, max(a.warmup, 3)
    train, val = make_datasets(a.model, gb, K, a.seq_len)
    warm, _ = make_datasets(a.model, gb, W, a.seq_len, seed=7)
    # import the REFERENCE's src package, not this repo's compatibility package of the same name
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or ".") != ROOT]
    for m in [m for m in sys.modules if m == "src" or m.startswith("src.")]:
        del sys.modules[m]
    sys.path[:0] = [ref_dir, os.path.join(ref_dir, "_shims")]
    try:
        import structlog  # noqa: F401
    except Exception:
        pass
    from src.trainer import Trainer as RefTrainer

    torch.cuda.set_device(local_rank)      # before Trainer: the reference sets the device after the DDP wrap
    dev = torch.device("cuda", local_rank)
    if a.model == "mlmodel":
        from src.model import MLModel as RefModel

        model = RefModel()
    else:
        sys.path.append(ROOT)
        from bench.ref_models import build_reference_model

        model = build_reference_model(a.model, cfg["dtype"])
    import logging

    logging.disable(logging.INFO)
    os.environ.setdefault("TQDM_DISABLE", "1")
    tr = RefTrainer(model, datasets=(warm, val), epochs=1, batch_size=gb, is_parallel=world > 1, save_history=False,
                    optimizer=a.optimizer or cfg["optimizer"], lr=cfg["lr"], momentum=0.9, metric=a.metric,
                    pred_function="softmax" if a.metric else None, backend="nccl", seed=32,
                    model_dir=os.path.join("/tmp", f"mlt_ref_{os.getpid()}"))
    os.makedirs(tr.model_dir, exist_ok=True)
    if cfg["dtype"] == "bf16":
        _wrap_bf16_inputs(tr)
    tr._train_one_epoch(1)                  # warm-up through the reference's own hot loop
    # timed: K batches through the same loop (src/trainer.py:174-204); swap the public train_loader attribute
    from src.dataloader import Loader as RefLoader

    sampler = None
    per_rank = gb
    if world > 1:
        import torch.distributed as dist
        from torch.utils.data import distributed as tud

        sampler = tud.DistributedSampler(train, num_replicas=world, rank=dist.get_rank())
        per_rank = max(gb // world, 1)
    tr.train_loader = RefLoader(train, batch_size=per_rank, shuffle=sampler is None, sampler=sampler)
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    barrier_sync(dev)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    tr._train_one_epoch(2)
    e.record()
    barrier_sync(dev)
    ms = reduce_max(s.elapsed_time(e), dev)
    clk = clocks.stop() if rank == 0 else None
    if rank != 0:
        return
    steps = len(tr.train_loader)
    sps = gb * steps / (ms / 1e3)
    x0 = train[0][0]
    out = {
        "metric": "train samples/sec (device-timed, max over ranks)", "value": round(sps, 2), "unit": "samples/s",
        "n_gpus": world, "steps": steps, "warmup": W, "ms_per_step": round(ms / steps, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": round(sps / PUBLISHED[a.model], 3) if a.model in PUBLISHED else None,
        "dtype": cfg["dtype"], "data": "synthetic (random inputs/labels of the named shape, random-init weights)",
        "impl": "reference",
        "config": {"model": a.model, "global_batch": gb, "per_rank_batch": per_rank, "parallelism": f"dp{world}",
                   "optimizer": a.optimizer or cfg["optimizer"], "metric": a.metric,
                   "stack": "reference Trainer._train_one_epoch (src/trainer.py:174-204), stock DDP+NCCL, torch.optim",
                   "l2": "no flush (the reference loop cannot be interleaved); inputs arrive from host memory every step"},
        "e2e": {"value": round(sps, 2), "unit": "samples/s",
                "h2d_bytes_per_step": int(per_rank * (x0.numel() * x0.element_size() + 8)), "d2h_bytes_per_step": 8,
                "note": "the reference loop is end-to-end by construction (pageable H2D + loss.item() every step)"},
        "gpu_launches": 0, "clocks": clk,
    }
    print(json.dumps(out), flush=True)


def _wrap_bf16_inputs(tr):
    """bf16 legs: the reference has no AMP; feed it a pre-cast model and bf16 inputs."""
    import torch

    tr.model = tr.model.to(torch.bfloat16)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
