#!/usr/bin/env python
"""Headline benchmark: training samples/sec of the reference's model/config, device-timed, max over ranks.

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference] [--model mlmodel|resnet50|bert_base]
                    [--secondary resnet50,bert_base|none] [--repeats R] [--scaling weak|strong]

Contract (ONE JSON line on rank 0; see DESIGN.md "bench.py"):
  value     whole-job samples/s from CUDA-event timing of K training steps with the inputs already staged on the
            device (kernel-only view).  The K-step region is timed R (>= 5) times; the MEDIAN is reported.
  e2e       the same metric through the public Trainer API (``Trainer.train_epoch``), every step including the H2D
            copy of that step's inputs from pinned host memory and a D2H read of the step's loss.  Also a median of
            R runs of K steps.  This is the number to compare with the reference arm.
  secondary the same record for the driver-named workloads (BASELINE.json configs 3-4): ResNet-50 bf16 batch 256 / GPU
            and BERT-base bf16 batch 32 / GPU at sequence length 512, measured in the same invocation.
  --impl reference runs the UNMODIFIED reference Trainer from baseline/_ref (stock DDP + NCCL + cuDNN/cuBLAS)
            on the same configs and prints the same line with "impl": "reference".

Primary config = the reference's own model and published setting: MLModel, batch 32 per GPU,
SGD lr 1e-3 momentum 0.9, fp32, cross-entropy.  Weak scaling (default): the reference's batch_size is GLOBAL and is
divided by the world size (src/trainer.py:63), so both arms are handed batch_size = 32 * N and every rank keeps 32
samples.  ``--scaling strong`` keeps the global batch fixed instead (the reference's literal semantics).

A watchdog prints whatever has been measured and exits 0 if the secondaries would overrun ``--budget-s``.
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
METRIC = "train samples/sec (device-timed, max over ranks)"
DATA = "synthetic (random inputs/labels of the named shape, random-init weights)"
L2_NOTE = ("every timed step consumes a batch it has not read before: our device-timed value walks a device pool larger "
           "than L2 (or flushes L2 with a 1 GiB fill before each step, not timed); e2e and the reference arm receive a "
           "fresh batch from host memory every step")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="mlmodel", choices=["mlmodel", "resnet50", "bert_base"])
    ap.add_argument("--secondary", default=None,
                    help="comma list of further models measured in the same run (default: resnet50,bert_base when "
                         "--model is mlmodel, else none); 'none' disables")
    ap.add_argument("--repeats", type=int, default=5, help="times the K-step region is timed; the median is reported")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--budget-s", dest="budget_s", type=float, default=780.0,
                    help="wall-clock budget of the whole invocation; the watchdog prints the line when it is reached")
    ap.add_argument("--no-pool", dest="no_pool", action="store_true",
                    help="single-kernel models: time one launch per step with an L2 flush before each, instead of one multi-step launch")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch size (0 = the model's default); global = batch * N")
    ap.add_argument("--metric", default=None, help="None (main.py default) or accuracy")
    ap.add_argument("--optimizer", default=None)
    ap.add_argument("--backend", default=None)
    ap.add_argument("--comm-algo", dest="comm_algo", default=None)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--child", action="store_true", help=argparse.SUPPRESS)      # a secondary model measured in its own process
    a = ap.parse_args()
    if a.secondary is None:
        a.secondary = "resnet50,bert_base" if a.model == "mlmodel" and not a.batch else "none"
    a.secondary = [m for m in a.secondary.split(",") if m and m != "none" and m != a.model]
    a.repeats = max(a.repeats, 1)
    return a


MODEL_CFG = {
    # model: per-GPU batch, compute dtype, optimizer, lr
    "mlmodel": dict(batch=32, dtype="fp32", optimizer="sgd", lr=1e-3),
    "resnet50": dict(batch=256, dtype="bf16", optimizer="sgd", lr=1e-3),
    "bert_base": dict(batch=32, dtype="bf16", optimizer="adamw", lr=2e-5),
}


def batches_for(a, model, world):
    """(global batch handed to the Trainer, per-rank batch).  weak: per-GPU work fixed; strong: global batch fixed."""
    per = a.batch or MODEL_CFG[model]["batch"]
    if a.scaling == "strong":
        return per, max(per // world, 1)
    return per * world, per


def shared_config(a, model, world, gb, per_rank):
    """The part of ``config`` both arms print identically (arm-specific notes go under ``detail``)."""
    cfg = MODEL_CFG[model]
    return {"model": model, "global_batch": gb, "per_rank_batch": per_rank,
            "seq_len": a.seq_len if model == "bert_base" else None, "parallelism": f"dp{world}",
            "optimizer": a.optimizer or cfg["optimizer"], "lr": cfg["lr"], "metric": a.metric, "dtype": cfg["dtype"],
            "scaling": a.scaling, "repeats": a.repeats, "l2": L2_NOTE}


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
    if "ml_trainer_b200" not in sys.modules:
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




def median(xs):
    return float(statistics.median(xs)) if xs else float("nan")


def measure_in_child(a, model, index, timeout_s):
    """Measure ``model`` with this same script in a fresh process per rank (its own CUDA context, allocator and process group),
    so nothing one model leaves behind - graph pools, page-locked datasets, communicators - can touch the next.  Every rank
    of a multi-GPU job spawns its own child with the same rank environment and a rendezvous port of its own."""
    env = dict(os.environ)
    rank, _, world = env_rank()
    if world > 1:
        env["MASTER_PORT"] = str(int(env.get("MASTER_PORT", "29500")) + 97 * (index + 1))
        env.pop("TORCHELASTIC_USE_AGENT_STORE", None)      # the child ranks rendezvous on their own store, not torchrun's
    cmd = [sys.executable, os.path.abspath(__file__), "--gpus", str(a.gpus), "--steps", str(a.steps), "--warmup", str(a.warmup),
           "--impl", a.impl, "--model", model, "--secondary", "none", "--repeats", str(a.repeats), "--scaling", a.scaling,
           "--seq-len", str(a.seq_len), "--budget-s", str(max(timeout_s - 20.0, 30.0)), "--child"]
    for flag, val in (("--metric", a.metric), ("--optimizer", a.optimizer), ("--backend", a.backend), ("--comm-algo", a.comm_algo)):
        if val:
            cmd += [flag, str(val)]
    for flag, on in (("--no-graph", a.no_graph), ("--no-flush", a.no_flush), ("--no-pool", a.no_pool)):
        if on:
            cmd.append(flag)
    t0 = time.time()
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout_s)
    except subprocess.TimeoutExpired:
        return {"error": f"timed out after {timeout_s:.0f} s"}
    if r.stderr:
        sys.stderr.write(r.stderr[-4000:])
    if rank != 0:
        return None
    rec = None
    for line in reversed(r.stdout.strip().splitlines()):
        if line.startswith("{"):
            try:
                rec = json.loads(line)
                break
            except ValueError:
                pass
    if rec is None:
        return {"error": f"child exited {r.returncode} without a record", "stderr_tail": r.stderr[-600:]}
    rec["wall_s"] = round(time.time() - t0, 1)
    return rec


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def measure_ours(a, model_name):
    """Build the Trainer for ``model_name`` and measure it.  Returns the record (rank 0) or None (other ranks)."""
    import torch

    rank, local_rank, world = env_rank()
    dev = torch.device("cuda", local_rank)
    from ml_trainer_b200 import Trainer, build_model
    from ml_trainer_b200.ops import _lib

    cfg = MODEL_CFG[model_name]
    gb, per = batches_for(a, model_name, world)
    K, W, R = a.steps, max(a.warmup, 3), a.repeats
    warm, val = make_datasets(model_name, gb, W + 4, a.seq_len, seed=7)
    timed, _ = make_datasets(model_name, gb, K, a.seq_len)
    model = build_model(model_name)
    extra = {"comm_algo": a.comm_algo} if a.comm_algo else {}
    tr = Trainer(model, datasets=(warm, val), epochs=1, batch_size=gb, is_parallel=world > 1,
                 optimizer=a.optimizer or cfg["optimizer"], lr=cfg["lr"], momentum=0.9, metric=a.metric,
                 dtype=cfg["dtype"], backend=a.backend or "p2p", graph=not a.no_graph, progress=False,
                 model_dir=os.path.join("/tmp", f"mlt_bench_{os.getpid()}"), seed=32, **extra)
    per_rank = tr.batch_size
    assert per_rank == per, (per_rank, per)
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
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    pool_mode = eng.fused is not None and hasattr(eng.fused, "run_pool") and not a.no_pool
    dev_ms_runs = []
    if pool_mode:
        # Single-kernel models: the K timed steps are ONE launch of the step kernel walking a device-resident pool of P distinct
        # batches (P x 393 KB = 201 MB > the 126 MB L2, every step reads a batch it has not seen); weights stay where
        # steady-state training keeps them.  L2 is flushed before every launch.
        P = 512
        xp = torch.randn(P, per_rank, 3, 32, 32, device=dev)
        yp = torch.randint(0, 10, (P, per_rank), device=dev)
        tr.optimizer.begin_step(tr.reducer.grad_scale)
        eng.fused.run_pool(xp, yp, max(W, 8), True, acc3=eng.acc, want_acc=a.metric == "accuracy")
        for r in range(R):
            if flush is not None:
                flush.fill_(r + 1)
            barrier_sync(dev)
            if r == R - 1:
                _lib.launches.reset()
            s0, e0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            eng.fused.run_pool(xp, yp, K, True, acc3=eng.acc, want_acc=a.metric == "accuracy")
            e0.record()
            barrier_sync(dev)
            dev_ms_runs.append(reduce_max(s0.elapsed_time(e0), dev))
        gpu_launches = _lib.launches.n
        del xp, yp
    else:
        replays0 = 0
        for r in range(R):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
            barrier_sync(dev)
            if r == R - 1:
                _lib.launches.reset()
                replays0 = eng.graph_replays
            for i in range(K):
                x, y = slots[i % 2]
                if flush is not None:
                    flush.fill_(i & 0xFF)                # evict L2 (1 GiB >> 126 MB); not timed
                ev[i][0].record()
                eng.train_step(x, y)
                ev[i][1].record()
            barrier_sync(dev)
            dev_ms_runs.append(reduce_max(sum(s.elapsed_time(e) for s, e in ev), dev))
        gpu_launches = _lib.launches.n + eng.launches_per_step * (eng.graph_replays - replays0 if eng.use_graph and eng.fused is None else 0)
    by_kernel = dict(_lib.launches.by_name)
    if not pool_mode and eng.use_graph and eng.fused is None:
        # the timed steps are replays of a captured step: the launches are the captured step's kernels, K times
        by_kernel = {k: int(v * K) for k, v in getattr(eng, "kernels_per_step", {}).items()} or \
            {"kernels_inside_graph_replays": int(eng.launches_per_step * K)}
    del flush
    dev_ms = median(dev_ms_runs)

    # ---- (2) end to end through the Trainer's public epoch loop ------------------------------------
    e2e = run_e2e_ours(tr, timed, K, R, dev, world, rank)
    clk = clocks.stop() if rank == 0 else None
    comm_name = tr.comm.name
    step_kind = "single fused kernel" if eng.fused is not None else ("cuda graph" if eng.use_graph else "eager")
    try:
        tr.comm.check_errors()
    finally:
        tr.close()
    del tr, eng, model, slots, feed
    import gc

    gc.collect()
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    sps = gb * K / (dev_ms / 1e3)
    return {
        "metric": METRIC, "value": round(sps, 2), "unit": "samples/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(dev_ms / K, 5), "higher_is_better": True,
        "scaling": a.scaling, "vs_baseline": round(sps / PUBLISHED[model_name], 3) if model_name in PUBLISHED else None,
        "dtype": cfg["dtype"], "data": DATA, "impl": "ours",
        "config": shared_config(a, model_name, world, gb, per_rank),
        "detail": {"step": step_kind, "comm": comm_name,
                   "value_runs_ms": [round(x, 4) for x in dev_ms_runs],
                   "l2": ("the K steps are one launch of the step kernel over a device pool of 512 distinct batches "
                          "(201 MB > 126 MB L2), L2 flushed before each launch") if pool_mode else
                         ("no flush" if a.no_flush else "1 GiB fill before every timed step, excluded from timing")},
        "e2e": e2e, "gpu_launches": int(gpu_launches), "launches_by_kernel": by_kernel,
        "clocks": clk,
    }


def run_e2e_ours(tr, timed, K, R, dev, world, rank):
    """R runs of K steps of Trainer.train_epoch: host dataset -> pinned staging -> H2D -> step -> D2H of the loss, every step."""
    import torch
    from torch.utils.data import distributed as tud

    from ml_trainer_b200 import Loader

    sampler = tud.DistributedSampler(timed, num_replicas=world, rank=rank) if world > 1 else None
    tr.train_sampler = sampler
    tr.train_loader = Loader(timed, batch_size=tr.batch_size, shuffle=sampler is None, sampler=sampler,
                             pin_memory=True, num_workers=0)
    steps = len(tr.train_loader)
    ms_runs, wall_runs = [], []
    stamps = None
    if os.environ.get("MLT_BENCH_STAMPS") == "1" and getattr(tr.engine, "fused", None) is not None:
        # analysis only: {start, end} %globaltimer stamps of every step of the LAST run (csrc/lenet_step.cu, CTA 0)
        import ctypes

        from ml_trainer_b200.ops import _lib
        stamps = torch.zeros(2 * steps, dtype=torch.int64, device=dev)
    for r in range(R):
        if stamps is not None and r == R - 1:
            _lib.lib().mlt_lenet_set_stamps(ctypes.c_void_p(stamps.data_ptr()), ctypes.c_int(steps))
        barrier_sync(dev)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        s.record()
        tr.train_epoch(2 + r)
        e.record()
        barrier_sync(dev)
        wall_runs.append(reduce_max((time.perf_counter() - t0) * 1e3, dev))
        ms_runs.append(reduce_max(max(s.elapsed_time(e), 0.0), dev))
    ms, wall_ms = median(ms_runs), median(wall_runs)
    gpu_steps = None
    if stamps is not None:
        _lib.lib().mlt_lenet_set_stamps(None, ctypes.c_int(0))
        t = stamps.view(steps, 2).cpu().double()
        if float(t.min()) > 0:
            gpu_steps = {"body_us": [round(v, 1) for v in ((t[:, 1] - t[:, 0]) / 1e3).tolist()],
                         "gap_us": [round(v, 1) for v in ((t[1:, 0] - t[:-1, 1]) / 1e3).tolist()]}
    st = getattr(tr, "last_native_stats", {}) or {}
    losses = tr.last_step_losses
    gb = tr.global_batch_size
    return {"value": round(gb * steps / (ms / 1e3), 2), "unit": "samples/s", "ms_per_step": round(ms / steps, 5),
            "wall_ms_per_step": round(wall_ms / steps, 5), "steps": steps, "runs_ms": [round(x, 4) for x in ms_runs],
            "h2d_bytes_per_step": int(st.get("h2d_bytes", 0) / max(steps, 1)),
            "d2h_bytes_per_step": int(st.get("d2h_bytes", 0) / max(steps, 1)),
            "api": "Trainer.train_epoch (what fit() runs): in-memory dataset -> pinned staging -> copy-stream H2D -> step",
            "executor": "native C++ epoch executor" if (tr._native_ok(tr.train_loader, tr.model) or
                                                        tr._native_graph_ok(tr.train_loader, tr.model)) else "python loop + PinnedPrefetcher",
            "executor_host_us": {k: round(v, 1) for k, v in (st.get("host_us") or {}).items()},
            "cpus_available": len(os.sched_getaffinity(0)),
            **({"gpu_steps_last_run": gpu_steps} if gpu_steps else {}),
            "final_loss": float(losses[steps - 1]) if losses is not None else None}


class LineKeeper:
    """Holds the record measured so far; prints it exactly once (normally at the end, or from the watchdog)."""

    def __init__(self, budget_s, is_printer):
        self.lock, self.done, self.rec, self.is_printer = threading.Lock(), False, None, is_printer
        self.timer = threading.Timer(max(budget_s, 5.0), self._expire)
        self.timer.daemon = True
        self.timer.start()
        self.t0 = time.time()

    def emit(self, note=None):
        with self.lock:
            if self.done:
                return
            self.done = True
            if self.is_printer and self.rec is not None:
                if note:
                    self.rec.setdefault("notes", []).append(note)
                print(json.dumps(self.rec), file=OUT, flush=True)

    def _expire(self):
        self.emit("watchdog: time budget reached before every secondary model finished")
        sys.stdout.flush()
        os._exit(0)

    def left(self):
        return self.timer.interval - (time.time() - self.t0)


def run_ours(a):
    import torch

    rank, local_rank, world = env_rank()
    assert world == a.gpus or (a.gpus == 1 and world == 1), f"--gpus {a.gpus} but WORLD_SIZE={world}"
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local_rank)
    sys.path.insert(0, ROOT)
    keeper = LineKeeper(a.budget_s, rank == 0)
    rec = measure_ours(a, a.model)
    keeper.rec = rec
    dev = torch.device("cuda", local_rank)
    # each secondary model runs in its own process (see measure_in_child); skip what cannot fit in the time left
    need = {"resnet50": 150.0, "bert_base": 120.0}
    for i, m in enumerate(a.secondary):
        left = -reduce_max(-keeper.left(), dev)           # the same decision on every rank
        if left < need.get(m, 120.0):
            if rec is not None:
                rec.setdefault("secondary", {})[m] = {"skipped": "time budget"}
            continue
        sub = measure_in_child(a, m, i, left - 15.0)
        if rec is not None and sub is not None:
            rec.setdefault("secondary", {})[m] = sub
    keeper.emit()
    keeper.timer.cancel()


# --------------------------------------------------------------------------------------------------
# reference arm: the unmodified reference Trainer (stock DDP/NCCL/cuDNN/cuBLAS), same metric and configs
# --------------------------------------------------------------------------------------------------
def _reference_imports():
    """Make ``import src`` resolve to the REFERENCE's package (baseline/_ref), not this repo's compatibility package."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    os.environ["TQDM_DISABLE"] = "1"          # before tqdm is imported by the reference: no progress bars in the timed region
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or ".") != ROOT]
    for m in [m for m in sys.modules if m == "src" or m.startswith("src.")]:
        del sys.modules[m]
    sys.path[:0] = [ref_dir, os.path.join(ref_dir, "_shims")]
    try:
        import structlog  # noqa: F401
    except Exception:
        pass
    import logging

    logging.disable(logging.INFO)
    from src.trainer import Trainer as RefTrainer
    from src.dataloader import Loader as RefLoader
    from src.model import MLModel as RefModel

    return RefTrainer, RefLoader, RefModel


def measure_reference(a, model_name, keeper, RefTrainer, RefLoader, RefModel):
    """Returns (record or None, seconds of budget left as agreed by all ranks)."""
    import torch

    rank, local_rank, world = env_rank()
    dev = torch.device("cuda", local_rank)
    cfg = MODEL_CFG[model_name]
    gb, per = batches_for(a, model_name, world)
    K, W, R = a.steps, max(a.warmup, 3), a.repeats
    train, val = make_datasets(model_name, gb, K, a.seq_len)
    warm, _ = make_datasets(model_name, gb, W, a.seq_len, seed=7)
    if model_name == "mlmodel":
        model = RefModel()
    else:
        if ROOT not in sys.path:
            sys.path.append(ROOT)
        from bench.ref_models import build_reference_model

        model = build_reference_model(model_name, cfg["dtype"])
    tr = RefTrainer(model, datasets=(warm, val), epochs=1, batch_size=gb, is_parallel=world > 1, save_history=False,
                    optimizer=a.optimizer or cfg["optimizer"], lr=cfg["lr"], momentum=0.9, metric=a.metric,
                    pred_function="softmax" if a.metric else None, backend="nccl", seed=32,
                    model_dir=os.path.join("/tmp", f"mlt_ref_{os.getpid()}"))
    os.makedirs(tr.model_dir, exist_ok=True)
    if cfg["dtype"] == "bf16":
        tr.model = tr.model.to(torch.bfloat16)      # bf16 legs: the reference has no AMP; pre-cast model
    tr._train_one_epoch(1)                  # warm-up through the reference's own hot loop
    # timed: K batches through the same loop (src/trainer.py:174-204); swap the public train_loader attribute
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
    ms_runs = []
    for r in range(R):
        barrier_sync(dev)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        tr._train_one_epoch(2 + r)
        e.record()
        barrier_sync(dev)
        ms_runs.append(reduce_max(s.elapsed_time(e), dev))
    ms = median(ms_runs)
    clk = clocks.stop() if rank == 0 else None
    steps = len(tr.train_loader)
    x0 = train[0][0]
    left = -reduce_max(-keeper.left(), dev)
    del tr, model
    import gc

    gc.collect()
    torch.cuda.empty_cache()
    if rank != 0:
        return None, left
    sps = gb * steps / (ms / 1e3)
    return {
        "metric": METRIC, "value": round(sps, 2), "unit": "samples/s",
        "n_gpus": world, "steps": steps, "warmup": W, "ms_per_step": round(ms / steps, 5), "higher_is_better": True,
        "scaling": a.scaling, "vs_baseline": round(sps / PUBLISHED[model_name], 3) if model_name in PUBLISHED else None,
        "dtype": cfg["dtype"], "data": DATA, "impl": "reference",
        "config": shared_config(a, model_name, world, gb, per_rank),
        "detail": {"stack": "reference Trainer._train_one_epoch (src/trainer.py:174-204), stock DDP+NCCL, torch.optim",
                   "value_runs_ms": [round(x, 4) for x in ms_runs],
                   "l2": "no flush (the reference loop cannot be interleaved); inputs arrive from host memory every step"},
        "e2e": {"value": round(sps, 2), "unit": "samples/s", "ms_per_step": round(ms / steps, 5), "steps": steps,
                "h2d_bytes_per_step": int(per_rank * (x0.numel() * x0.element_size() + 8)), "d2h_bytes_per_step": 8,
                "note": "the reference loop is end-to-end by construction (pageable H2D + loss.item() every step)"},
        "gpu_launches": 0, "clocks": clk,
    }, left


def run_reference(a):
    os.environ["TQDM_DISABLE"] = "1"          # tqdm reads it when it is first imported: no progress bars in the timed region
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "src", "trainer.py")):
        try:
            sys.path.insert(0, os.path.join(ROOT, "baseline"))
            import install_reference

            install_reference.install()
        except Exception as ex:  # noqa: BLE001
            print(json.dumps({"impl": "reference", "unavailable": f"reference not installed: {ex}"[:300]}), file=OUT, flush=True)
            return
    import torch

    rank, local_rank, world = env_rank()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "reference", "unavailable": "no CUDA device"}), file=OUT, flush=True)
        return
    sys.path.insert(0, ROOT)
    from ml_trainer_b200.data.loader import SyntheticDataset  # noqa: F401  (datasets come from this repo; import before the path swap)

    refs = _reference_imports()
    torch.cuda.set_device(local_rank)      # before Trainer: the reference sets the device after the DDP wrap
    dev = torch.device("cuda", local_rank)
    keeper = LineKeeper(a.budget_s, rank == 0)
    rec, left = measure_reference(a, a.model, keeper, *refs)
    keeper.rec = rec
    need = {"resnet50": 150.0, "bert_base": 100.0}
    for i, m in enumerate(a.secondary):
        if i > 0:
            left = -reduce_max(-keeper.left(), dev)
        if left < need.get(m, 120.0):
            if rec is not None:
                rec.setdefault("secondary", {})[m] = {"skipped": "time budget"}
            continue
        sub = measure_in_child(a, m, i, left - 15.0)
        if rec is not None and sub is not None:
            rec.setdefault("secondary", {})[m] = sub
    keeper.emit()
    keeper.timer.cancel()


OUT = sys.stdout


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL announces its version on fd 1 when a process
    group comes up), so keep a private handle on the real stdout for the record and point fd 1 at stderr for everybody else."""
    global OUT
    sys.stdout.flush()
    OUT = os.fdopen(os.dup(1), "w", buffering=1)
    os.dup2(2, 1)


if __name__ == "__main__":
    args = parse()
    claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
