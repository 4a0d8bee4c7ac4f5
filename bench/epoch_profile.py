"""Where does a short native epoch spend its host time?  K-step epochs of MLModel through Trainer.train_epoch with
wall-clock timers around the Python stages and the C++ executor's own timestamps.  Usage: python bench/epoch_profile.py [K] [R]"""
import json
import sys
import time

import torch

sys.path.insert(0, ".")
import ml_trainer_b200.runtime as rt
from ml_trainer_b200 import Loader, SyntheticDataset, Trainer, build_model

import os  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 20
R = int(sys.argv[2]) if len(sys.argv) > 2 else 7
WORLD, RANK = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
if WORLD > 1:        # under torchrun: one process per GPU, 32 samples per rank (weak scaling, as bench.py)
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
warm = SyntheticDataset((3, 32, 32), 10, length=32 * 24 * WORLD, seed=7, pool=8192)
timed = SyntheticDataset((3, 32, 32), 10, length=32 * K * WORLD, seed=32, pool=min(32 * K * WORLD, 8192))
tr = Trainer(build_model("mlmodel"), datasets=(warm, warm), epochs=1, batch_size=32 * WORLD, optimizer="sgd", lr=1e-3, momentum=0.9,
             metric=None, progress=False, model_dir=f"/tmp/ep{RANK}", seed=32, is_parallel=WORLD > 1, backend="p2p")
tr.train_epoch(1)
if WORLD > 1:
    from torch.utils.data.distributed import DistributedSampler

    tr.train_sampler = DistributedSampler(timed, num_replicas=WORLD, rank=RANK)
    tr.train_loader = Loader(timed, batch_size=32, sampler=tr.train_sampler, pin_memory=True, num_workers=0)
else:
    tr.train_loader = Loader(timed, batch_size=32, shuffle=True, pin_memory=True, num_workers=0)
stages = {}


def timed_fn(name, fn):
    def w(*a, **k):
        t = time.perf_counter()
        r = fn(*a, **k)
        stages.setdefault(name, []).append((time.perf_counter() - t) * 1e6)
        return r
    return w


rt.epoch_indices = timed_fn("epoch_indices", rt.epoch_indices)
rt.NativeLoop.run_lenet = timed_fn("run_lenet", rt.NativeLoop.run_lenet)
tr.optimizer.hyper_rows = timed_fn("hyper_rows", tr.optimizer.hyper_rows)
tr._reduce_epoch = timed_fn("reduce_epoch", tr._reduce_epoch)
tr.comm.reduce_scalars = timed_fn("reduce_scalars", tr.comm.reduce_scalars)
tr.comm.reduce_scalars_host = timed_fn("reduce_scalars_host", tr.comm.reduce_scalars_host)
tr.comm.check_errors = timed_fn("check_errors", tr.comm.check_errors)
tr.train_epoch = timed_fn("train_epoch", tr.train_epoch)
tr._run_native_epoch = timed_fn("run_native_epoch", tr._run_native_epoch)
import ctypes  # noqa: E402

from ml_trainer_b200.ops import _lib  # noqa: E402

stamps = torch.zeros(2 * K * (R + 1), dtype=torch.int64, device="cuda")
LIB = _lib.lib()
rows = []
for ep in range(2, 2 + R):
    stamps.zero_()
    LIB.mlt_lenet_set_stamps(ctypes.c_void_p(stamps.data_ptr()), ctypes.c_int(K))
    torch.cuda.synchronize()
    if WORLD > 1:
        tr.comm.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t = time.perf_counter()
    s.record()
    tr.train_epoch(ep)
    e.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t) * 1e6
    st = stamps[:2 * K].view(K, 2).cpu().double()
    gpu = {}
    if float(st[:, 0].min()) > 0:
        dur = (st[:, 1] - st[:, 0]) / 1e3
        gap = (st[1:, 0] - st[:-1, 1]) / 1e3
        gpu = {"step_us_mean": round(dur.mean().item(), 2), "step_us_max": round(dur.max().item(), 2),
               "gap_us_mean": round(gap.mean().item(), 2), "gap_us_max": round(gap.max().item(), 2),
               "first_start_to_last_end_us": round(((st[-1, 1] - st[0, 0]) / 1e3).item(), 1),
               "gaps_us": [round(g, 1) for g in gap.tolist()]}
    rows.append({"wall_us": wall, "event_us": s.elapsed_time(e) * 1e3,
                 "native": tr.last_native_stats.get("host_us"), "gpu_steps": gpu})
LIB.mlt_lenet_set_stamps(None, ctypes.c_int(0))
out = {"K": K, "world": WORLD, "rank": RANK, "epochs": rows, "stages_us": {k: [round(x, 1) for x in v] for k, v in stages.items()}}
if RANK == 0:
    print(json.dumps(out, indent=1))
if WORLD > 1:
    tr.close()
