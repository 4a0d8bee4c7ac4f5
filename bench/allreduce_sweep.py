"""Gradient all-reduce bucket sweep (BASELINE.json config 5): fused P2P / NVLS all-reduce + optimizer update
(this repository) against ``ncclAllReduce`` (+ the separate optimizer step the reference needs), 1 KB .. 1 GB.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/allreduce_sweep.py [--max-mb 256]

Every number is device time from CUDA events, max over ranks, median of ``--iters`` launches after warm-up.
Bus bandwidth uses the all-reduce convention 2(N-1)/N * bytes / time.  Roofline per variant (bytes that must cross
NVLink per GPU / 770 GB/s measured peer bandwidth, see B200_PROFILING): one-shot (N-1)S, two-shot 2(N-1)/N S, NVLS (1+1/N) S.
"""
import argparse
import json
import os
import statistics
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def med_time(fn, iters, warm=5):
    """Device time per call: `iters` back-to-back calls captured in ONE CUDA graph (no host launch latency in the
    number), replayed 3 times; the median replay / iters, max over ranks."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(iters):
                fn()
    ts = []
    for _ in range(3):
        dist.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3 / iters)
    t = torch.tensor([statistics.median(ts)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=1024)
    ap.add_argument("--step", type=int, default=2, help="ratio between consecutive sizes")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--out", default="gpurun_out/allreduce_sweep.json")
    a = ap.parse_args()
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from ml_trainer_b200.optim import FusedOptimizer
    from ml_trainer_b200.parallel.comm import init_process_group
    from ml_trainer_b200.parallel.p2p import P2PComm

    init_process_group("p2p", dev)
    world = dist.get_world_size()
    gdt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
    esz = 2 if a.dtype == "bf16" else 4
    sizes = []
    b = 1024
    while b <= a.max_mb * 2 ** 20:
        sizes.append(b)
        b *= a.step
    nmax = sizes[-1] // esz
    comm = P2PComm(dev, oneshot_max_bytes=4 << 20)
    comm.prepare(nmax + 4096, gdt, n_states=1)
    p = torch.nn.Parameter(torch.zeros(nmax, device=dev))
    opt = FusedOptimizer([p], kind="sgd", lr=1e-3, momentum=0.9, compute_dtype=gdt, allocator=comm.alloc)
    opt.grad.normal_()
    opt.begin_step(grad_scale=1.0 / world)
    # NCCL baseline state: same dtype gradients, fp32-free stock path (bf16 params + torch SGD foreach), as the reference would run
    g_nccl = torch.randn(nmax, device=dev).to(gdt)
    p_nccl = torch.nn.Parameter(torch.zeros(nmax, device=dev, dtype=gdt))
    rows = []
    for nbytes in sizes:
        n = nbytes // esz // 128 * 128
        if n == 0:
            n = 128
        row = {"bytes": n * esz, "world": world, "dtype": a.dtype}
        for algo in ("oneshot", "twoshot", "nvls"):
            if algo == "oneshot" and n * esz > comm.oneshot_max:
                continue
            if algo == "nvls" and not comm.arena.multicast:
                continue
            comm.algo = algo
            comm._oneshot_off.clear()            # every size is its own "bucket plan": start the receive-slot layout afresh
            comm._oneshot_used = 0
            # every run starts from the same state on every rank: the sharded variants leave fp32 master / momentum current
            # only on the owning rank, which a following one-shot run (it updates everything everywhere) would inherit
            opt.master[:n].zero_(); opt.s1[:n].zero_(); opt.model[:n].zero_()
            t = med_time(lambda: comm.allreduce_update(opt, 0, n), a.iters)
            # values check: every replica holds bit-identical weights after the timed updates
            parts = [torch.empty_like(opt.model[:n]) for _ in range(world)]
            dist.all_gather(parts, opt.model[:n].contiguous())
            row[f"{algo}_replicas_identical"] = all(torch.equal(parts[0], q) for q in parts[1:])
            del parts
            row[f"{algo}_fused_us"] = round(t, 2)
            row[f"{algo}_busbw_GBs"] = round(2 * (world - 1) / world * n * esz / t / 1e3, 1)
        gs = g_nccl[:n]
        t_ar = med_time(lambda: dist.all_reduce(gs), a.iters)
        pv = torch.nn.Parameter(p_nccl.data[:n])
        pv.grad = gs
        sgd = torch.optim.SGD([pv], lr=1e-3, momentum=0.9)

        def nccl_step():
            dist.all_reduce(gs)
            gs.div_(world)
            sgd.step()
        t_full = med_time(nccl_step, a.iters)
        row["nccl_allreduce_us"] = round(t_ar, 2)
        row["nccl_busbw_GBs"] = round(2 * (world - 1) / world * n * esz / t_ar / 1e3, 1)
        row["nccl_allreduce_plus_sgd_us"] = round(t_full, 2)
        best = min(v for k, v in row.items() if k.endswith("_fused_us"))
        row["speedup_vs_nccl_plus_sgd"] = round(t_full / best, 2)
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        # crossover points: the largest size up to which each variant is the fastest fused one (what pick_algo should use)
        best = [(r["bytes"], min((r[k], k.split("_")[0]) for k in r if k.endswith("_fused_us"))[1]) for r in rows]
        cross = {"oneshot_max_bytes": max([b for b, w in best if w == "oneshot"], default=0),
                 "twoshot_max_bytes": max([b for b, w in best if w in ("oneshot", "twoshot")], default=0), "winners": best}
        with open(a.out, "w") as f:
            json.dump({"world": world, "dtype": a.dtype, "multicast": bool(comm.arena.multicast), "crossover": cross, "rows": rows}, f, indent=1)
    comm.check_errors()
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
