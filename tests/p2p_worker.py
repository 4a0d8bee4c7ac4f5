"""Rank program of the real multi-GPU tests (launched by ml_trainer_b200.parallel.launch, one process per GPU).

    python p2p_worker.py <out_dir> <command> [args...]

commands
  kernels      fused all-reduce + update kernels (one-shot / two-shot / NVLS; fp32 and bf16 gradients; sizes 1 KB .. 256 MB,
               bucket lengths that do not divide by world x 8) against dist.all_reduce + the Python update rule
  stress       10^5 back-to-back scalar reductions, broadcasts and one-shot all-reduces; flags must stay clean
  train        Trainer(world=N, native collectives through the Reducer) == Trainer(world=N, NCCL all-reduce + the same rule)
  ckpt         sharded update kernels: model.pth / trainer_state.pth hold every rank's slices; resume round trip
  init         ranks start from DIFFERENT weights; after Trainer.__init__ every replica equals rank 0
  fault        rank 1 dies silently in the middle of an epoch; rank 0 must stop with the time-out diagnostic
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ml_trainer_b200 import SyntheticDataset, Trainer, build_model  # noqa: E402
from ml_trainer_b200.ops import _lib  # noqa: E402
from ml_trainer_b200.optim import rules  # noqa: E402
from ml_trainer_b200.optim.fused import FusedOptimizer  # noqa: E402
from ml_trainer_b200.parallel.comm import init_process_group  # noqa: E402


def setup():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    init_process_group("p2p", dev)
    return rank, dist.get_world_size(), dev


def gather_equal(t: torch.Tensor) -> bool:
    """True when every rank holds bit-identical ``t``."""
    world = dist.get_world_size()
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t.contiguous())
    return all(torch.equal(parts[0], p) for p in parts[1:])


# ------------------------------------------------------------------------------------------------ kernels
def cmd_kernels(out_dir):
    from ml_trainer_b200.parallel.p2p import P2PComm

    rank, world, dev = setup()
    sizes = [256, 128 * 3, 128 * 1001, (1 << 20) + 128, (1 << 24) + 128 * 7, 1 << 26]   # 1 KB .. 256 MB of fp32 gradient
    quick = os.environ.get("MLT_TEST_QUICK", "0") == "1"
    if quick:
        sizes = sizes[:4]
    report, failures = [], []
    for gdt in (torch.float32, torch.bfloat16):
        for kind in ("sgd", "adamw"):
            comm = P2PComm(dev, oneshot_max_bytes=1 << 22, algo="auto")
            nmax = max(sizes)
            comm.prepare(nmax + 128, gdt, n_states=rules.RULE_STATES[kind])
            for n in sizes:
                for algo in ("oneshot", "twoshot", "nvls"):
                    if algo == "oneshot" and n * (2 if gdt == torch.bfloat16 else 4) > comm.oneshot_max:
                        continue
                    if algo == "nvls" and not comm.arena.multicast:
                        continue
                    heap0 = comm._heap
                    p = torch.nn.Parameter(torch.empty(n, device=dev))
                    g0 = torch.Generator(device=dev).manual_seed(1234)            # same initial weights everywhere
                    with torch.no_grad():
                        p.copy_(torch.randn(n, generator=g0, device=dev))
                    opt = FusedOptimizer([p], kind=kind, lr=0.05, momentum=0.9, weight_decay=0.01, compute_dtype=gdt,
                                         allocator=comm.alloc)
                    opt.s1.zero_()                     # the heap is recycled between cases; a fresh arena is zero-filled
                    if opt.s2.numel() > 1:
                        opt.s2.zero_()
                    w0 = opt.master.clone()
                    gr = torch.Generator(device=dev).manual_seed(99 + rank)       # different gradients per rank
                    grads = []
                    exp_p, exp_s1, exp_s2 = w0.clone(), torch.zeros_like(w0), torch.zeros_like(w0)
                    comm.algo = algo
                    comm._oneshot_off.clear()
                    comm._oneshot_used = 0
                    for step in range(3):
                        g = torch.randn(opt.numel, generator=gr, device=dev).to(gdt)
                        opt.grad.copy_(g)
                        grads.append(g)
                        h = opt.begin_step(1.0 / world)
                        torch.cuda.synchronize()
                        dist.barrier()
                        comm.allreduce_update(opt, 0, opt.numel)
                        torch.cuda.synchronize()
                        # oracle: library all-reduce of the up-cast gradients, then the Python statement of the rule
                        gs = g.float().clone()
                        dist.all_reduce(gs, op=dist.ReduceOp.SUM)
                        if algo == "nvls" and gdt == torch.bfloat16:
                            gs = gs.to(torch.bfloat16).float()       # multimem.ld_reduce accumulates in fp32 and RETURNS bf16
                        rules.apply_rule(kind, exp_p, gs, exp_s1, exp_s2 if rules.RULE_STATES[kind] > 1 else exp_s2[:1], h)
                    comm.check_errors()
                    # sharded variants keep master/state only on the owner: pull the slices, then compare everywhere
                    class _B:  # noqa: D401  (a one-bucket plan)
                        lo, hi = 0, opt.numel
                    comm.consolidate(opt, [_B])
                    tol = dict(rtol=2e-5, atol=2e-6) if kind == "sgd" else dict(rtol=1e-4, atol=1e-4)
                    if kind != "sgd" and world > 2:
                        # more than two addends: the kernels' fixed rank order and the library's tree differ in the last bit of
                        # the sum, and Adam turns a sign flip of a ~0 gradient into a full +-lr step: compare in the mean
                        tol = dict(rtol=1e-3, atol=3 * 2 * 0.05)
                    if algo == "nvls" and gdt == torch.bfloat16:         # the switch may round the bf16 result differently from RNE
                        tol = dict(rtol=2e-3, atol=3 * 0.05 * 8 * 2.0 ** -8)
                    mean_tol = 1e-3 if (algo == "nvls" and gdt == torch.bfloat16) else 1e-4
                    ok_master = torch.allclose(opt.master, exp_p, **tol) and float((opt.master - exp_p).abs().mean()) < mean_tol
                    # the compute-dtype weights are exactly the rounded master weights (checked after the slices were gathered)
                    ok_model = torch.equal(opt.model.float(), opt.master.to(gdt).float())
                    tol_s = dict(tol)
                    if algo == "nvls" and gdt == torch.bfloat16:
                        tol_s = dict(rtol=2e-2, atol=0.08)               # the momentum buffer accumulates the raw (bf16-rounded) sums
                    ok_s1 = torch.allclose(opt.s1, exp_s1, **tol_s)
                    same = gather_equal(opt.model) and gather_equal(opt.master) and gather_equal(opt.s1)
                    rec = dict(dtype=str(gdt), rule=kind, n=n, algo=algo, master=ok_master, model=ok_model, s1=ok_s1, replicas_identical=same,
                               max_err=float((opt.master - exp_p).abs().max()))
                    report.append(rec)
                    if not (ok_master and ok_model and ok_s1 and same):
                        failures.append(rec)
                    del opt, p, w0, exp_p, exp_s1, exp_s2, grads
                    comm._heap = heap0                 # recycle the arena heap for the next case (set-up is collective and symmetric)
            comm.close()
    if rank == 0:
        json.dump({"cases": len(report), "failures": failures, "launches": dict(_lib.launches.by_name), "report": report},
                  open(os.path.join(out_dir, "kernels.json"), "w"), indent=1)
    dist.barrier()
    return 1 if failures else 0


# ------------------------------------------------------------------------------------------------ stress
def cmd_stress(out_dir):
    from ml_trainer_b200.parallel.p2p import P2PComm

    rank, world, dev = setup()
    iters = int(os.environ.get("MLT_STRESS_ITERS", "100000"))
    comm = P2PComm(dev, algo="oneshot")
    n = 4096
    comm.prepare(n + 128, torch.float32, n_states=1)
    p = torch.nn.Parameter(torch.zeros(n, device=dev))
    opt = FusedOptimizer([p], kind="sgd", lr=1.0, momentum=0.0, compute_dtype=torch.float32, allocator=comm.alloc)
    buf = comm.alloc(1024, torch.float32, dev)
    bad = []
    # scalar reductions: every rank contributes rank+1 (+ the iteration), the sum is known in closed form
    t0 = time.time()
    vals = torch.tensor([float(rank + 1), 0.0, 1.0], device=dev)
    for i in range(iters):
        vals[1] = float(i % 97)
        out = comm.reduce_scalars(vals)
        if i % 1000 == 999 or i < 3:
            got = out.tolist()
            exp = [world * (world + 1) / 2, world * float(i % 97), float(world)]
            if got != exp:
                bad.append(("scalars", i, got, exp))
    t_scalars = time.time() - t0
    # broadcasts from a rotating root
    t0 = time.time()
    for i in range(iters // 10):
        root = i % world
        buf.fill_(float(rank * 1000 + i))
        comm.broadcast(buf, src=root)
        if i % 500 == 499 or i < 3:
            v = buf.unique().tolist()
            if v != [float(root * 1000 + i)]:
                bad.append(("broadcast", i, v[:4], root * 1000 + i))
    t_bcast = time.time() - t0
    # one-shot all-reduce + update: lr 1, no momentum, grad = 1 on every rank -> the weights drop by exactly 1 per call
    t0 = time.time()
    opt.grad.fill_(1.0)
    opt.begin_step(1.0 / world)
    k = iters // 10
    for i in range(k):
        comm.allreduce_update(opt, 0, opt.numel)
    torch.cuda.synchronize()
    t_ar = time.time() - t0
    if not bool((opt.master == -float(k)).all()):
        bad.append(("oneshot", k, float(opt.master.min()), float(opt.master.max())))
    same = gather_equal(opt.master)
    err = None
    try:
        comm.check_errors()
    except RuntimeError as ex:
        err = str(ex)
    if rank == 0:
        json.dump({"iters": iters, "bad": bad[:10], "replicas_identical": same, "error": err,
                   "seconds": {"scalars": t_scalars, "broadcast": t_bcast, "oneshot": t_ar}},
                  open(os.path.join(out_dir, "stress.json"), "w"), indent=1)
    dist.barrier()
    comm.close()
    return 1 if (bad or err or not same) else 0


# ------------------------------------------------------------------------------------------------ train
def _make_trainer(model_name, backend, algo, dtype, bucket_mb, out_dir, fused_step=False, epochs=2, optimizer="adamw", seed_model=32,
                  **model_kw):
    torch.manual_seed(seed_model)
    model = build_model(model_name, **model_kw)
    if model_name.startswith("bert"):
        tr = SyntheticDataset((64,), 2, length=64, seed=11, kind="tokens", vocab=1000)       # bert_tiny's vocabulary
        va = SyntheticDataset((64,), 2, length=32, seed=12, kind="tokens", vocab=1000)
        bs = 16
    elif model_name.startswith("resnet"):
        tr = SyntheticDataset((3, 64, 64), 16, length=64, seed=11)
        va = SyntheticDataset((3, 64, 64), 16, length=32, seed=12)
        bs = 16
    else:
        tr = SyntheticDataset((3, 32, 32), 10, length=128, seed=11)
        va = SyntheticDataset((3, 32, 32), 10, length=64, seed=12)
        bs = 32
    return Trainer(model, datasets=(tr, va), epochs=epochs, batch_size=bs, is_parallel=True, save_history=False,
                   model_dir=out_dir, backend=backend, optimizer=optimizer, lr=1e-3, progress=False, metric="accuracy",
                   comm_algo=algo, dtype=dtype, bucket_mb=bucket_mb, fused_step=fused_step, graph=False, seed=32)


def cmd_train(out_dir, model_name, algo, dtype, bucket_mb):
    """native collectives (through the Reducer, several buckets) vs the library all-reduce + the same update rule."""
    rank, world, dev = setup()
    _lib.launches.reset()
    t = _make_trainer(model_name, "p2p", algo, dtype, float(bucket_mb), out_dir)
    nb = len(t.reducer.buckets)
    t.fit()
    by_name = dict(_lib.launches.by_name)
    t.consolidate()
    ours = t.optimizer.master.detach().float().cpu().clone()
    ours_model = t.optimizer.model.detach().float().cpu().clone()
    same = gather_equal(t.optimizer.model) and gather_equal(t.optimizer.master)
    hist = (list(t.train_losses), list(t.val_losses))
    t.close()
    t2 = _make_trainer(model_name, "nccl", "auto", dtype, float(bucket_mb), out_dir)
    t2.fit()
    ref = t2.optimizer.master.detach().float().cpu().clone()
    hist2 = (list(t2.train_losses), list(t2.val_losses))
    t2.close()
    if rank == 0:
        torch.save({"ours": ours, "ours_model": ours_model, "ref": ref, "launches": by_name, "buckets": nb, "replicas_identical": same,
                    "hist": hist, "hist_ref": hist2}, os.path.join(out_dir, "train.pt"))
    dist.barrier()
    return 0


# ------------------------------------------------------------------------------------------------ ckpt
def cmd_ckpt(out_dir, algo, dtype):
    from ml_trainer_b200.utils.utils import strip_module_prefix

    rank, world, dev = setup()
    nodrop = dict(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)      # masks are a function of a per-device step counter
    t = _make_trainer("bert_tiny", "p2p", algo, dtype, 0.05, out_dir, epochs=1, **nodrop)
    sharded = sum(1 for bk in t.reducer.buckets if t.comm.owned_slices(t.optimizer, bk.lo, bk.hi))
    t.fit()                                           # saves model.pth + trainer_state.pth on rank 0 after consolidating
    dist.barrier()
    live = {k: v.detach().clone() for k, v in t.model.state_dict().items()}
    saved = strip_module_prefix(torch.load(os.path.join(out_dir, "model.pth"), map_location=dev))
    bad = []
    for name, p in t.model.named_parameters():
        if not torch.equal(saved[name].to(p.dtype), live[name]):
            bad.append(name)
    # resume: a second trainer built from the files must hold the same master / state, and train on identically
    t2 = _make_trainer("bert_tiny", "p2p", algo, dtype, 0.05, out_dir, epochs=2, **nodrop)
    t2.load(out_dir)
    t.consolidate()
    same_master = torch.equal(t.optimizer.master, t2.optimizer.master)
    same_s1 = torch.equal(t.optimizer.s1, t2.optimizer.s1)
    same_s2 = torch.equal(t.optimizer.s2, t2.optimizer.s2)
    t._train_one_epoch(2)
    t2._train_one_epoch(2)
    torch.cuda.synchronize()
    same_after = torch.equal(t.optimizer.model, t2.optimizer.model)
    max_after = float((t.optimizer.model.float() - t2.optimizer.model.float()).abs().max())
    res = dict(sharded_buckets=sharded, buckets=len(t.reducer.buckets), bad=bad[:8], same_master=same_master, same_s1=same_s1,
               same_s2=same_s2, same_after=same_after, max_after=max_after, start_epoch=t2.start_epoch)
    # the continued and the resumed run see the same batches; embedding gradients fold with atomics, so "same" is a tolerance
    ok = sharded > 0 and not bad and same_master and same_s1 and same_s2 and (same_after or max_after < 3e-2)
    oks = [None] * world
    dist.all_gather_object(oks, (ok, res))
    if rank == 0:
        json.dump({"ok": all(o for o, _ in oks), "per_rank": [r for _, r in oks]}, open(os.path.join(out_dir, "ckpt.json"), "w"), indent=1)
    dist.barrier()
    t.close()
    t2.close()
    return 0 if all(o for o, _ in oks) else 1


# ------------------------------------------------------------------------------------------------ init
def cmd_init(out_dir, model_name):
    rank, world, dev = setup()
    t = _make_trainer(model_name, "p2p", "auto", "fp32" if model_name == "mlmodel" else "bf16", 25.0, out_dir, fused_step=True,
                      seed_model=1000 + rank)        # every rank builds DIFFERENT initial weights (and BN buffers stay default)
    torch.manual_seed(1000)                           # what rank 0 built
    ref = build_model(model_name)
    exp = torch.cat([p.detach().float().flatten() for p in ref.parameters()]).sum().item()
    got = sum(t.optimizer.master_view(i).float().sum().item() for i in range(len(t.optimizer.params_in_order)))
    same = gather_equal(t.optimizer.master) and gather_equal(t.optimizer.model)
    for b in t.model.buffers():
        same = same and gather_equal(b)
    ok = same and abs(got - exp) < 1e-3 * max(1.0, abs(exp))
    oks = [None] * world
    dist.all_gather_object(oks, (ok, same, got, exp))
    if rank == 0:
        json.dump({"ok": all(o[0] for o in oks), "per_rank": oks}, open(os.path.join(out_dir, "init.json"), "w"), indent=1)
    dist.barrier()
    t.close()
    return 0 if all(o[0] for o in oks) else 1


# ------------------------------------------------------------------------------------------------ fault
def cmd_fault(out_dir, fused):
    """Rank 1 leaves silently (exit code 0, so no launcher steps in) in the middle of an epoch.  Rank 0's next collective
    waits MLT_P2P_TIMEOUT_S, stops without touching the weights, and the Trainer raises with the diagnostic."""
    rank, world, dev = setup()
    t0 = time.time()
    t = _make_trainer("mlmodel", "p2p", "auto", "fp32", 25.0, out_dir, fused_step=bool(int(fused)), epochs=1, optimizer="sgd")
    t.cfg.native_loop = False                         # Python loop: the hook below runs between steps
    steps = {"n": 0}
    orig = t.engine.train_step

    def step(x, y):
        steps["n"] += 1
        if rank == 1 and steps["n"] == 3:
            torch.cuda.synchronize()
            os._exit(0)
        return orig(x, y)

    t.engine.train_step = step
    before = None
    try:
        t._train_one_epoch(1)
        msg, code = "no error raised", 0
    except RuntimeError as ex:
        msg, code = str(ex), 3
    torch.cuda.synchronize()
    with open(os.path.join(out_dir, f"fault_rank{rank}.json"), "w") as f:
        json.dump({"message": msg, "seconds": time.time() - t0, "steps": steps["n"],
                   "weights_finite": bool(torch.isfinite(t.optimizer.master).all())}, f)
    os._exit(code)                                    # no clean shutdown: the peer is gone


# ------------------------------------------------------------------------------------------------ lenet
def cmd_lenet(out_dir, steps="200"):
    """The single-kernel MLModel step with its in-kernel exchange, back to back through the native executor and as one
    multi-step launch: replicas must stay bit-identical and the loss must not run away."""
    rank, world, dev = setup()
    n = int(steps)
    tr = SyntheticDataset((3, 32, 32), 10, length=32 * world * n, seed=11, pool=4096)
    va = SyntheticDataset((3, 32, 32), 10, length=64 * world, seed=12)
    t = Trainer(build_model("mlmodel"), datasets=(tr, va), epochs=1, batch_size=32 * world, is_parallel=True, model_dir=out_dir,
                backend="p2p", optimizer="sgd", lr=0.01, progress=False, metric=None, seed=32)
    assert t.engine.fused is not None
    t.train_epoch(1)
    torch.cuda.synchronize()
    losses = t.last_step_losses.clone()
    same_exec = gather_equal(t.optimizer.master)
    # one multi-step launch over a device pool (what bench.py times)
    xp = torch.randn(64, 32, 3, 32, 32, device=dev)
    yp = torch.randint(0, 10, (64, 32), device=dev)
    t.optimizer.begin_step(t.reducer.grad_scale)
    t.engine.fused.run_pool(xp, yp, n, True)
    torch.cuda.synchronize()
    same_pool = gather_equal(t.optimizer.master)
    err = None
    try:
        t.comm.check_errors()
    except RuntimeError as ex:
        err = str(ex)
    finite = bool(torch.isfinite(t.optimizer.master).all())
    res = dict(same_after_executor=same_exec, same_after_pool=same_pool, finite=finite, error=err,
               loss_first=float(losses[:10].mean()), loss_last=float(losses[-10:].mean()), steps=n)
    ok = same_exec and same_pool and finite and err is None
    oks = [None] * world
    dist.all_gather_object(oks, (ok, res))
    if rank == 0:
        json.dump({"ok": all(o for o, _ in oks), "per_rank": [r for _, r in oks]}, open(os.path.join(out_dir, "lenet.json"), "w"), indent=1)
    dist.barrier()
    t.close()
    return 0 if all(o for o, _ in oks) else 1


def main():
    out_dir, cmd, args = sys.argv[1], sys.argv[2], sys.argv[3:]
    rc = {"kernels": cmd_kernels, "stress": cmd_stress, "train": cmd_train, "ckpt": cmd_ckpt, "init": cmd_init,
          "fault": cmd_fault, "diag": cmd_diag, "lenet": cmd_lenet}[cmd](out_dir, *args)
    if dist.is_initialized():
        dist.destroy_process_group()
    sys.exit(rc)




# ------------------------------------------------------------------------------------------------ diag
def cmd_diag(out_dir, model_name="bert_tiny", algo="twoshot", bucket_mb="0.05", steps="3"):
    """Step-by-step replica comparison with per-bucket / per-slice detail (analysis aid for multi-bucket runs)."""
    rank, world, dev = setup()
    t = _make_trainer(model_name, "p2p", algo, "bf16", float(bucket_mb), out_dir, epochs=1)
    opt, comm = t.optimizer, t.comm
    rep = {"buckets": [(b.lo, b.hi, comm.algo_for(b.lo, b.hi, opt.grad.element_size())) for b in t.reducer.buckets], "steps": []}

    def cmp(tensor):
        parts = [torch.empty_like(tensor) for _ in range(world)]
        dist.all_gather(parts, tensor.contiguous())
        out = []
        for b in t.reducer.buckets:
            d = (parts[0][b.lo:b.hi].float() - parts[1][b.lo:b.hi].float()).abs()
            nz = d.nonzero().flatten()
            out.append(None if nz.numel() == 0 else dict(lo=b.lo, hi=b.hi, n=int(nz.numel()), first=int(nz[0]) + b.lo, last=int(nz[-1]) + b.lo,
                                                       max=float(d.max())))
        return out

    t.model.train()
    feed = iter(t._feed(t.train_loader))
    for s in range(int(steps)):
        x, y = next(feed)
        t.engine.train_step(x, y)
        torch.cuda.synchronize()
        dist.barrier()
        rep["steps"].append({"model": cmp(opt.model), "grad": cmp(opt.grad)})
    before = cmp(opt.master)
    t.consolidate()
    torch.cuda.synchronize()
    rep["master_before_consolidate"] = before
    rep["master_after_consolidate"] = cmp(opt.master)
    rep["s1_after_consolidate"] = cmp(opt.s1)
    rep["master_vs_model"] = float((opt.master.to(torch.bfloat16).float() - opt.model.float()).abs().max())
    slices = [comm.owned_slices(opt, b.lo, b.hi) for b in t.reducer.buckets]
    rep["slices"] = slices[:4]
    if rank == 0:
        json.dump(rep, open(os.path.join(out_dir, f"diag_{model_name}_{algo}_{bucket_mb}.json"), "w"), indent=1, default=str)
    dist.barrier()
    t.close()
    return 0


if __name__ == "__main__":
    main()
