"""Worker for the world_size=2 CPU/gloo equivalence test (BASELINE.json config 2, CPU leg).

Each rank trains MLModel through the Trainer in parallel mode on a DistributedSampler shard, then the same
shard through stock DistributedDataParallel + torch.optim, and rank 0 dumps both parameter sets.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ml_trainer_b200 import Loader, MLModel, SyntheticDataset, Trainer  # noqa: E402


def main():
    torch.backends.cudnn.allow_tf32 = False          # the DDP oracle must be plain fp32
    torch.backends.cuda.matmul.allow_tf32 = False
    out_dir, optimizer = sys.argv[1], sys.argv[2]
    backend = sys.argv[3] if len(sys.argv) > 3 else "gloo"
    tr = SyntheticDataset((3, 32, 32), 10, length=128, seed=11)
    va = SyntheticDataset((3, 32, 32), 10, length=64, seed=12)
    torch.manual_seed(32)
    t = Trainer(MLModel(), datasets=(tr, va), epochs=2, batch_size=32, is_parallel=True, save_history=True,
                model_dir=out_dir, backend=backend, optimizer=optimizer, lr=0.01, progress=False, metric="accuracy",
                comm_algo=os.environ.get("MLT_COMM_ALGO", "auto"), fused_step=os.environ.get("MLT_FUSED_STEP", "1") == "1")
    rank, world = dist.get_rank(), dist.get_world_size()
    assert t.batch_size == 32 // world
    assert len(t.train_loader.sampler) == 128 // world
    from ml_trainer_b200.ops import _lib

    _lib.launches.reset()
    t.fit()
    launches = dict(_lib.launches.by_name)
    ours = {k: v.detach().cpu().clone() for k, v in t.model.state_dict().items()}

    # oracle: stock DDP + torch.optim on the same shards, same order (set_epoch like ours)
    dev = t.device
    torch.manual_seed(32)
    ref = MLModel().to(dev)
    # make the oracle use the torch statement of the ops on any device
    ddp = torch.nn.parallel.DistributedDataParallel(ref, device_ids=[dev.index] if dev.type == "cuda" else None)
    mk = {"sgd": lambda p: torch.optim.SGD(p, lr=0.01, momentum=0.9),
          "adamw": lambda p: torch.optim.AdamW(p, lr=0.01, weight_decay=0.0),
          "adam": lambda p: torch.optim.Adam(p, lr=0.01)}[optimizer]
    opt = mk(ddp.parameters())
    sampler = torch.utils.data.distributed.DistributedSampler(tr, num_replicas=world, rank=rank)
    loader = Loader(tr, batch_size=32 // world, sampler=sampler)
    for epoch in (1, 2):
        sampler.set_epoch(epoch)
        for x, y in loader:
            x, y = x.to(dev), y.to(dev)
            opt.zero_grad()
            torch.nn.functional.cross_entropy(ddp(x), y).backward()
            opt.step()
    theirs = {k: v.detach().cpu().clone() for k, v in ref.state_dict().items()}
    hist = t.history
    # the epoch-end form for host values: sums over the ranks, error check included
    host_sum = t.comm.reduce_scalars_host([float(rank + 1), 2.5, 10.0 * (rank + 1)])
    gathered = [None] * world
    dist.all_gather_object(gathered, hist)
    if rank == 0:
        torch.save({"ours": ours, "ddp": theirs, "hist": gathered, "world": world, "launches": launches, "host_sum": host_sum},
                   os.path.join(out_dir, "result.pt"))
    dist.barrier()
    t.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
