"""One-process-per-GPU launcher (replaces the SageMaker estimator + smdistributed/mpirun launch,
reference 02_ML_Training_SageMaker_distributed.ipynb:115-158).

Env contract handed to every rank (same as torchrun, so either launcher works):
``RANK, LOCAL_RANK, WORLD_SIZE, MASTER_ADDR, MASTER_PORT`` plus optional ``SM_*`` pass-through.

Failure handling: the first rank that exits non-zero takes the whole group
down (SIGTERM, then SIGKILL after a grace period) instead of leaving peers spinning on a
flag that will never be written.

    python -m ml_trainer_b200.parallel.launch --nproc 8 main.py --synthetic --epochs 1
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List, Optional


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(cmd: List[str], nproc: int, master_addr: str = "127.0.0.1", master_port: Optional[int] = None,
           env_extra: Optional[dict] = None, grace_s: float = 10.0) -> int:
    port = master_port or free_port()
    procs = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nproc), "LOCAL_WORLD_SIZE": str(nproc),
                    "MASTER_ADDR": master_addr, "MASTER_PORT": str(port)})
        env.update(env_extra or {})
        procs.append(subprocess.Popen(cmd, env=env, start_new_session=True))
    rc = 0
    alive = set(range(nproc))
    try:
        while alive:
            for r in list(alive):
                code = procs[r].poll()
                if code is None:
                    continue
                alive.discard(r)
                if code != 0 and rc == 0:
                    rc = code
                    sys.stderr.write(f"[launch] rank {r} exited with {code}; stopping the other ranks\n")
                    _stop(procs, alive, grace_s)
                    alive.clear()
            time.sleep(0.05)
    except KeyboardInterrupt:
        _stop(procs, alive, grace_s)
        rc = 130
    return rc


def _stop(procs, alive, grace_s):
    for r in alive:
        try:
            os.killpg(procs[r].pid, signal.SIGTERM)     # each rank is its own session/process group
        except ProcessLookupError:
            pass
    t0 = time.time()
    while time.time() - t0 < grace_s and any(procs[r].poll() is None for r in alive):
        time.sleep(0.1)
    for r in alive:
        if procs[r].poll() is None:
            try:
                os.killpg(procs[r].pid, signal.SIGKILL)
            except ProcessLookupError:
                pass


def spawn_self(nproc: int) -> int:
    """Re-run the current command line under ``nproc`` ranks (used by ``main.py --nproc N``)."""
    return launch([sys.executable, *sys.argv], nproc)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--nproc", type=int, default=0, help="ranks to spawn (default: visible GPUs, else 1)")
    ap.add_argument("--master-addr", default="127.0.0.1")
    ap.add_argument("--master-port", type=int, default=0)
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    n = a.nproc
    if n <= 0:
        try:
            import torch

            n = max(torch.cuda.device_count(), 1)
        except Exception:
            n = 1
    return launch([sys.executable, a.script, *a.script_args], n, a.master_addr, a.master_port or None)


if __name__ == "__main__":
    raise SystemExit(main())
