#!/bin/bash
# N-GPU headline bench only (no secondaries), default settings and gather-thread / ring variants
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nproc; python -c "import os; print('affinity', len(os.sched_getaffinity(0)))"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29741 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_new.json 2> gpurun_out/bench_ours_n${N}_new.err; echo "bench rc=$?"
MLT_GATHER_THREADS=1 timeout 400 $TR --master-port 29742 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_g1.json 2> gpurun_out/bench_ours_n${N}_g1.err; echo "bench g1 rc=$?"
python - <<P
import json
for n in ("new","g1"):
    try:
        d=json.load(open(f"gpurun_out/bench_ours_n${N}_{n}.json")); e=d["e2e"]; print(n, d["value"], d["ms_per_step"], e["value"], e["ms_per_step"], e.get("runs_ms"), e.get("executor_host_us"), e.get("cpus_available"))
    except Exception as ex: print(n, ex)
P
