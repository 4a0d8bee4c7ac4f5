#!/bin/bash
# eight-GPU check of the headline: default settings (the record), the same with per-step GPU stamps, old waits / one gather
# thread, and a deep ring
N=${1:-8}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nproc
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
B="bench.py --gpus $N --steps 20 --warmup 5 --secondary none"
timeout 300 $TR --master-port 29711 $B > gpurun_out/bench_ours_n${N}_default.json 2> gpurun_out/bench_ours_n${N}_default.err; echo "default rc=$?"
MLT_BENCH_STAMPS=1 timeout 300 $TR --master-port 29713 $B > gpurun_out/bench_ours_n${N}_stamps.json 2> gpurun_out/bench_ours_n${N}_stamps.err; echo "stamps rc=$?"
MLT_WAIT_SPIN=1 MLT_GATHER_THREADS=1 timeout 300 $TR --master-port 29715 $B > gpurun_out/bench_ours_n${N}_oldwaits.json 2> gpurun_out/bench_ours_n${N}_oldwaits.err; echo "oldwaits rc=$?"
MLT_RING_SLOTS=12 MLT_BENCH_STAMPS=1 timeout 300 $TR --master-port 29717 $B > gpurun_out/bench_ours_n${N}_ring12b.json 2> gpurun_out/bench_ours_n${N}_ring12b.err; echo "ring12 rc=$?"
python - <<P
import json
for n in ("default","stamps","oldwaits","ring12b"):
    try:
        d=json.load(open(f"gpurun_out/bench_ours_n${N}_{n}.json")); e=d["e2e"]
        print(n, d["value"], d["ms_per_step"], e["value"], e["ms_per_step"], e.get("runs_ms"), e.get("executor_host_us"), e.get("cpus_available"))
        g=e.get("gpu_steps_last_run")
        if g: print("   body", g["body_us"]); print("   gap ", g["gap_us"])
    except Exception as ex: print(n, ex)
P
