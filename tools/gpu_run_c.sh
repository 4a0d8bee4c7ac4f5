#!/bin/bash
# N-GPU headline, both arms, no secondary models
N=${1:-4}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29741 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_headline.json 2> gpurun_out/bench_ours_n${N}_headline.err; echo "ours rc=$?"
timeout 400 $TR --master-port 29743 bench.py --impl reference --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ref_n${N}_headline.json 2> gpurun_out/bench_ref_n${N}_headline.err; echo "ref rc=$?"
python - <<P
import json
for n in ("ours","ref"):
    try:
        d=json.load(open(f"gpurun_out/bench_{n}_n${N}_headline.json")); e=d["e2e"]; print(n, d["value"], d["ms_per_step"], e["value"], e["ms_per_step"], e.get("cpus_available"))
    except Exception as ex: print(n, ex)
P
