#!/bin/bash
# N-GPU check after executor changes: multi-GPU tests, the full bench line of our arm, strong scaling
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_multigpu.py -q > gpurun_out/pytest_multigpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_multigpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ours_n$N.json 2> gpurun_out/bench_ours_n$N.err; echo "bench ours rc=$?"
timeout 600 $TR --master-port 29621 bench.py --gpus $N --steps 20 --warmup 5 --scaling strong --secondary none > gpurun_out/bench_ours_strong_n$N.json 2> gpurun_out/bench_ours_strong_n$N.err; echo "strong rc=$?"
python - <<P
import json
def row(d):
    e=d.get("e2e",{}); return (d.get("value"), d.get("ms_per_step"), e.get("value"), e.get("ms_per_step"), e.get("executor_host_us"), e.get("cpus_available"))
o=json.load(open("gpurun_out/bench_ours_n$N.json")); print("mlmodel", row(o))
for k,v in o.get("secondary",{}).items(): print(k, row(v))
print("strong", row(json.load(open("gpurun_out/bench_ours_strong_n$N.json"))))
P
