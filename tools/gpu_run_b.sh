#!/bin/bash
# 2 ranks: multi-GPU tests, where an epoch spends its host time, headline bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_multigpu.py -q > gpurun_out/pytest_multigpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_multigpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29811 bench/epoch_profile.py 20 7 > gpurun_out/epoch_profile_n2.json 2> gpurun_out/epoch_profile_n2.err; echo "rc=$?"
python - <<'P'
import json
d=json.load(open("gpurun_out/epoch_profile_n2.json"))
print({k:[round(x) for x in v[2:7]] for k,v in d["stages_us"].items()})
for e in d["epochs"][2:6]:
    g=e["gpu_steps"]; print(round(e["wall_us"]), round(e["event_us"]), {k:round(v) for k,v in e["native"].items()}, {k:v for k,v in g.items() if k!="gaps_us"})
P
timeout 400 $TR --master-port 29741 bench.py --gpus 2 --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n2_new.json 2> gpurun_out/bench_ours_n2_new.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.load(open("gpurun_out/bench_ours_n2_new.json")); e=d["e2e"]; print(d["value"], d["ms_per_step"], e["value"], e["ms_per_step"], e.get("runs_ms"), e.get("executor_host_us"))
P
