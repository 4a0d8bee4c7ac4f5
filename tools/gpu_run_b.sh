#!/bin/bash
# where a 2-rank epoch spends its host time
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29811 bench/epoch_profile.py 20 7 > gpurun_out/epoch_profile_n2.json 2> gpurun_out/epoch_profile_n2.err; echo "rc=$?"
tail -5 gpurun_out/epoch_profile_n2.err
python - <<'P'
import json
d=json.load(open("gpurun_out/epoch_profile_n2.json"))
print({k:[round(x) for x in v[2:7]] for k,v in d["stages_us"].items()})
for e in d["epochs"][2:6]:
    g=e["gpu_steps"]; print(round(e["wall_us"]), round(e["event_us"]), {k:round(v) for k,v in e["native"].items()}, {k:v for k,v in g.items() if k!="gaps_us"})
P
