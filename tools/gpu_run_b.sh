#!/bin/bash
# where an end-to-end MLModel step goes on the GPU: per-step %globaltimer stamps, gather threads 1 / 2 / 3
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_trainer.py tests/test_gpu_ops.py -q -x -k "native or executor or loop or augment or epoch" > gpurun_out/pytest_native.log 2>&1; echo "native tests rc=$?"; tail -3 gpurun_out/pytest_native.log
for g in 1 2 3; do
MLT_GATHER_THREADS=$g timeout 300 python bench/epoch_profile.py 20 7 > gpurun_out/epoch_profile_g$g.json 2> gpurun_out/epoch_profile_g$g.err; echo "g$g rc=$?"
done
python - <<'P'
import json
for n in ("g1","g2","g3"):
    try:
        d=json.load(open(f"gpurun_out/epoch_profile_{n}.json"))
        for e in d["epochs"][2:6]:
            g=e["gpu_steps"]; print(n, round(e["wall_us"]), {k:round(v) for k,v in e["native"].items()}, {k:v for k,v in g.items() if k!="gaps_us"}, g.get("gaps_us"))
    except Exception as ex: print(n, "ERR", ex)
P
for g in 1 2 3; do
MLT_GATHER_THREADS=$g timeout 300 python bench.py --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n1_g$g.json 2> gpurun_out/bench_ours_n1_g$g.err; echo "bench g$g rc=$?"
done
python - <<'P'
import json
for n in ("g1","g2","g3"):
    try:
        d=json.load(open(f"gpurun_out/bench_ours_n1_{n}.json")); print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("runs_ms"))
    except Exception as ex: print(n, ex)
P
