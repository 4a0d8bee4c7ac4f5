#!/bin/bash
# attention dropout check: transformer tests, attention timing, fp32-split GEMM test, 1-GPU ring-depth comparison
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_transformer.py -q > gpurun_out/pytest_transformer.log 2>&1; echo "transformer rc=$?"; tail -15 gpurun_out/pytest_transformer.log
timeout 300 python -m pytest tests/test_gpu_gemm.py -q -k "split or f32" > gpurun_out/pytest_gemm_f32.log 2>&1; echo "gemm f32 rc=$?"; tail -3 gpurun_out/pytest_gemm_f32.log
timeout 300 python bench/attention_time.py > gpurun_out/attention_time.json 2> gpurun_out/attention_time.err; echo "attn rc=$?"; cat gpurun_out/attention_time.json
timeout 300 python bench.py --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n1_ring3.json 2> gpurun_out/bench_ours_n1_ring3.err; echo "ring3 rc=$?"
MLT_RING_SLOTS=12 timeout 300 python bench.py --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n1_ring12.json 2> gpurun_out/bench_ours_n1_ring12.err; echo "ring12 rc=$?"
python - <<'P'
import json
for n in ("ring3","ring12"):
    try:
        d=json.load(open(f"gpurun_out/bench_ours_n1_{n}.json")); print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("runs_ms"))
    except Exception as ex: print(n, ex)
P
