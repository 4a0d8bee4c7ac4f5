#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -8 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_quick.json 2> gpurun_out/bench_ours_quick.err; tail -c 300 gpurun_out/bench_ours_quick.json
