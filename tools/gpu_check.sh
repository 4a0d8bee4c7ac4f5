#!/bin/bash
# First-contact GPU run: tests, smoke, short bench of both arms.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest gpu" 
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
echo "== smoke"
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
echo "== bench ours"
timeout 600 python bench.py --steps 200 --warmup 20 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "rc=$?"; cat gpurun_out/bench_ours.json; tail -5 gpurun_out/bench_ours.err
echo "== bench ours (metric=accuracy, no graph)"
timeout 600 python bench.py --steps 100 --warmup 10 --no-graph --metric accuracy > gpurun_out/bench_ours_eager.json 2> gpurun_out/bench_ours_eager.err; echo "rc=$?"; cat gpurun_out/bench_ours_eager.json; tail -3 gpurun_out/bench_ours_eager.err
echo "== bench reference"
timeout 600 python bench.py --impl reference --steps 200 --warmup 20 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?"; cat gpurun_out/bench_ref.json; tail -5 gpurun_out/bench_ref.err
