#!/bin/bash
# single-GPU check: GPU test-suite, both bench arms (with the secondary models), epoch host-time profile
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
tail -c 1000 gpurun_out/bench_ours.json
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
tail -c 600 gpurun_out/bench_ref.json
timeout 300 python bench/epoch_profile.py 20 7 > gpurun_out/epoch_profile.json 2> gpurun_out/epoch_profile.err; echo "profile rc=$?"
MLT_LENET_PERSISTENT=1 timeout 300 python bench/epoch_profile.py 20 7 > gpurun_out/epoch_profile_persistent.json 2> gpurun_out/epoch_profile_persistent.err; echo "profile2 rc=$?"
