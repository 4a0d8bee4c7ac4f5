#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
for R in 4 6 8; do
MLT_BN_R_STATS=$R timeout 600 python bench.py --model resnet50 --steps 10 --warmup 3 --repeats 3 > gpurun_out/bench_resnet_bnr$R.json 2> gpurun_out/bench_resnet_bnr$R.err; tail -c 250 gpurun_out/bench_resnet_bnr$R.json; echo
done
timeout 300 python bench/profile_step.py bert_base > gpurun_out/profile_bert.txt 2>&1
timeout 300 python bench/attention_time.py > gpurun_out/attention_time.json 2>&1; tail -1 gpurun_out/attention_time.json
