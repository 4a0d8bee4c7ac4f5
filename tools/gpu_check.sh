#!/bin/bash
# quick single-GPU check after a kernel change: the transformer / ops tests and the BERT-base step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_transformer.py tests/test_gpu_ops.py -q > gpurun_out/pytest_check.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/pytest_check.log
timeout 600 python bench.py --model bert_base --steps 20 --warmup 5 --secondary none > gpurun_out/bench_bert_check.json 2> gpurun_out/bench_bert_check.err; echo "bert rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_bert_check.json')); print('bert', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
timeout 300 python bench/profile_step.py bert_base > gpurun_out/profile_bert.txt 2>&1; head -30 gpurun_out/profile_bert.txt
