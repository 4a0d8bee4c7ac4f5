#!/bin/bash
# quick single-GPU check after a kernel change: transformer tests, BERT-base step with and without the skip-gradient fusion
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_transformer.py -q > gpurun_out/pytest_check.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/pytest_check.log
for v in 1 0; do
MLT_SKIP_GRAD_FUSION=$v timeout 600 python bench.py --model bert_base --steps 20 --warmup 5 --secondary none > gpurun_out/bench_bert_skip$v.json 2> gpurun_out/bench_bert_skip$v.err; echo "bert skip=$v rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_bert_skip$v.json')); print('bert skip=$v', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d.get('launches_by_kernel'))"
done
