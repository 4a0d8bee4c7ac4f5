#!/bin/bash
# BatchNorm with batched loads: correctness, per-shape times under grid / row-split variants, ResNet-50 step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_resnet.py tests/test_gpu_conv.py -q > gpurun_out/pytest_bn.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/pytest_bn.log
for v in "8 4" "3 3" "6 3" "4 3"; do
set -- $v
echo "== grid mult $1, stats rows mult $2"
MLT_BN_GRID_MULT=$1 MLT_BN_R_STATS=$2 timeout 300 python bench/bn_shapes.py gpurun_out/bn_shapes_g$1_s$2.json 2>&1 | grep -E "M': (3211264|802816|200704), 'C': (64|256|512)|totals"
done
for v in "8 4" "3 3"; do
set -- $v
MLT_BN_GRID_MULT=$1 MLT_BN_R_STATS=$2 timeout 600 python bench.py --model resnet50 --steps 20 --warmup 5 --secondary none > gpurun_out/bench_resnet_g$1_s$2.json 2> gpurun_out/bench_resnet_g$1_s$2.err; echo "resnet $v rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_resnet_g$1_s$2.json')); print('resnet', '$v', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
