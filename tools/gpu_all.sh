#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -8
for M in bert_base resnet50; do
  timeout 900 python bench.py --model $M --steps 10 --warmup 4 2> gpurun_out/${M}_ours.err | tail -1 > gpurun_out/${M}_ours.json; echo "$M rc=$?"; python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${M}_ours.json')); print('$M', 'value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['executor'], d['e2e']['h2d_bytes_per_step'])
except Exception as e:
    print('bad json', e); print(open('gpurun_out/${M}_ours.err').read()[-1500:])
PY
done
