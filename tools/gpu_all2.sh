#!/bin/bash
# full single-GPU record: GPU tests, then both arms of every bench config -> gpurun_out/rec_<model>_<impl>.json
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
for M in mlmodel bert_base resnet50; do
  for I in reference ours; do
    S=10; [ "$M" = "mlmodel" ] && S=200
    timeout 900 python bench.py --model $M --impl $I --steps $S --warmup 4 2> gpurun_out/rec_${M}_${I}.err | tail -1 > gpurun_out/rec_${M}_${I}.json
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/rec_${M}_${I}.json')); e=d.get('e2e',{})
    print('$M $I', 'value', d.get('value'), 'ms', d.get('ms_per_step'), 'e2e', e.get('value'), e.get('ms_per_step'), e.get('executor'), 'clk', d.get('clocks',{}).get('sm_mhz'))
except Exception as ex:
    print('$M $I bad json', ex); print(open('gpurun_out/rec_${M}_${I}.err').read()[-800:])
PY
  done
done
