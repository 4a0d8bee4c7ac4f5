#!/bin/bash
# N-GPU check of every bench config (ours), plus the multi-GPU tests.  usage: gpu_multi4.sh N
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -x -q -m gpu 2>&1 | tail -3
PORT=29730
for M in mlmodel bert_base resnet50; do
  S=10; [ "$M" = "mlmodel" ] && S=300
  PORT=$((PORT+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --model $M --steps $S --warmup 4 2> gpurun_out/m4_${M}_$N.err | tail -1 > gpurun_out/m4_${M}_$N.json
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/m4_${M}_$N.json')); e=d.get('e2e',{})
    print('$M N=$N value', d.get('value'), 'ms', d.get('ms_per_step'), 'e2e', e.get('value'), e.get('ms_per_step'), d['config'].get('comm'), d['config'].get('parallelism'))
except Exception as ex:
    print('$M bad json', ex); print(open('gpurun_out/m4_${M}_$N.err').read()[-1500:])
PY
done
