#!/bin/bash
# 8-GPU weak-scaling check: mlmodel both arms, bert / resnet ours.  usage: gpu_scale8.sh N
N=${1:-8}
mkdir -p gpurun_out
PORT=29800
run() { # model impl steps
  M=$1; I=$2; S=$3
  PORT=$((PORT+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --model $M --impl $I --steps $S --warmup 4 2> gpurun_out/s8_${M}_${I}_$N.err | tail -1 > gpurun_out/s8_${M}_${I}_$N.json
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/s8_${M}_${I}_$N.json')); e=d.get('e2e',{})
    print('$M $I N=$N value', d.get('value'), 'ms', d.get('ms_per_step'), 'e2e', e.get('value'), d.get('config',{}).get('comm'))
except Exception as ex:
    print('$M $I bad json', ex); print(open('gpurun_out/s8_${M}_${I}_$N.err').read()[-1200:])
PY
}
run mlmodel ours 300
run mlmodel reference 100
run bert_base ours 10
run resnet50 ours 8
