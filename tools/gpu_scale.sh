#!/bin/bash
# Full scaling table on one 8-GPU box: bench (both arms) and the all-reduce sweep at N = 8, 4, 2 (+ ours at 1).
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
PORT=29600
run() { # N impl extra...
  N=$1; IMPL=$2; shift 2
  PORT=$((PORT+1))
  if [ "$N" = "1" ]; then
    timeout 600 python bench.py --gpus 1 --impl $IMPL --steps 200 --warmup 20 "$@" 2> gpurun_out/scale_${IMPL}_$N.err | tail -1 > gpurun_out/scale_${IMPL}_$N.json
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --impl $IMPL --steps 200 --warmup 20 "$@" 2> gpurun_out/scale_${IMPL}_$N.err | tail -1 > gpurun_out/scale_${IMPL}_$N.json
  fi
  echo "N=$N $IMPL rc=$? $(cut -c1-330 gpurun_out/scale_${IMPL}_$N.json)"
}
for N in 8 4 2 1; do run $N ours; run $N reference; done
for N in 8 4 2; do
  PORT=$((PORT+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench/allreduce_sweep.py --max-mb 256 --out gpurun_out/allreduce_sweep_$N.json > gpurun_out/sweep_$N.log 2>&1
  echo "sweep N=$N rc=$?"; tail -4 gpurun_out/sweep_$N.log | cut -c1-400
done
