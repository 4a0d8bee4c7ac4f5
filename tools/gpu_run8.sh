#!/bin/bash
# eight-GPU validation: headline bench (with secondaries), executor variant, collectives at world 8, bucket sweep to 1 GB
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29711 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ours_n$N.json 2> gpurun_out/bench_ours_n$N.err; echo "bench ours rc=$?"
tail -c 600 gpurun_out/bench_ours_n$N.json
MLT_LENET_PERSISTENT=1 timeout 400 $TR --master-port 29713 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_persistent.json 2> gpurun_out/bench_ours_n${N}_persistent.err; echo "bench persistent rc=$?"
tail -c 300 gpurun_out/bench_ours_n${N}_persistent.json
MLT_TEST_WORLD=$N timeout 900 python -m pytest tests/test_multigpu.py -q -k "kernels or stress or initial_state" > gpurun_out/pytest_multigpu_w$N.log 2>&1; tail -5 gpurun_out/pytest_multigpu_w$N.log
timeout 600 $TR --master-port 29719 bench/allreduce_sweep.py --max-mb 1024 --out gpurun_out/allreduce_sweep_${N}gpu.json > gpurun_out/sweep_n$N.log 2>&1; echo "sweep rc=$?"
tail -2 gpurun_out/sweep_n$N.log | cut -c1-300
