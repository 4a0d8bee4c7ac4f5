#!/bin/bash
# multi-GPU check: N = number of GPUs on the box
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
echo "== pytest multigpu"
timeout 900 python -m pytest tests/test_multigpu.py -x -q -m gpu > gpurun_out/pytest_multigpu.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_multigpu.log
echo "== bench ours N=$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/bench_ours_$N.json 2> gpurun_out/bench_ours_$N.err; echo "rc=$?"; cat gpurun_out/bench_ours_$N.json; tail -5 gpurun_out/bench_ours_$N.err
echo "== bench reference N=$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 200 --warmup 20 > gpurun_out/bench_ref_$N.json 2> gpurun_out/bench_ref_$N.err; echo "rc=$?"; cat gpurun_out/bench_ref_$N.json; tail -3 gpurun_out/bench_ref_$N.err | cut -c1-300
