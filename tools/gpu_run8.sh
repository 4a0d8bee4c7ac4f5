#!/bin/bash
# eight-GPU validation: single-kernel step replicas, collectives at world 8, executor ring depth, bucket sweep to 1 GB
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
MLT_TEST_WORLD=$N timeout 600 python -m pytest tests/test_multigpu.py -q -k "single_kernel or kernels" > gpurun_out/pytest_multigpu_w$N.log 2>&1; tail -5 gpurun_out/pytest_multigpu_w$N.log
timeout 400 $TR --master-port 29713 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_ring12.json 2> gpurun_out/bench_ours_n${N}_ring12.err; echo "bench ring12 rc=$?"
tail -c 300 gpurun_out/bench_ours_n${N}_ring12.json
MLT_RING_SLOTS=3 timeout 400 $TR --master-port 29715 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_ring3.json 2> gpurun_out/bench_ours_n${N}_ring3.err; echo "bench ring3 rc=$?"
tail -c 300 gpurun_out/bench_ours_n${N}_ring3.json
timeout 400 $TR --master-port 29717 bench.py --gpus $N --steps 200 --warmup 5 --secondary none --repeats 3 > gpurun_out/bench_ours_n${N}_k200.json 2> gpurun_out/bench_ours_n${N}_k200.err; echo "bench k200 rc=$?"
timeout 600 $TR --master-port 29719 bench/allreduce_sweep.py --max-mb 1024 --out gpurun_out/allreduce_sweep_${N}gpu.json > gpurun_out/sweep_n$N.log 2>&1; echo "sweep rc=$?"
