#!/bin/bash
# two-GPU check: the whole GPU test-suite (multi-GPU tests included) and optional bench / sweep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
MODE=${2:-tests}
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_multi.log
tail -12 gpurun_out/pytest_gpu_multi.log
if [ "$MODE" = "all" ]; then
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ours_n$N.json 2> gpurun_out/bench_ours_n$N.err; echo "bench ours rc=$?"
tail -c 700 gpurun_out/bench_ours_n$N.json
timeout 900 $TR --master-port 29617 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "bench ref rc=$?"
timeout 900 $TR --master-port 29619 bench/allreduce_sweep.py --max-mb 1024 --out gpurun_out/allreduce_sweep_${N}gpu.json > gpurun_out/sweep_n$N.log 2>&1; echo "sweep rc=$?"
fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 20 --warmup 5 --scaling strong --secondary none > gpurun_out/bench_ours_strong_n$N.json 2> gpurun_out/bench_ours_strong_n$N.err; echo "strong rc=$?"
