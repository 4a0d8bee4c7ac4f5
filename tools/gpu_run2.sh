#!/bin/bash
# two-GPU check: the whole GPU test-suite (multi-GPU tests included), bench at N=2 for both arms, executor variants, sweep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_multi.log
tail -8 gpurun_out/pytest_gpu_multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ours_n$N.json 2> gpurun_out/bench_ours_n$N.err; echo "bench ours rc=$?"
tail -c 700 gpurun_out/bench_ours_n$N.json
MLT_LENET_PERSISTENT=1 timeout 600 $TR --master-port 29613 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_persistent.json 2> gpurun_out/bench_ours_n${N}_persistent.err; echo "bench persistent rc=$?"
MLT_MC_FLAGS=0 timeout 600 $TR --master-port 29615 bench.py --gpus $N --steps 20 --warmup 5 --secondary none > gpurun_out/bench_ours_n${N}_unicast.json 2> gpurun_out/bench_ours_n${N}_unicast.err; echo "bench unicast rc=$?"
timeout 900 $TR --master-port 29617 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "bench ref rc=$?"
tail -c 500 gpurun_out/bench_ref_n$N.json
timeout 900 $TR --master-port 29619 bench/allreduce_sweep.py --max-mb 1024 --out gpurun_out/allreduce_sweep_${N}gpu.json > gpurun_out/sweep_n$N.log 2>&1; echo "sweep rc=$?"
