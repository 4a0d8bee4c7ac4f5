#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 200 --warmup 20 2> gpurun_out/w_ours_$N.err | tail -1 | tee gpurun_out/w_ours_$N.json | cut -c1-900
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --impl reference --gpus $N --steps 200 --warmup 20 2> gpurun_out/w_ref_$N.err | tail -1 | tee gpurun_out/w_ref_$N.json | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29713 bench/allreduce_sweep.py --max-mb 64 --out gpurun_out/allreduce_sweep_graph_$N.json 2>&1 | grep bytes | cut -c1-420
