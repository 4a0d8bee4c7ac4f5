#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py tests/test_gpu_lenet.py -x -q -m gpu 2>&1 | tail -3
python bench.py --steps 300 --warmup 20 2>/dev/null | tail -1 | tee gpurun_out/c_ours_1.json | python -c "import json,sys; d=json.load(sys.stdin); print('N=1 value',d['value'],d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus $N --steps 300 --warmup 20 2> gpurun_out/c_ours_$N.err | tail -1 | tee gpurun_out/c_ours_$N.json | python -c "import json,sys; d=json.load(sys.stdin); print('N=$N value',d['value'],d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'])"
tail -2 gpurun_out/c_ours_$N.err | cut -c1-300
