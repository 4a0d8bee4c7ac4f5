#!/bin/bash
# ncu captures (--set full, source import) of the kernels under study; each command first runs plain and must exit 0
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/ncu
NCU="ncu --set full --clock-control none --import-source on"
python bench/gemm_one.py 802816 256 64 fwd > gpurun_out/ncu/gemm_plain.log 2>&1 && \
  $NCU -k regex:gemm_bf16_tcgen05 -s 6 -c 1 -o gpurun_out/ncu/gemm_shortk -f python bench/gemm_one.py 802816 256 64 fwd > gpurun_out/ncu/gemm_ncu.log 2>&1
echo "gemm ncu rc=$?"
python bench/attention_time.py > gpurun_out/ncu/attn_plain.log 2>&1 && \
  $NCU -k regex:attention_ -s 12 -c 4 -o gpurun_out/ncu/attention -f python bench/attention_time.py > gpurun_out/ncu/attn_ncu.log 2>&1
echo "attention ncu rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ncu/smoke_plain.log 2>&1 && \
  $NCU -k regex:lenet_step -c 2 -o gpurun_out/ncu/lenet_step -f python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ncu/smoke_ncu.log 2>&1
echo "lenet ncu rc=$?"
tail -3 gpurun_out/ncu/*_ncu.log
timeout 600 python -m pytest tests/test_gpu_resnet.py -q -x > gpurun_out/pytest_resnet.log 2>&1; tail -3 gpurun_out/pytest_resnet.log
