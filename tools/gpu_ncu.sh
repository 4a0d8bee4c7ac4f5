#!/bin/bash
# ncu captures (--set full, source import) of the kernels under study; each command first runs plain and must exit 0.
# (The single-kernel LeNet step is NOT profiled here: a cooperative cluster launch does not survive ncu's replay on this stack.)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/ncu
NCU="ncu --set full --clock-control none --import-source on"
python bench/gemm_one.py 802816 256 64 fwd > gpurun_out/ncu/gemm_plain.log 2>&1 && \
  $NCU -k regex:gemm_bf16_tcgen05 -s 6 -c 1 -o gpurun_out/ncu/gemm_shortk_lean -f python bench/gemm_one.py 802816 256 64 fwd > gpurun_out/ncu/gemm_ncu.log 2>&1
echo "gemm ncu rc=$?"
python bench/bn_one.py 802816 256 relu > gpurun_out/ncu/bn_plain.log 2>&1 && \
  $NCU -k regex:bn_ -s 8 -c 4 -o gpurun_out/ncu/bn_802816x256 -f python bench/bn_one.py 802816 256 relu > gpurun_out/ncu/bn_ncu.log 2>&1
echo "bn ncu rc=$?"
python bench/bn_one.py 802816 64 relu > gpurun_out/ncu/bn64_plain.log 2>&1 && \
  $NCU -k regex:bn_stats -s 2 -c 1 -o gpurun_out/ncu/bn_802816x64 -f python bench/bn_one.py 802816 64 relu > gpurun_out/ncu/bn64_ncu.log 2>&1
echo "bn64 ncu rc=$?"
tail -2 gpurun_out/ncu/*_ncu.log
timeout 300 python bench/bn_shapes.py gpurun_out/bn_shapes.json > gpurun_out/bn_shapes.log 2>&1; echo "bn shapes rc=$?"; tail -20 gpurun_out/bn_shapes.log
