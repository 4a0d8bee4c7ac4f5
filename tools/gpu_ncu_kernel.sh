#!/bin/bash
# usage: gpu_ncu_kernel.sh <kernel-regex> <outname> <cmd...>
PAT=$1; OUT=$2; shift 2
mkdir -p gpurun_out
"$@" > gpurun_out/plain_$OUT.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$PAT -s 4 -c 2 -o gpurun_out/$OUT -f "$@" > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$OUT.log
