#!/bin/bash
# usage: gpu_ncu_kernel2.sh <kernel-regex> <outname> <skip> <count> <cmd...>
PAT=$1; OUT=$2; SKIP=$3; CNT=$4; shift 4
mkdir -p gpurun_out
"$@" > gpurun_out/plain_$OUT.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$PAT -s $SKIP -c $CNT -o gpurun_out/$OUT -f "$@" > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$OUT.log
