#!/bin/bash
# usage: gpu_ncu_list2.sh <out> <skip> <count> <cmd...>
OUT=$1; SKIP=$2; CNT=$3; shift 3
mkdir -p gpurun_out
"$@" > gpurun_out/plain_$OUT.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c $CNT --csv --log-file gpurun_out/$OUT.csv "$@" > gpurun_out/ncu_$OUT.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$OUT.log | cut -c1-300
