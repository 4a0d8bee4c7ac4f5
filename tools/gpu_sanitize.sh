#!/bin/bash
# compute-sanitizer passes over a subset of the GPU tests.  Output -> gpurun_out/sanitizer_*.log
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() { # name tool tests...
  NAME=$1; TOOL=$2; shift 2
  timeout 900 $CS --tool $TOOL --error-exitcode 9 --launch-timeout 0 python -m pytest "$@" -q -m gpu -x -p no:cacheprovider > gpurun_out/sanitizer_$NAME.log 2>&1
  echo "$NAME ($TOOL) rc=$? : $(grep -E 'passed|failed|error' gpurun_out/sanitizer_$NAME.log | tail -1) | $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitizer_$NAME.log | tail -1)"
}
run ops_memcheck memcheck tests/test_gpu_ops.py
run lenet_memcheck memcheck tests/test_gpu_lenet.py -k "matches or kernel"
run gemm_memcheck memcheck tests/test_gpu_gemm.py -k "majors_and_tiles or epilogues or ffn"
run attn_memcheck memcheck tests/test_gpu_transformer.py -k "fused_attention and 256"
run conv_memcheck memcheck tests/test_gpu_im2col_tma.py tests/test_gpu_resnet.py -k "im2col or conv2d or batchnorm"
run ops_racecheck racecheck tests/test_gpu_ops.py
run comm_memcheck memcheck tests/test_gpu_comm.py -k "oneshot or twoshot"
