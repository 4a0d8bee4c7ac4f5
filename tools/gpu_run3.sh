#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_resnet.py tests/test_gpu_transformer.py tests/test_gpu_ops.py -q -x > gpurun_out/pytest_gemm.log 2>&1; tail -4 gpurun_out/pytest_gemm.log
timeout 600 python bench/resnet_gemm_shapes.py gpurun_out/resnet_gemm_shapes_db.json > gpurun_out/resnet_gemm_shapes_db.log 2>&1; tail -2 gpurun_out/resnet_gemm_shapes_db.log
timeout 600 python bench.py --model resnet50 --steps 10 --warmup 3 --repeats 3 > gpurun_out/bench_resnet_db.json 2> gpurun_out/bench_resnet_db.err; tail -c 400 gpurun_out/bench_resnet_db.json
timeout 600 python bench.py --model bert_base --steps 10 --warmup 3 --repeats 3 > gpurun_out/bench_bert_db.json 2> gpurun_out/bench_bert_db.err; tail -c 400 gpurun_out/bench_bert_db.json
