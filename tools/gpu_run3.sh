#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/ncu
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_resnet.py tests/test_gpu_transformer.py tests/test_gpu_ops.py tests/test_gpu_im2col_tma.py -q > gpurun_out/pytest_gemm.log 2>&1; tail -6 gpurun_out/pytest_gemm.log
timeout 600 python bench/resnet_gemm_shapes.py gpurun_out/resnet_gemm_shapes_lean.json > gpurun_out/resnet_gemm_shapes_lean.log 2>&1; tail -1 gpurun_out/resnet_gemm_shapes_lean.log
timeout 600 python bench/gemm_shapes.py > gpurun_out/bert_gemm_shapes.log 2>&1; tail -20 gpurun_out/bert_gemm_shapes.log
timeout 600 python bench.py --model resnet50 --steps 10 --warmup 3 --repeats 3 > gpurun_out/bench_resnet_lean.json 2> gpurun_out/bench_resnet_lean.err; tail -c 300 gpurun_out/bench_resnet_lean.json
timeout 600 python bench.py --model bert_base --steps 10 --warmup 3 --repeats 3 > gpurun_out/bench_bert_lean.json 2> gpurun_out/bench_bert_lean.err; tail -c 300 gpurun_out/bench_bert_lean.json
timeout 300 python bench/profile_step.py resnet50 > gpurun_out/profile_resnet50.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ncu/smoke_plain.log 2>&1 && \
  ncu --set full --clock-control none --import-source on --replay-mode application -k regex:lenet_step -c 1 -o gpurun_out/ncu/lenet_step_app -f python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ncu/smoke_ncu_app.log 2>&1
echo "lenet ncu (application replay) rc=$?"
