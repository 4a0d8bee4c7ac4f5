#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/diag
L="python -m ml_trainer_b200.parallel.launch --nproc 2 tests/p2p_worker.py gpurun_out/diag"
MLT_LOG_LEVEL=WARNING timeout 300 $L diag bert_tiny twoshot 0.05 3 > gpurun_out/diag/a.log 2>&1; echo "a rc=$?"
mv gpurun_out/diag/diag_bert_tiny_twoshot_0.05.json gpurun_out/diag/a_mc.json
MLT_MC_FLAGS=0 MLT_LOG_LEVEL=WARNING timeout 300 $L diag bert_tiny twoshot 0.05 3 > gpurun_out/diag/b.log 2>&1; echo "b rc=$?"
mv gpurun_out/diag/diag_bert_tiny_twoshot_0.05.json gpurun_out/diag/b_nomc.json
MLT_LOG_LEVEL=WARNING timeout 300 $L diag bert_tiny twoshot 25 3 > gpurun_out/diag/c.log 2>&1; echo "c rc=$?"
MLT_LOG_LEVEL=WARNING timeout 300 $L diag resnet_tiny twoshot 0.05 3 > gpurun_out/diag/d.log 2>&1; echo "d rc=$?"
MLT_LOG_LEVEL=WARNING timeout 300 $L diag bert_tiny nvls 0.05 3 > gpurun_out/diag/e.log 2>&1; echo "e rc=$?"
timeout 600 python bench/resnet_gemm_shapes.py gpurun_out/resnet_gemm_shapes.json > gpurun_out/resnet_gemm_shapes.log 2>&1; echo "gemm rc=$?"
timeout 300 python -m pytest tests/test_gpu_transformer.py -q -x > gpurun_out/pytest_transformer.log 2>&1; tail -3 gpurun_out/pytest_transformer.log
