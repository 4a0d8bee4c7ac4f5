#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_transformer.py -q -m gpu 2>&1 | tail -5
echo "== bert ours"
timeout 900 python bench.py --model bert_base --steps 8 --warmup 3 > gpurun_out/bert_ours.json 2> gpurun_out/bert_ours.err; echo rc=$?; cat gpurun_out/bert_ours.json | cut -c1-1500; tail -5 gpurun_out/bert_ours.err
echo "== bert reference"
timeout 900 python bench.py --model bert_base --impl reference --steps 8 --warmup 3 > gpurun_out/bert_ref.json 2> gpurun_out/bert_ref.err; echo rc=$?; cat gpurun_out/bert_ref.json | cut -c1-1200; tail -5 gpurun_out/bert_ref.err | cut -c1-400
