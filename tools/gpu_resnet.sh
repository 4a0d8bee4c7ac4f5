#!/bin/bash
mkdir -p gpurun_out
echo "== resnet ours"
timeout 900 python bench.py --model resnet50 --steps 6 --warmup 3 > gpurun_out/resnet_ours.json 2> gpurun_out/resnet_ours.err; echo rc=$?; cat gpurun_out/resnet_ours.json | cut -c1-1300; tail -5 gpurun_out/resnet_ours.err
echo "== resnet reference"
timeout 900 python bench.py --model resnet50 --impl reference --steps 6 --warmup 3 > gpurun_out/resnet_ref.json 2> gpurun_out/resnet_ref.err; echo rc=$?; cat gpurun_out/resnet_ref.json | cut -c1-700; tail -3 gpurun_out/resnet_ref.err | cut -c1-300
