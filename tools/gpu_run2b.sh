#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/test_multigpu.py -q > gpurun_out/pytest_multigpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multigpu.log
tail -12 gpurun_out/pytest_multigpu.log
