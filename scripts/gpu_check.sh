#!/bin/bash
# run on the GPU box: python -m pytest on the gpu tests with output captured under gpurun_out/
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout ${1:-600} python -m pytest tests -m gpu -x -q -s 2>&1 | tail -60 > gpurun_out/pytest_gpu.log
echo "exit=$?" >> gpurun_out/pytest_gpu.log
cat gpurun_out/pytest_gpu.log
