#!/bin/bash
# GPU visit for row A9: cv1 parity tests first, then the whole gpu suite, smoke and the bench.  Usage: gpu_cv1.sh <tag>
cd "$(dirname "$0")/.."
TAG=${1:-cv1}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cv1_gpu.py -q -s 2>&1 | tail -40 > gpurun_out/${TAG}_pytest_cv1.log
echo "cv1_exit=${PIPESTATUS[0]}" >> gpurun_out/${TAG}_pytest_cv1.log
cat gpurun_out/${TAG}_pytest_cv1.log
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_cv1_gpu.py 2>&1 | tail -15 > gpurun_out/${TAG}_pytest_gpu.log
echo "pytest_exit=${PIPESTATUS[0]}" >> gpurun_out/${TAG}_pytest_gpu.log
tail -6 gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
echo "smoke_exit=$?" | tee -a gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench_exit=$?"
cat gpurun_out/${TAG}_bench.json | cut -c1-600
