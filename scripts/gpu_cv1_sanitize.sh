#!/bin/bash
# compute-sanitizer memcheck over the row-A9 parity tests (bounded by timeout).  Usage: gpu_cv1_sanitize.sh <tag>
cd "$(dirname "$0")/.."
TAG=${1:-san}
mkdir -p gpurun_out
timeout 420 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_cv1_gpu.py -q -x -k "match_reference or missing" \
  > gpurun_out/${TAG}_memcheck.log 2>&1
echo "memcheck_exit=$?" | tee -a gpurun_out/${TAG}_memcheck.log
grep -E "ERROR SUMMARY|Invalid|passed|failed" gpurun_out/${TAG}_memcheck.log | head -20
