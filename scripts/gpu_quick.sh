#!/bin/bash
# quick GPU visit: device unit tests, gpu parity tests, bench with per-shape profile; no ncu.  Usage: gpu_quick.sh <tag>
cd "$(dirname "$0")/.."
TAG=${1:-q}
mkdir -p gpurun_out
L=moss_speech_decoder_cosy_b200/lib
for t in test_ffn test_gemm; do
  if [ -x $L/$t ]; then timeout 300 $L/$t > gpurun_out/${TAG}_$t.log 2>&1; echo "$t exit=$?" | tee -a gpurun_out/${TAG}_$t.log; tail -8 gpurun_out/${TAG}_$t.log; fi
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 > gpurun_out/${TAG}_pytest_gpu.log
echo "pytest_exit=${PIPESTATUS[0]}" >> gpurun_out/${TAG}_pytest_gpu.log
tail -12 gpurun_out/${TAG}_pytest_gpu.log
MOSSDEC_PROFILE_DUMP=1 timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench_exit=$?"
tail -3 gpurun_out/${TAG}_bench.err | cut -c1-300
cat gpurun_out/${TAG}_bench.json
