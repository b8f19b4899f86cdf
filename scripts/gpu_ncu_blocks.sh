#!/bin/bash
# quick visit, then one ncu --set full capture of two transformer blocks (q|k|v GEMM, attention, out-projection, FFN)
cd "$(dirname "$0")/.."
TAG=${1:-q}
bash scripts/gpu_quick.sh $TAG
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_tc_kernel|attention_ws|attention_tc|ffn_tc' \
  -s 500 -c 8 -o gpurun_out/${TAG}_blocks python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_blocks.log 2>&1
echo "ncu exit=$?"
find gpurun_out -size +40M -print -delete
