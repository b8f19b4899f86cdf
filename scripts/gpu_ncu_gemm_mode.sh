#!/bin/bash
# ncu --set full of one launch of a gemm_tc_kernel<mode> instantiation inside test_gemm's bench cases.
# usage: gpu_ncu_gemm_mode.sh <tag> <mode> <skip>
cd "$(dirname "$0")/.."
TAG=$1; MODE=$2; SKIP=${3:-20}
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
GEMM_BENCH=1 timeout 100 moss_speech_decoder_cosy_b200/lib/test_gemm > gpurun_out/${TAG}_plain.log 2>&1 || exit 1
GEMM_BENCH=1 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  --kernel-name "regex:gemm_tc_kernel<\(int\)${MODE}>" -s $SKIP -c 1 -o gpurun_out/${TAG}_mode${MODE} \
  moss_speech_decoder_cosy_b200/lib/test_gemm > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit=$?"; ls -la gpurun_out/*.ncu-rep
find gpurun_out -size +40M -delete
