#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-q}
mkdir -p gpurun_out
L=moss_speech_decoder_cosy_b200/lib
$L/test_ffn > gpurun_out/${TAG}_ffn_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ffn_tc -s 4 -c 1 -o gpurun_out/${TAG}_ffn $L/test_ffn > gpurun_out/${TAG}_ffn_ncu.log 2>&1
echo "ncu exit=$?"
