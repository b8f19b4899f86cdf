#!/bin/bash
# A/B of several builds of libmossdec.so on the same box, alternating: gpu_ab.sh <tag> <lib>... [-- bench_flow args]
cd "$(dirname "$0")/.."
TAG=$1; shift; LIBS=""
while [ $# -gt 0 ] && [ "$1" != "--" ]; do LIBS="$LIBS $1"; shift; done
[ "$1" == "--" ] && shift
mkdir -p gpurun_out
for r in 1 2; do
  for L in $LIBS; do
    echo "== $L" >> gpurun_out/${TAG}_ab.log
    MOSSDEC_LIB=$PWD/$L timeout 300 python scripts/bench_flow.py "$@" >> gpurun_out/${TAG}_ab.log 2>&1
  done
done
cat gpurun_out/${TAG}_ab.log
