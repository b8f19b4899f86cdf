#!/bin/bash
# A/B of several builds of libmossdec.so with the whole bench step (same box, alternating): gpu_ab_bench.sh <tag> <lib>...
cd "$(dirname "$0")/.."
TAG=$1; shift
mkdir -p gpurun_out
for r in 1 2 3; do
  for L in "$@"; do
    MOSSDEC_LIB=$PWD/$L timeout 300 python bench.py --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$L', round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2), d['clocks']['sm_mhz'])" | tee -a gpurun_out/${TAG}_abbench.log
  done
done
