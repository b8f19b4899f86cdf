#!/bin/bash
# Row A9 measurement: plain timed run, then (after it exited 0) the ncu launch list with durations and DRAM bytes.
cd "$(dirname "$0")/.."
TAG=${1:-cv1p}
mkdir -p gpurun_out
if [ -n "$2" ]; then
  timeout 600 python -m pytest tests/test_cv1_gpu.py -q -s 2>&1 | tail -30 > gpurun_out/${TAG}_pytest_cv1.log
  echo "cv1_exit=${PIPESTATUS[0]}" >> gpurun_out/${TAG}_pytest_cv1.log
  cat gpurun_out/${TAG}_pytest_cv1.log
fi
timeout 300 python scripts/bench_cv1.py --iters 10 > gpurun_out/${TAG}_bench_cv1.jsonl 2> gpurun_out/${TAG}_bench_cv1.err
BE=$?
echo "bench_cv1_exit=$BE"; cat gpurun_out/${TAG}_bench_cv1.jsonl; tail -3 gpurun_out/${TAG}_bench_cv1.err
if [ $BE -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:'gn_|cv1_|gemm_tc' --csv --log-file gpurun_out/${TAG}_launches.csv python scripts/bench_cv1.py --iters 1 \
    > gpurun_out/${TAG}_ncu.log 2>&1
  echo "ncu_exit=$?"
  wc -l gpurun_out/${TAG}_launches.csv
fi
