#!/bin/bash
# One GPU-box visit: gpu tests, smoke, bench (plain), then ncu launch list and one --set full capture (after the plain
# runs exited 0).  Usage: scripts/gpu_round.sh <tag> [skip_tests]
cd "$(dirname "$0")/.."
TAG=${1:-run}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
if [ -z "$2" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/${TAG}_pytest_gpu.log
  echo "pytest_exit=${PIPESTATUS[0]}" >> gpurun_out/${TAG}_pytest_gpu.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
  echo "smoke_exit=$?" >> gpurun_out/${TAG}_smoke.log
fi
MOSSDEC_PROFILE_DUMP=1 timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
BE=$?
echo "bench_exit=$BE"
if [ $BE -eq 0 ]; then
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline \
    > gpurun_out/${TAG}_ncu_launches.log 2>&1
  echo "ncu_launches_exit=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_tc_kernel|attention_ws|attention_tc|ffn_tc' \
    -s 500 -c 8 -o gpurun_out/${TAG}_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline \
    > gpurun_out/${TAG}_ncu_full.log 2>&1
  echo "ncu_full_exit=$?"
fi
# keep gpurun_out under the 64 MiB return limit
find gpurun_out -size +40M -print -delete
du -sh gpurun_out
tail -3 gpurun_out/${TAG}_pytest_gpu.log gpurun_out/${TAG}_smoke.log 2>/dev/null
cat gpurun_out/${TAG}_bench.json
