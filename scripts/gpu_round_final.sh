#!/bin/bash
# Round-end GPU visit: gpu tests, smoke, bench (all sections), then -- after the plain runs exited 0 -- the ncu launch list
# (one metric, bounded launch count) and two --set full captures (estimator kernels; the vocoder residual-conv epilogue).
# Usage: scripts/gpu_round_final.sh <tag>
cd "$(dirname "$0")/.."
TAG=${1:-final}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/${TAG}_pytest_gpu.log
echo "pytest_exit=${PIPESTATUS[0]}" >> gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
echo "smoke_exit=$?" >> gpurun_out/${TAG}_smoke.log
timeout 300 moss_speech_decoder_cosy_b200/lib/test_gemm > gpurun_out/${TAG}_test_gemm.log 2>&1
echo "test_gemm_exit=$?" >> gpurun_out/${TAG}_test_gemm.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
BE=$?
echo "bench_exit=$BE"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err
echo "ref_exit=$?"
if [ $BE -eq 0 ]; then
  ARGS="--steps 1 --warmup 3 --no-cpu-baseline --no-extras"
  timeout 300 python bench.py $ARGS > gpurun_out/${TAG}_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 26000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py $ARGS > gpurun_out/${TAG}_ncu_launches.log 2>&1
  echo "ncu_launches_exit=$?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gemm_tc_kernel|attention_ws|ffn_tc' \
    -s 500 -c 8 -f -o gpurun_out/${TAG}_full python bench.py $ARGS > gpurun_out/${TAG}_ncu_full.log 2>&1
  echo "ncu_full_exit=$?"
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'gemm_tc_kernel<\(int\)5>' -s 10 -c 2 -f -o gpurun_out/${TAG}_full_vocres python bench.py $ARGS \
    > gpurun_out/${TAG}_ncu_full_vocres.log 2>&1
  echo "ncu_vocres_exit=$?"
fi
find gpurun_out -size +40M -print -delete
du -sh gpurun_out
tail -3 gpurun_out/${TAG}_pytest_gpu.log gpurun_out/${TAG}_smoke.log gpurun_out/${TAG}_test_gemm.log 2>/dev/null
cat gpurun_out/${TAG}_bench.json | cut -c1-600
