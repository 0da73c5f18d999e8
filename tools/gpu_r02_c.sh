#!/bin/bash
# round-2 call C: tests after the tail kernels / rounding change, step timings A/B, prefill knob experiments, ncu launch lists
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_engine.py tests/test_gpu_linear.py tests/test_gpu_gemm_tc.py tests/test_gpu_sglang_model.py tests/test_gpu_kv_int8.py tests/test_gpu_loader.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -s -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$b.log >> gpurun_out/summary.txt
done
grep -h "^config0\|config0 /" gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
for t in 1 0; do
  echo "WF_TAIL=$t" >> gpurun_out/summary.txt
  for bc in "1 512" "4 512" "8 512" "16 2048" "32 2048" "64 4096"; do
    set -- $bc
    WF_TAIL=$t timeout 300 python tools/profile_step.py --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt
  done
done
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 1 --ctx 512 > gpurun_out/r02_trace_tail_bs1.txt 2>&1
bash tools/prefill_experiments.sh > gpurun_out/prefill_experiments.log 2>&1
NCU="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
python tools/profile_step.py --batch 1 --ctx 512 > gpurun_out/plain1.log 2>&1 && $NCU --log-file gpurun_out/r02_launches_bs1_ctx512.csv python tools/profile_step.py --batch 1 --ctx 512 > gpurun_out/ncu1.log 2>&1
python tools/profile_step.py --batch 64 --ctx 4096 > gpurun_out/plain64.log 2>&1 && $NCU --log-file gpurun_out/r02_launches_bs64_ctx4096.csv python tools/profile_step.py --batch 64 --ctx 4096 > gpurun_out/ncu64.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
