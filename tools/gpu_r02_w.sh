#!/bin/bash
# round-2 call W: flag-in-data tail (arming off the epilogue path): engine parity, timing A/B, trace
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 1500 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/test_gpu_engine.log 2>&1
echo "test_gpu_engine exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
P="python tools/profile_step.py"
for lam in 1 0 1 0; do
  echo "WF_TAIL_LAM=$lam" >> gpurun_out/summary.txt
  for cfg in "2b 1 512" "2b 1 4096"; do set -- $cfg
    WF_TAIL_LAM=$lam timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 50 2>&1 | tail -1 | cut -c1-150 >> gpurun_out/summary.txt
  done
done
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 1 --ctx 512 > gpurun_out/r02w_trace_tail_bs1.txt 2>&1
tail -12 gpurun_out/r02w_trace_tail_bs1.txt >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
