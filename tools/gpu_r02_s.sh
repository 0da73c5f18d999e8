#!/bin/bash
# round-2 call S: flag-in-data hand-offs in the batch-1 tail kernel (parity, timing A/B, trace)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 1200 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -x -p no:cacheprovider > gpurun_out/test_gpu_engine.log 2>&1
echo "test_gpu_engine exit=$?" >> gpurun_out/summary.txt
tail -4 gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
P="python tools/profile_step.py"
for lam in 1 0; do
  echo "WF_TAIL_LAM=$lam" >> gpurun_out/summary.txt
  for cfg in "2b 1 512" "2b 1 4096" "8b 1 512"; do set -- $cfg
    WF_TAIL_LAM=$lam timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 50 2>&1 | tail -1 >> gpurun_out/summary.txt
  done
done
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 1 --ctx 512 > gpurun_out/r02s_trace_tail_bs1.txt 2>&1
tail -12 gpurun_out/r02s_trace_tail_bs1.txt >> gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_gpu_sglang_model.py tests/test_gpu_loader.py tests/test_gpu_kv_int8.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/test_rest.log 2>&1
echo "other engine users exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_rest.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
