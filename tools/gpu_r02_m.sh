#!/bin/bash
# round-2 call M: batched decode linears at the 8B / 70B shapes (TP = 1): one-shot mma.sync GEMM vs tcgen05 split-K form vs the batched tail kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "tcgen05 or batched_equals" > gpurun_out/test_gpu_engine.log 2>&1
echo "test_gpu_engine(subset) exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
P="python tools/profile_step.py"
for min in 33 9; do
  echo "WF_DECODE_TC_MIN=$min" >> gpurun_out/summary.txt
  for cfg in "2b 16 2048" "2b 32 2048" "8b 16 2048" "8b 32 2048" "70b 16 8192"; do set -- $cfg
    WF_DECODE_TC_MIN=$min timeout 600 $P --shape $1 --batch $2 --ctx $3 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
  done
done
echo "WF_TAIL_BATCH=1 (2b only fits)" >> gpurun_out/summary.txt
for cfg in "2b 16 2048" "2b 32 2048"; do set -- $cfg
  WF_TAIL_BATCH=1 timeout 600 $P --shape $1 --batch $2 --ctx $3 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
