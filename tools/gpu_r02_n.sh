#!/bin/bash
# round-2 call N: batched tail kernel at the 8B / 70B shapes (rows of up to 28672 elements), TP = 1
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for tb in 1 0; do
  echo "WF_TAIL_BATCH=$tb WF_DECODE_TC_MIN=9" >> gpurun_out/summary.txt
  for cfg in "8b 16 2048" "8b 32 2048" "8b 64 2048" "70b 16 8192"; do set -- $cfg
    WF_DECODE_TC_MIN=9 WF_TAIL_BATCH=$tb timeout 600 $P --shape $1 --batch $2 --ctx $3 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
  done
done
echo "small batches at 8b (WF_SMALL_MAX 1 vs 8; tail)" >> gpurun_out/summary.txt
for sm in 1 8; do for cfg in "8b 1 512" "8b 2 512" "8b 4 512" "8b 8 512" "70b 1 512" "70b 4 512"; do set -- $cfg
  echo -n "small_max=$sm " >> gpurun_out/summary.txt
  WF_SMALL_MAX=$sm timeout 600 $P --shape $1 --batch $2 --ctx $3 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
done; done
cat gpurun_out/summary.txt
