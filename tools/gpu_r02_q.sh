#!/bin/bash
# round-2 call Q: sweep of the tcgen05 decode form knobs at the 70B / 8B shapes; attention chunk rebalancing
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for big in 40 16 100000; do for ah in 12 24 6 0; do
  echo -n "big_mb=$big l2_ahead=$ah " >> gpurun_out/summary.txt
  WF_TC_BIG_MB=$big WF_TC_L2_AHEAD=$ah timeout 600 $P --shape 70b --batch 16 --ctx 8192 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
done; done
for big in 40 16; do
  echo -n "big_mb=$big " >> gpurun_out/summary.txt
  WF_TC_BIG_MB=$big timeout 600 $P --shape 8b --batch 32 --ctx 2048 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
  echo -n "big_mb=$big " >> gpurun_out/summary.txt
  WF_TC_BIG_MB=$big timeout 600 $P --shape 8b --batch 16 --ctx 2048 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
done
for ah in 12 0; do
  echo -n "l2_ahead=$ah " >> gpurun_out/summary.txt
  WF_TC_L2_AHEAD=$ah timeout 600 $P --shape 8b --batch 64 --ctx 2048 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
