#!/bin/bash
# round-2 call AJ: feasibility of overlapping sub-batches (attention of one under the linears of another)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for parts in 1 2 4; do
  timeout 300 python tools/exp_overlap.py --parts $parts >> gpurun_out/summary.txt 2>&1
done
timeout 300 python tools/exp_overlap.py --parts 2 --skew-us 100 >> gpurun_out/summary.txt 2>&1
timeout 300 python tools/exp_overlap.py --parts 3 --batch 63 >> gpurun_out/summary.txt 2>&1
timeout 300 python tools/exp_overlap.py --parts 2 --ctx 2048 >> gpurun_out/summary.txt 2>&1
timeout 300 python tools/exp_overlap.py --parts 1 --ctx 2048 >> gpurun_out/summary.txt 2>&1
cat gpurun_out/summary.txt
