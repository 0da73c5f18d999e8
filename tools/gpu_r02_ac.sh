#!/bin/bash
# round-2 call AC: more forced chunk counts (is one chunk per pair right whenever pairs fill half the slots?)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for cfg in "2b 64 4096 1" "2b 64 4096 2" "2b 64 4096 4" "70b 16 8192 1" "2b 64 2048 1" "2b 64 2048 2" "2b 64 2048 4" "2b 48 4096 1" "2b 48 4096 2" "2b 48 4096 4" "2b 24 4096 1" "2b 24 4096 2" "2b 24 4096 3" "2b 64 8192 1" "2b 64 8192 4" "2b 64 8192 8"; do set -- $cfg
  echo -n "nchunks=$4 " >> gpurun_out/summary.txt
  WF_ATTN_NCHUNKS=$4 timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 15 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
