#!/bin/bash
# round-2 call AZ: batch 1..4 at the 8B / 70B widths: persistent tail kernel vs phase kernels (GEMV-style) vs row glue + streaming GEMM
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "70b 1 2048" "70b 2 2048" "8b 1 512" "8b 2 512" "2b 1 512"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_X=0
  run WF_TAIL=0
  run WF_TAIL=0 WF_SMALL_MAX=0
  run WF_TAIL=0 WF_SMALL_MAX=8
done
cat gpurun_out/summary.txt
