#!/bin/bash
# round-2 call BC: streaming GEMM at 17..24 rows (two-pass 16-warp form) vs the default forms
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "70b 24 4096" "70b 20 4096" "8b 20 2048" "8b 24 2048" "2b 20 2048" "2b 24 2048"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_DECODE_STREAM_MAX=16
  run WF_DECODE_STREAM_MAX=24
done
cat gpurun_out/summary.txt
