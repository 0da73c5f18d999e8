#!/bin/bash
# round-2 call AK: int8-KV attention chunk counts (the bf16-calibrated wave model regressed it: 73.7 us/layer in round 1, 84 us now)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py --kv int8 --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "64 4096" "64 2048" "32 2048" "48 4096" "16 2048"; do set -- $cfg; B=$1; C=$2
  run WF_X=0
  run WF_ATTN_WANT=4
  for n in 1 2 3 4 6 8; do run WF_ATTN_TWO=0 WF_ATTN_NCHUNKS=$n; done
done
cat gpurun_out/summary.txt
