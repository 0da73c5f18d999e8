#!/bin/bash
# round-2 call AE: two-class attention grid (whole pairs fill the slots, leftover pairs cut fine): parity + timing A/B
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for two in 1 0 1 0; do
  echo "WF_ATTN_TWO=$two" >> gpurun_out/summary.txt
  for cfg in "2b 64 4096" "2b 64 2048" "2b 64 8192" "2b 62 4096"; do set -- $cfg
    WF_ATTN_TWO=$two timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 15 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
  done
done
for f in tests/test_gpu_engine.py tests/test_gpu_kv_int8.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -2 gpurun_out/$b.log >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
