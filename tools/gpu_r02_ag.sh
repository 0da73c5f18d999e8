#!/bin/bash
# round-2 call AG: two-class grid dispatch order (short CTAs first vs last)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for ord in 1 0 1 0; do
  echo "WF_ATTN_TWO_ORDER=$ord" >> gpurun_out/summary.txt
  for cfg in "2b 64 4096" "2b 64 2048" "2b 64 8192"; do set -- $cfg
    WF_ATTN_TWO_ORDER=$ord timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 15 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
  done
done
WF_ATTN_TWO_ORDER=1 timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "two_class" > gpurun_out/test_two.log 2>&1
echo "two-class parity (short first) exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
