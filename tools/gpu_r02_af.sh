#!/bin/bash
# round-2 call AF: two-class grid: how full should the short wave be; parity test
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for fill in 8 6 4 3; do
  echo "WF_ATTN_TWO_FILL=$fill" >> gpurun_out/summary.txt
  for cfg in "2b 64 4096" "2b 64 2048" "2b 64 8192"; do set -- $cfg
    WF_ATTN_TWO_FILL=$fill timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 15 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
  done
done
timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "two_class or even_split" > gpurun_out/test_two.log 2>&1
echo "two-class parity exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_two.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
