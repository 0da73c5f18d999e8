#!/bin/bash
# round-2 call AL: int8-KV attention after (a) the even form / experiment orders compiled out of the int8 kernel (spills 80 B -> 4 B),
# (b) the chunk model with R = slots for the int8 form
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_kv_int8.py tests/test_gpu_attention.py tests/test_gpu_engine.py; do
  b=$(basename $f .py)
  timeout 900 python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt; tail -2 gpurun_out/$b.log >> gpurun_out/summary.txt
done
P="python tools/profile_step.py --steps 20"
for cfg in "2b 64 4096" "2b 64 2048" "2b 32 2048" "2b 48 4096" "2b 16 2048" "2b 24 4096" "2b 8 8192" "8b 32 2048" "8b 16 2048" "70b 16 8192"; do set -- $cfg
  timeout 600 $P --kv int8 --shape $1 --batch $2 --ctx $3 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt
done
for cfg in "2b 64 4096" "2b 32 2048"; do set -- $cfg
  timeout 600 $P --shape $1 --batch $2 --ctx $3 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
