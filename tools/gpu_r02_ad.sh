#!/bin/bash
# round-2 call AD: bandwidth-aware attention chunk model: timings across configs + attention / engine / int8 parity + bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for cfg in "2b 64 4096" "2b 64 2048" "2b 64 1024" "2b 64 8192" "2b 48 4096" "2b 32 2048" "2b 24 4096" "2b 16 2048" "2b 8 8192" "2b 1 16384" "2b 1 512" "8b 32 2048" "8b 16 2048" "8b 64 2048" "70b 16 8192"; do set -- $cfg
  timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 15 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
done
for f in tests/test_gpu_attention.py tests/test_gpu_engine.py tests/test_gpu_kv_int8.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -2 gpurun_out/$b.log >> gpurun_out/summary.txt
done
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02ad_bench.json 2> gpurun_out/r02ad_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
