#!/bin/bash
# round-2 call R: full GPU suite on the current tree + step timings + bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
RUN_BENCH=0 TEST_TIMEOUT=1500 bash tools/gpu_ci.sh > gpurun_out/ci.log 2>&1
P="python tools/profile_step.py"
for cfg in "2b 1 512" "2b 16 2048" "2b 32 2048" "2b 64 4096" "8b 1 512" "8b 16 2048" "8b 32 2048" "70b 16 8192"; do set -- $cfg
  timeout 600 $P --shape $1 --batch $2 --ctx $3 --steps 20 2>&1 | tail -1 >> gpurun_out/summary.txt
done
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02r_bench.json 2> gpurun_out/r02r_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
