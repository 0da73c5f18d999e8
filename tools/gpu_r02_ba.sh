#!/bin/bash
# round-2 call BA: batch 1 at the 8B / 70B widths through the row glue + streaming GEMM by default: engine tests, step times
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 1200 python -m pytest tests/test_gpu_engine.py tests/test_gpu_sglang_model.py tests/test_gpu_loader.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/engine_tests.log 2>&1
echo "engine/model/loader tests exit=$?" >> gpurun_out/summary.txt; grep -E "^FAILED|passed|failed" gpurun_out/engine_tests.log | cut -c1-200 | head >> gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
for cfg in "8b 1 512" "8b 1 4096" "70b 1 2048" "2b 1 512"; do set -- $cfg
  timeout 300 $P --shape $1 --batch $2 --ctx $3 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
