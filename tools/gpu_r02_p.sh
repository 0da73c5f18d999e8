#!/bin/bash
# round-2 call P: L2 prefetch-ahead in the tcgen05 decode form (parity + timing at the 2B / 8B / 70B shapes)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_engine.py tests/test_gpu_gemm_tc.py -m gpu -q --no-header -rf -p no:cacheprovider -k "tcgen05 or batched_equals or gemm_tc or phase_kernels_equal_per_op_kernels_2b" > gpurun_out/test_subset.log 2>&1
echo "tests(subset) exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_subset.log >> gpurun_out/summary.txt
P="python tools/profile_step.py"
for cfg in "2b 64 4096" "2b 40 2048" "8b 16 2048" "8b 32 2048" "8b 64 2048" "70b 16 8192" "70b 32 4096"; do set -- $cfg
  timeout 600 $P --shape $1 --batch $2 --ctx $3 --steps 10 2>&1 | tail -1 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
