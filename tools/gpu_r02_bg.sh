#!/bin/bash
# repeat the single-GPU streaming GEMM parity cases (race hunting)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
fails=0
for i in $(seq 1 10); do
  timeout 300 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -p no:cacheprovider -k "streaming_decode" > gpurun_out/rep_$i.log 2>&1 || { fails=$((fails+1)); grep -E "^FAILED" gpurun_out/rep_$i.log >> gpurun_out/summary.txt; }
done
echo "engine streaming parity: $fails failing runs of 10" >> gpurun_out/summary.txt
fails=0
for i in 1 2 3; do
  timeout 300 python -m pytest tests/test_gpu_linear.py -m gpu -q --no-header -p no:cacheprovider -k "streaming_gemm" > gpurun_out/repu_$i.log 2>&1 || { fails=$((fails+1)); grep -E "^FAILED" gpurun_out/repu_$i.log | head -3 >> gpurun_out/summary.txt; }
done
echo "unit streaming parity: $fails failing runs of 3" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
