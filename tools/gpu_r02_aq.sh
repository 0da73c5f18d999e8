#!/bin/bash
# round-2 call AQ: streaming GEMM unit parity vs the oracle (localise the 8B / 16-row mismatch of call AP)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_linear.py -m gpu -q --no-header -rf -p no:cacheprovider -k "streaming_gemm" > gpurun_out/stream_unit.log 2>&1
echo "stream unit exit=$?" >> gpurun_out/summary.txt; tail -40 gpurun_out/stream_unit.log | cut -c1-220 >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
