#!/bin/bash
# round-2 call AR: localise the 8B / 16-row mismatch of the streaming GEMM inside the engine (PDL on / off, repeated), unit test with the wide matrices
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_linear.py -m gpu -q --no-header -rf -p no:cacheprovider -k "streaming_gemm" > gpurun_out/stream_unit.log 2>&1
echo "stream unit exit=$?" >> gpurun_out/summary.txt; tail -6 gpurun_out/stream_unit.log | cut -c1-220 >> gpurun_out/summary.txt
for pdl in 1 0; do for rep in 1 2; do
  WF_PDL=$pdl timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "streaming_decode" > gpurun_out/stream_engine_pdl${pdl}_$rep.log 2>&1
  echo "engine pdl=$pdl rep=$rep exit=$?" >> gpurun_out/summary.txt; grep -E "^FAILED|passed|failed" gpurun_out/stream_engine_pdl${pdl}_$rep.log | cut -c1-200 >> gpurun_out/summary.txt
done; done
cat gpurun_out/summary.txt
