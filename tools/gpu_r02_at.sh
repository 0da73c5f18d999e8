#!/bin/bash
# round-2 call AT: streaming GEMM with the 16-warp two-pass form for 17..32 rows: parity, then step times (0 = off, 1 = up to 16 rows, 2 = up to 32)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_linear.py tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "streaming" > gpurun_out/stream_tests.log 2>&1
echo "stream tests exit=$?" >> gpurun_out/summary.txt; tail -5 gpurun_out/stream_tests.log | cut -c1-220 >> gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "8b 32 2048" "8b 24 2048" "2b 32 2048" "2b 24 4096" "70b 32 4096" "70b 16 8192" "8b 16 2048" "2b 8 512"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_DECODE_STREAM=0
  run WF_DECODE_STREAM=1
  run WF_DECODE_STREAM=2
done
cat gpurun_out/summary.txt
