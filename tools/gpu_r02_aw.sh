#!/bin/bash
# round-2 call AW: streaming GEMM with the conflict-free shared tile and rotating RED duty
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_linear.py tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "streaming" > gpurun_out/stream_tests.log 2>&1
echo "stream tests exit=$?" >> gpurun_out/summary.txt; tail -5 gpurun_out/stream_tests.log | cut -c1-220 >> gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "8b 16 2048" "8b 8 2048" "8b 4 512" "70b 16 8192" "70b 8 4096" "2b 16 2048" "2b 8 512" "2b 2 512" "8b 24 2048"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_DECODE_STREAM=2
done
cat gpurun_out/summary.txt
