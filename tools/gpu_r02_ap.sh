#!/bin/bash
# round-2 call AP: streaming batched GEMM (gemm_stream.cu): parity, then step times with WF_DECODE_STREAM = 0 / 1 / 2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -x -k "streaming_decode" > gpurun_out/stream_test.log 2>&1
echo "stream test exit=$?" >> gpurun_out/summary.txt; tail -12 gpurun_out/stream_test.log | cut -c1-200 >> gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "70b 16 8192" "8b 32 2048" "8b 16 2048" "8b 8 2048" "8b 4 512" "2b 32 2048" "2b 16 2048" "2b 8 512" "2b 4 512" "2b 2 512"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_DECODE_STREAM=0
  run WF_DECODE_STREAM=1
  run WF_DECODE_STREAM=2
done
cat gpurun_out/summary.txt
