#!/bin/bash
# round-2 call AY: one-warp-per-unit streaming GEMM (v2): parity (default selection and forced on every shape), step times v2 off / default / forced
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for v in 12 1; do
  WF_GS_V2=$v timeout 900 python -m pytest tests/test_gpu_linear.py tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "streaming" > gpurun_out/stream_tests_v$v.log 2>&1
  echo "stream tests WF_GS_V2=$v exit=$?" >> gpurun_out/summary.txt; grep -E "^FAILED|passed|failed" gpurun_out/stream_tests_v$v.log | cut -c1-200 | head -8 >> gpurun_out/summary.txt
done
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "70b 16 8192" "70b 8 4096" "8b 16 2048" "8b 8 2048" "8b 4 512" "2b 16 2048" "2b 8 512"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_GS_V2=0
  run WF_GS_V2=12
  run WF_GS_V2=4
done
cat gpurun_out/summary.txt
