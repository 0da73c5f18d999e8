#!/bin/bash
# round-2 call Y: new hand-off parity test, forced cluster merge on the large attention grid
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "handoff or layer_tail" > gpurun_out/test_gpu_engine.log 2>&1
echo "test_gpu_engine (tail tests) exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
P="python tools/profile_step.py"
for cl in 1 2 1 2; do
  echo "WF_ATTN_CLUSTER=$cl" >> gpurun_out/summary.txt
  for cfg in "2b 64 4096" "2b 32 2048" "8b 32 2048"; do set -- $cfg
    WF_ATTN_CLUSTER=$cl timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 20 2>&1 | tail -1 | cut -c1-150 >> gpurun_out/summary.txt
  done
done
cat gpurun_out/summary.txt
