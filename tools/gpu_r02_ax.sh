#!/bin/bash
# round-2 call AX: row glue as a cluster of CTAs per row (long rows, small batch)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "70b 16 8192" "70b 8 4096" "70b 32 4096"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_GLUE_CLUSTER=0
  run WF_GLUE_CLUSTER=1
done
for cfg in "8b 16 2048" "8b 32 2048" "2b 64 4096" "2b 16 2048"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_GLUE_CLUSTER=0
  run WF_GLUE_CLUSTER=2
  run WF_GLUE_CLUSTER=4
done
WF_GLUE_CLUSTER=2 timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/engine_cl2.log 2>&1
echo "engine tests with WF_GLUE_CLUSTER=2 exit=$?" >> gpurun_out/summary.txt; grep -E "^FAILED|passed|failed" gpurun_out/engine_cl2.log | cut -c1-200 >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
