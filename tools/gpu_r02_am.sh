#!/bin/bash
# round-2 call AM: per-sequence (adaptive) chunking vs fixed chunks of the hint; int8 model (R = slots, C0 = 30, no old-rule tie)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_kv_int8.py tests/test_gpu_attention.py tests/test_gpu_engine.py; do
  b=$(basename $f .py)
  timeout 900 python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt; tail -2 gpurun_out/$b.log >> gpurun_out/summary.txt
done
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --kv $KV --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for KV in int8 bf16; do
for cfg in "2b 64 4096" "2b 64 2048" "2b 32 2048" "2b 48 4096" "2b 48 3000" "2b 24 4096" "8b 32 2048" "8b 16 2048"; do set -- $cfg; S=$1; B=$2; C=$3
  run WF_ATTN_ADAPT=1
  run WF_ATTN_ADAPT=0
done
done
KV=int8
for cfg in "2b 32 2048" "2b 32 4000" "2b 24 4000"; do set -- $cfg; S=$1; B=$2; C=$3
  for n in 1 2 3 4; do run WF_ATTN_TWO=0 WF_ATTN_NCHUNKS=$n; done
done
cat gpurun_out/summary.txt
