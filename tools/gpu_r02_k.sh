#!/bin/bash
# round-2 call K: attention cluster/DSMEM chunk merge for small grids (parity + timing A/B), new batch 2..8 routing
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_engine.py tests/test_gpu_attention.py tests/test_gpu_kv_int8.py tests/test_gpu_sglang_model.py tests/test_gpu_loader.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -4 gpurun_out/$b.log >> gpurun_out/summary.txt
done
P="python tools/profile_step.py"
for cl in 1 0; do
  echo "WF_ATTN_CLUSTER=$cl" >> gpurun_out/summary.txt
  for bc in "1 512" "1 4096" "1 16384" "4 512" "8 2048" "16 2048"; do set -- $bc; WF_ATTN_CLUSTER=$cl timeout 300 $P --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt; done
done
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 1 --ctx 512 > gpurun_out/r02k_trace_tail_bs1.txt 2>&1
tail -6 gpurun_out/r02k_trace_tail_bs1.txt >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
