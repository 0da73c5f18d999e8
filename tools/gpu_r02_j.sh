#!/bin/bash
# round-2 call J: per-CTA-flag grid barrier in the tail kernels (parity + bs1 timing + trace), batch 2..8 routing, attention dispatch order
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_engine.py tests/test_gpu_attention.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -4 gpurun_out/$b.log >> gpurun_out/summary.txt
done
P="python tools/profile_step.py"
echo "default routing" >> gpurun_out/summary.txt
for bc in "1 512" "2 512" "4 512" "8 512" "1 4096"; do set -- $bc; timeout 300 $P --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt; done
echo "WF_SMALL_MAX=1" >> gpurun_out/summary.txt
for bc in "2 512" "4 512" "8 512"; do set -- $bc; WF_SMALL_MAX=1 timeout 300 $P --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt; done
echo "WF_SMALL_MAX=1 WF_TAIL_BATCH=1" >> gpurun_out/summary.txt
for bc in "2 512" "4 512" "8 512" "64 4096"; do set -- $bc; WF_SMALL_MAX=1 WF_TAIL_BATCH=1 timeout 300 $P --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt; done
for cm in 1 0; do for want in 4 6 8; do
  echo -n "chunk_major=$cm want=$want " >> gpurun_out/summary.txt
  WF_ATTN_CHUNK_MAJOR=$cm WF_ATTN_WANT=$want timeout 300 $P --batch 64 --ctx 4096 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt
done; done
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 1 --ctx 512 > gpurun_out/r02j_trace_tail_bs1.txt 2>&1
cat gpurun_out/summary.txt
