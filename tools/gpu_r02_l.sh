#!/bin/bash
# round-2 call L: engine parity after the path-switch fix, cluster merge v2 timings, tcgen05 decode linears at batch 64, bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_engine.py tests/test_gpu_kv_int8.py tests/test_gpu_sglang_model.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -4 gpurun_out/$b.log >> gpurun_out/summary.txt
done
P="python tools/profile_step.py"
for bc in "1 512" "1 4096" "1 16384" "2 512" "4 512" "8 2048" "16 2048" "32 2048" "64 4096"; do set -- $bc; timeout 300 $P --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt; done
echo "WF_DECODE_TC=1" >> gpurun_out/summary.txt
for bc in "40 2048" "64 4096"; do set -- $bc; WF_DECODE_TC=1 timeout 300 $P --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt; done
echo "WF_DECODE_TC=0" >> gpurun_out/summary.txt
for bc in "40 2048"; do set -- $bc; timeout 300 $P --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt; done
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 1 --ctx 512 > gpurun_out/r02l_trace_tail_bs1.txt 2>&1
tail -6 gpurun_out/r02l_trace_tail_bs1.txt >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02l_bench.json 2> gpurun_out/r02l_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
