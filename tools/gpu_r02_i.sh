#!/bin/bash
# round-2 call I: attention even work split (parity + step timing A/B), prefill GEMM with the lean unpack loop
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_attention.py tests/test_gpu_engine.py tests/test_gpu_kv_int8.py tests/test_gpu_gemm_tc.py tests/test_gpu_sglang_model.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -x -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -4 gpurun_out/$b.log >> gpurun_out/summary.txt
done
for ev in 1 0; do
  echo "WF_ATTN_EVEN=$ev" >> gpurun_out/summary.txt
  for bc in "64 4096" "32 2048" "16 2048" "64 1024" "8 8192"; do
    set -- $bc
    WF_ATTN_EVEN=$ev timeout 300 python tools/profile_step.py --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt
  done
done
for cl in 1 2; do
  WF_TC_CLUSTER=$cl timeout 300 python tools/bench_prefill.py > gpurun_out/r02i_prefill_cl$cl.json 2> gpurun_out/r02i_prefill_cl$cl.err
  echo "prefill cl=$cl exit=$?" >> gpurun_out/summary.txt
done
grep -h -o '"layer_TOPs[^}]*' gpurun_out/r02i_prefill_cl*.json >> gpurun_out/summary.txt
for dbg in 1 5 13; do
  echo -n "gate_up dbg=$dbg " >> gpurun_out/summary.txt
  WF_TC_DBG=$dbg timeout 120 python tools/bench_prefill.py --only gate_up --iters 5 2>/dev/null | grep -o '"layer_TOPs": [0-9.]*' >> gpurun_out/summary.txt
done
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
