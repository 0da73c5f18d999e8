#!/bin/bash
# round-2 call F: tcgen05 prefill GEMM with the 8-warp epilogue: parity tests, knob sweep (unpack groups, cluster multicast), full record
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_gemm_tc.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/test_gpu_gemm_tc.log 2>&1
echo "test_gpu_gemm_tc exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_gpu_gemm_tc.log >> gpurun_out/summary.txt
out=gpurun_out/r02f_prefill_experiments.jsonl
: > $out
for only in gate_up down; do
  for cfg in "0 2 1" "0 3 1" "0 2 2" "0 3 2" "0 2 4" "1 2 1" "1 2 2" "1 2 4"; do
    set -- $cfg
    echo "{\"only\": \"$only\", \"WF_TC_DBG\": $1, \"WF_TC_NU\": $2, \"WF_TC_CLUSTER\": $3, \"result\": " >> $out
    WF_TC_DBG=$1 WF_TC_NU=$2 WF_TC_CLUSTER=$3 timeout 120 python tools/bench_prefill.py --only $only --iters 5 >> $out 2>/dev/null
    echo "}" >> $out
  done
done
for cl in 1 2 4; do
  WF_TC_CLUSTER=$cl timeout 300 python tools/bench_prefill.py > gpurun_out/r02f_prefill_cl$cl.json 2> gpurun_out/r02f_prefill_cl$cl.err
  echo "prefill cl=$cl exit=$?" >> gpurun_out/summary.txt
done
cat $out | tr -d '\n' | sed 's/}{/}\n{/g' | grep -o '"only[^}]*TOPs": [0-9.]*' >> gpurun_out/summary.txt
grep -h -o '"layer_TOPs[^}]*' gpurun_out/r02f_prefill_cl*.json >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
