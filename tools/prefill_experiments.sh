#!/bin/bash
# prefill tcgen05 GEMM: where does the time go?  knobs of gemm_tc.cu (WF_TC_DBG: 1 = epilogue skips math/stores, 2 = also the TMEM loads;
# WF_TC_NU = unpack warp groups; WF_TC_CLUSTER = 2: activation tile multicast across a CTA pair)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/r02_prefill_experiments.jsonl
: > $out
for only in gate_up down; do
  for cfg in "0 3 1" "1 3 1" "2 3 1" "0 2 1" "0 1 1" "0 3 2" "0 2 2"; do
    set -- $cfg
    echo "{\"only\": \"$only\", \"WF_TC_DBG\": $1, \"WF_TC_NU\": $2, \"WF_TC_CLUSTER\": $3, \"result\": " >> $out
    WF_TC_DBG=$1 WF_TC_NU=$2 WF_TC_CLUSTER=$3 timeout 120 python tools/bench_prefill.py --only $only --iters 5 >> $out 2>/dev/null
    echo "}" >> $out
  done
done
cat $out | tr -d '\n' | sed 's/}{/}\n{/g'
