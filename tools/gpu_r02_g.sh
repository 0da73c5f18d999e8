#!/bin/bash
# round-2 call G: what paces the tcgen05 prefill main loop?  (a) SM clock / power under sustained load next to cuBLASLt,
# (b) stage-time experiments (WF_TC_DBG bit mask), (c) ncu --set full of the kernel as shipped
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 300 python tools/bench_prefill.py --clocks --only gate_up > gpurun_out/r02g_prefill_clocks_gate_up.json 2> gpurun_out/r02g_clocks.err
timeout 300 python tools/bench_prefill.py --clocks --only down > gpurun_out/r02g_prefill_clocks_down.json 2>> gpurun_out/r02g_clocks.err
cat gpurun_out/r02g_prefill_clocks_*.json >> gpurun_out/summary.txt
for only in gate_up down; do
  for dbg in 0 1 5 9 13; do
    echo -n "only=$only dbg=$dbg " >> gpurun_out/summary.txt
    WF_TC_DBG=$dbg timeout 120 python tools/bench_prefill.py --only $only --iters 5 2>/dev/null | grep -o '"layer_TOPs": [0-9.]*' >> gpurun_out/summary.txt
  done
done
timeout 300 python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/prefill_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ternary_gemm_tc -s 3 -c 1 -f -o gpurun_out/r02_full_gemm_tc \
    python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/ncu_tc.log 2>&1
echo "ncu gemm_tc exit=$?" >> gpurun_out/summary.txt
WF_TC_DBG=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:ternary_gemm_tc -s 3 -c 1 -f -o gpurun_out/r02_full_gemm_tc_noepi \
    python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/ncu_tc2.log 2>&1
echo "ncu gemm_tc (no epilogue) exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
