#!/bin/bash
# round-2 call H: tcgen05 prefill GEMM with the single-thread (elect.sync) issue loop: parity, knob sweep, record
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_gemm_tc.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/test_gpu_gemm_tc.log 2>&1
echo "test_gpu_gemm_tc exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_gpu_gemm_tc.log >> gpurun_out/summary.txt
for only in gate_up down; do
  for cfg in "0 2 1" "0 3 1" "0 2 2" "0 3 2" "1 2 1" "1 3 1" "1 2 2" "1 3 2" "13 2 1" "5 2 1" "9 2 1"; do
    set -- $cfg
    echo -n "only=$only dbg=$1 nu=$2 cl=$3 " >> gpurun_out/summary.txt
    WF_TC_DBG=$1 WF_TC_NU=$2 WF_TC_CLUSTER=$3 timeout 120 python tools/bench_prefill.py --only $only --iters 5 2>/dev/null | grep -o '"layer_TOPs": [0-9.]*' >> gpurun_out/summary.txt
  done
done
for cl in 1 2; do
  WF_TC_CLUSTER=$cl timeout 300 python tools/bench_prefill.py > gpurun_out/r02h_prefill_cl$cl.json 2> gpurun_out/r02h_prefill_cl$cl.err
  echo "prefill cl=$cl exit=$?" >> gpurun_out/summary.txt
done
grep -h -o '"layer_TOPs[^}]*' gpurun_out/r02h_prefill_cl*.json >> gpurun_out/summary.txt
timeout 300 python tools/bench_prefill.py --clocks --only gate_up > gpurun_out/r02h_prefill_clocks_gate_up.json 2> gpurun_out/r02h_clocks.err
cat gpurun_out/r02h_prefill_clocks_gate_up.json >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
