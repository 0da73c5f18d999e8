#!/bin/bash
# round-2 call A: full GPU parity suite (one process per file), bench (both arms), prefill record, ncu --set full of the tcgen05 GEMM
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
RUN_BENCH=0 TEST_TIMEOUT=900 bash tools/gpu_ci.sh > gpurun_out/ci.log 2>&1
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02a_bench_ref.json 2> gpurun_out/r02a_bench_ref.err
echo "bench_ref exit=$?" >> gpurun_out/summary.txt
timeout 300 python tools/bench_prefill.py > gpurun_out/r02a_prefill.json 2> gpurun_out/r02a_prefill.err
echo "prefill exit=$?" >> gpurun_out/summary.txt
timeout 300 python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/r02a_prefill_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ternary_gemm_tc -s 3 -c 1 -f -o gpurun_out/r02_full_gemm_tc \
    python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/r02a_ncu.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
