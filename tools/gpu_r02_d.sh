#!/bin/bash
# round-2 call D (first call after the container was re-created): full GPU parity suite, step-time A/B of the
# persistent tail kernels, bench (both arms), prefill record, launch lists, bs1 trace
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
RUN_BENCH=0 TEST_TIMEOUT=1500 bash tools/gpu_ci.sh > gpurun_out/ci.log 2>&1
grep -h "^config0\|config0 /" gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
for t in 1 0; do
  echo "WF_TAIL=$t" >> gpurun_out/summary.txt
  for bc in "1 512" "4 512" "8 512" "16 2048" "32 2048" "64 4096"; do
    set -- $bc
    WF_TAIL=$t timeout 300 python tools/profile_step.py --batch $1 --ctx $2 --steps 30 2>&1 | tail -1 >> gpurun_out/summary.txt
  done
done
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/r02d_bench.json 2> gpurun_out/r02d_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02d_bench_ref.json 2> gpurun_out/r02d_bench_ref.err
echo "bench_ref exit=$?" >> gpurun_out/summary.txt
timeout 300 python tools/bench_prefill.py > gpurun_out/r02d_prefill.json 2> gpurun_out/r02d_prefill.err
echo "prefill exit=$?" >> gpurun_out/summary.txt
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 1 --ctx 512 > gpurun_out/r02_trace_tail_bs1.txt 2>&1
WF_TRACE=1 timeout 120 python tools/trace_tail.py --batch 64 --ctx 4096 > gpurun_out/r02_trace_tail_bs64.txt 2>&1
NCU="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
python tools/profile_step.py --batch 1 --ctx 512 > gpurun_out/plain1.log 2>&1 && $NCU --log-file gpurun_out/r02_launches_bs1_ctx512.csv python tools/profile_step.py --batch 1 --ctx 512 > gpurun_out/ncu1.log 2>&1
python tools/profile_step.py --batch 64 --ctx 4096 > gpurun_out/plain64.log 2>&1 && $NCU --log-file gpurun_out/r02_launches_bs64_ctx4096.csv python tools/profile_step.py --batch 64 --ctx 4096 > gpurun_out/ncu64.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
