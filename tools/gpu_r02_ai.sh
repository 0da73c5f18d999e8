#!/bin/bash
# round-2 call AI: the allocation-free entry point (alone in a fresh process, then at the end of the whole suite in ONE process as the
# driver runs it), prefill GEMM re-measured with the 2-CTA multicast default, bench both arms
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/zz_alone.log 2>&1
echo "zz_alone exit=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/zz_alone.log >> gpurun_out/summary.txt
timeout 3000 python -m pytest tests -x -q -m gpu --no-header -rf -p no:cacheprovider > gpurun_out/suite_one_process.log 2>&1
echo "suite exit=$?" >> gpurun_out/summary.txt; tail -4 gpurun_out/suite_one_process.log >> gpurun_out/summary.txt
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit=$?" >> gpurun_out/summary.txt
for cl in 2 1; do
  WF_TC_CLUSTER=$cl timeout 600 python tools/bench_prefill.py --clocks > gpurun_out/r02ai_prefill_cl$cl.json 2> gpurun_out/prefill_cl$cl.err
  echo "prefill cl=$cl exit=$?" >> gpurun_out/summary.txt; tail -c 600 gpurun_out/r02ai_prefill_cl$cl.json >> gpurun_out/summary.txt; echo >> gpurun_out/summary.txt
done
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/r02ai_bench.json 2> gpurun_out/r02ai_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02ai_bench_ref.json 2> gpurun_out/r02ai_bench_ref.err
echo "bench_ref exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
