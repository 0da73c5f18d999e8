#!/bin/bash
# round-2 final evidence (tree with the two-class attention grid): full GPU suite, bench (both arms), launch list + ncu of the attention kernel at bs64/4K
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
RUN_BENCH=0 TEST_TIMEOUT=1500 bash tools/gpu_ci.sh > gpurun_out/ci.log 2>&1
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_final2_bench.json 2> gpurun_out/r02_final2_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_final2_bench_ref.json 2> gpurun_out/r02_final2_bench_ref.err
echo "bench_ref exit=$?" >> gpurun_out/summary.txt
P="python tools/profile_step.py"
L="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
$P --batch 64 --ctx 4096 > gpurun_out/plain64.log 2>&1 && $L --log-file gpurun_out/r02g_launches_bs64_ctx4096.csv $P --batch 64 --ctx 4096 > gpurun_out/ncu64.log 2>&1
$P --batch 1 --ctx 512 > gpurun_out/plain1.log 2>&1 && $L --log-file gpurun_out/r02g_launches_bs1_ctx512.csv $P --batch 1 --ctx 512 > gpurun_out/ncu1.log 2>&1
F="ncu --profile-from-start off --clock-control none --set full --import-source on"
timeout 600 $F -k regex:paged_decode_attn -s 5 -c 1 -f -o gpurun_out/r02g_full_attn_twoclass_bs64 $P --batch 64 --ctx 4096 > gpurun_out/n3.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
for cfg in "2b 1 512" "2b 16 2048" "2b 32 2048" "2b 64 4096" "8b 1 512" "8b 32 2048" "70b 16 8192"; do set -- $cfg
  timeout 600 $P --shape $1 --batch $2 --ctx $3 --steps 20 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
