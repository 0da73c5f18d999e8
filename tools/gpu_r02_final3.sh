#!/bin/bash
# round-2 final evidence on the last tree: whole GPU suite in ONE process (as the driver runs it), smoke, bench (both arms),
# launch lists + ncu --set full of the attention kernel at bs64/4K, step times of the other shapes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 3000 python -m pytest tests -x -q -m gpu --no-header -rf -p no:cacheprovider > gpurun_out/suite_one_process.log 2>&1
echo "suite exit=$?" >> gpurun_out/summary.txt; tail -4 gpurun_out/suite_one_process.log >> gpurun_out/summary.txt
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit=$?" >> gpurun_out/summary.txt
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_final3_bench.json 2> gpurun_out/r02_final3_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_final3_bench_ref.json 2> gpurun_out/r02_final3_bench_ref.err
echo "bench_ref exit=$?" >> gpurun_out/summary.txt
P="python tools/profile_step.py"
L="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
$P --batch 64 --ctx 4096 > gpurun_out/plain64.log 2>&1 && $L --log-file gpurun_out/r02h_launches_bs64_ctx4096.csv $P --batch 64 --ctx 4096 > gpurun_out/ncu64.log 2>&1
$P --batch 1 --ctx 512 > gpurun_out/plain1.log 2>&1 && $L --log-file gpurun_out/r02h_launches_bs1_ctx512.csv $P --batch 1 --ctx 512 > gpurun_out/ncu1.log 2>&1
F="ncu --profile-from-start off --clock-control none --set full --import-source on"
timeout 600 $F -k regex:paged_decode_attn -s 5 -c 1 -f -o gpurun_out/r02h_full_attn_bs64 $P --batch 64 --ctx 4096 > gpurun_out/n3.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
for cfg in "2b 1 512 bf16" "2b 16 2048 bf16" "2b 32 2048 bf16" "2b 64 4096 bf16" "2b 64 4096 int8" "2b 48 4096 int8" "8b 1 512 bf16" "8b 32 2048 bf16" "70b 16 8192 bf16"; do set -- $cfg
  timeout 600 $P --shape $1 --batch $2 --ctx $3 --kv $4 --steps 20 2>&1 | tail -1 | cut -c1-130 >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
