#!/bin/bash
# round-2 final evidence: full GPU suite, bench (both arms), launch lists, ncu --set full of the round-2 kernels as shipped
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
RUN_BENCH=0 TEST_TIMEOUT=1500 bash tools/gpu_ci.sh > gpurun_out/ci.log 2>&1
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err
echo "bench exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_final_bench_ref.json 2> gpurun_out/r02_final_bench_ref.err
echo "bench_ref exit=$?" >> gpurun_out/summary.txt
P="python tools/profile_step.py"
L="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
$P --batch 1 --ctx 512 > gpurun_out/plain1.log 2>&1 && $L --log-file gpurun_out/r02f_launches_bs1_ctx512.csv $P --batch 1 --ctx 512 > gpurun_out/ncu1.log 2>&1
$P --batch 64 --ctx 4096 > gpurun_out/plain64.log 2>&1 && $L --log-file gpurun_out/r02f_launches_bs64_ctx4096.csv $P --batch 64 --ctx 4096 > gpurun_out/ncu64.log 2>&1
echo "launch lists exit=$?" >> gpurun_out/summary.txt
F="ncu --profile-from-start off --clock-control none --set full --import-source on"
timeout 600 $F -k regex:layer_tail_kernel -s 5 -c 1 -f -o gpurun_out/r02f_full_layer_tail_handoff_bs1 $P --batch 1 --ctx 512 > gpurun_out/n1.log 2>&1
timeout 600 $F -k regex:paged_decode_attn -s 5 -c 1 -f -o gpurun_out/r02f_full_attn_cluster_bs1 $P --batch 1 --ctx 512 > gpurun_out/n2.log 2>&1
timeout 600 $F -k regex:paged_decode_attn -s 5 -c 1 -f -o gpurun_out/r02f_full_attn_bs64 $P --batch 64 --ctx 4096 > gpurun_out/n3.log 2>&1
timeout 600 $F -k regex:lm_head_tc -c 1 -f -o gpurun_out/r02f_full_lm_head_tc_bs64 $P --batch 64 --ctx 4096 > gpurun_out/n4.log 2>&1
timeout 600 $F -k regex:ternary_gemm_tc -s 9 -c 1 -f -o gpurun_out/r02f_full_gemm_tc_decode_bs64 $P --batch 64 --ctx 4096 > gpurun_out/n5.log 2>&1
echo "ncu decode captures exit=$?" >> gpurun_out/summary.txt
timeout 300 python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/prefill_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ternary_gemm_tc -s 3 -c 1 -f -o gpurun_out/r02f_full_gemm_tc_prefill \
    python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/n6.log 2>&1
echo "ncu prefill exit=$?" >> gpurun_out/summary.txt
timeout 300 python tools/bench_lm_head.py > gpurun_out/r02f_lm_head.txt 2>&1
cat gpurun_out/r02f_lm_head.txt >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
