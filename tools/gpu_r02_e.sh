#!/bin/bash
# round-2 call E: ncu --set full captures of the round-2 kernels (tail, batched tail, tcgen05 prefill GEMM, fused attention),
# the prefill knob experiments, compute-sanitizer memcheck
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
NCU="ncu --profile-from-start off --clock-control none --set full --import-source on"
$P --batch 1 --ctx 512 > gpurun_out/plain1.log 2>&1 &&
timeout 600 $NCU -k regex:layer_tail_kernel -s 5 -c 1 -f -o gpurun_out/r02_full_layer_tail_bs1 $P --batch 1 --ctx 512 > gpurun_out/ncu_tail1.log 2>&1
echo "ncu tail bs1 exit=$?" >> gpurun_out/summary.txt
WF_TAIL_BATCH=1 $P --batch 64 --ctx 4096 > gpurun_out/plain64.log 2>&1 &&
WF_TAIL_BATCH=1 timeout 600 $NCU -k regex:layer_tail_batch -s 5 -c 1 -f -o gpurun_out/r02_full_layer_tail_batch_bs64 $P --batch 64 --ctx 4096 > gpurun_out/ncu_tail64.log 2>&1
echo "ncu tail bs64 exit=$?" >> gpurun_out/summary.txt
timeout 600 $NCU -k regex:paged_decode_attn -s 5 -c 1 -f -o gpurun_out/r02_full_attn_fused_bs64 $P --batch 64 --ctx 4096 > gpurun_out/ncu_attn64.log 2>&1
echo "ncu attn bs64 exit=$?" >> gpurun_out/summary.txt
timeout 600 $NCU -k regex:paged_decode_attn -s 5 -c 1 -f -o gpurun_out/r02_full_attn_fused_bs1 $P --batch 1 --ctx 512 > gpurun_out/ncu_attn1.log 2>&1
echo "ncu attn bs1 exit=$?" >> gpurun_out/summary.txt
timeout 300 python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/prefill_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ternary_gemm_tc -s 3 -c 1 -f -o gpurun_out/r02_full_gemm_tc \
    python tools/bench_prefill.py --only gate_up --iters 2 > gpurun_out/ncu_tc.log 2>&1
echo "ncu gemm_tc exit=$?" >> gpurun_out/summary.txt
bash tools/prefill_experiments.sh > gpurun_out/prefill_experiments.log 2>&1
echo "prefill experiments exit=$?" >> gpurun_out/summary.txt
timeout 300 python tools/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 &&
timeout 1200 compute-sanitizer --tool memcheck --log-file gpurun_out/r02_sanitizer_memcheck.log python tools/sanitize_target.py > gpurun_out/sanitize_memcheck.out 2>&1
echo "memcheck exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/r02_sanitizer_memcheck.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
