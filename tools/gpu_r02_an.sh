#!/bin/bash
# round-2 call AN: new per-sequence-chunk parity test + ncu --set full of the int8 attention kernel at bs64/4K
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -p no:cacheprovider -k "per_sequence or two_class" > gpurun_out/adapt_test.log 2>&1
echo "adapt test exit=$?" >> gpurun_out/summary.txt; tail -2 gpurun_out/adapt_test.log >> gpurun_out/summary.txt
P="python tools/profile_step.py --kv int8 --batch 64 --ctx 4096"
$P > gpurun_out/plain.log 2>&1; tail -1 gpurun_out/plain.log >> gpurun_out/summary.txt
F="ncu --profile-from-start off --clock-control none --set full --import-source on"
timeout 600 $F -k regex:paged_decode_attn -s 5 -c 1 -f -o gpurun_out/r02h_full_attn_int8_bs64 $P > gpurun_out/ncu.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
