#!/bin/bash
# round-2 call AV: ncu --set full of the streaming GEMM inside a 70B-shaped batch-16 step (the four linears of one layer)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py --shape 70b --batch 16 --ctx 8192 --layers 4"
$P > gpurun_out/plain.log 2>&1; tail -1 gpurun_out/plain.log >> gpurun_out/summary.txt
L="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
$L --log-file gpurun_out/r02h_launches_70b_bs16.csv $P > gpurun_out/ncu_l.log 2>&1
F="ncu --profile-from-start off --clock-control none --set full --import-source on"
timeout 600 $F -k regex:gemm_stream -s 4 -c 4 -f -o gpurun_out/r02h_full_gemm_stream_70b_bs16 $P > gpurun_out/ncu.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
