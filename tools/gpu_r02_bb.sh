#!/bin/bash
# round-2 call BB: cluster row glue at batch 1..4 on the 8B / 70B widths (one 512-thread CTA per 8192..28672-element row otherwise)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py --steps 20"
run() { echo "== $*" >> gpurun_out/summary.txt; env "$@" timeout 300 $P --shape $S --batch $B --ctx $C 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/summary.txt; }
for cfg in "70b 1 2048" "70b 4 2048" "8b 1 512" "8b 4 512"; do set -- $cfg; S=$1; B=$2; C=$3
  for g in 0 1 4 8; do run WF_GLUE_CLUSTER=$g; done
done
L="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
$L --log-file gpurun_out/r02h_launches_70b_bs1.csv $P --shape 70b --batch 1 --ctx 2048 --layers 4 > gpurun_out/ncu_l.log 2>&1
cat gpurun_out/summary.txt
