#!/bin/bash
# round-2 call O: launch lists at the 8B / 70B shapes (where does a TP = 1 step go?)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
NCU="ncu --profile-from-start off --clock-control none --metrics gpu__time_duration.sum --csv"
for cfg in "70b 16 8192" "8b 32 2048" "8b 1 512"; do set -- $cfg
  WF_DECODE_TC_MIN=9 python tools/profile_step.py --shape $1 --batch $2 --ctx $3 > gpurun_out/plain_$1_$2.log 2>&1 &&
  WF_DECODE_TC_MIN=9 $NCU --log-file gpurun_out/r02_launches_$1_bs$2.csv python tools/profile_step.py --shape $1 --batch $2 --ctx $3 > gpurun_out/ncu_$1_$2.log 2>&1
  echo "$1 bs$2 exit=$?" >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
