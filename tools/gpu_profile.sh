#!/bin/bash
# Plain runs (the "exits 0 without ncu" gate), then the ncu launch lists and the --set full captures
# of the two dominant kernels.  Everything lands in gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
P="python tools/profile_step.py"
$P --batch 64 --ctx 4096 > gpurun_out/plain_bs64.log 2>&1 || { echo "plain bs64 failed"; tail -5 gpurun_out/plain_bs64.log; exit 1; }
$P --batch 1 --ctx 512 > gpurun_out/plain_bs1.log 2>&1 || { echo "plain bs1 failed"; tail -5 gpurun_out/plain_bs1.log; exit 1; }
WF_PDL=1 $P --batch 64 --ctx 4096 > gpurun_out/plain_bs64_pdl.log 2>&1
WF_PDL=1 $P --batch 1 --ctx 512 > gpurun_out/plain_bs1_pdl.log 2>&1
cat gpurun_out/plain_*.log
NCU="ncu --profile-from-start off --clock-control none"
$NCU --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches_bs64.csv $P --batch 64 --ctx 4096 > gpurun_out/ncu_bs64.log 2>&1
$NCU --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches_bs1.csv $P --batch 1 --ctx 512 > gpurun_out/ncu_bs1.log 2>&1
if [ "${FULL:-1}" = "1" ]; then
  $NCU --set full --import-source on -k regex:${K64:-paged_decode_attn} -c 2 -f -o gpurun_out/full_bs64 $P --batch 64 --ctx 4096 > gpurun_out/ncu_full_bs64.log 2>&1
  $NCU --set full --import-source on -k regex:${K1:-gemv_stream} -c 4 -f -o gpurun_out/full_bs1 $P --batch 1 --ctx 512 > gpurun_out/ncu_full_bs1.log 2>&1
fi
ls -la gpurun_out | tail -20
