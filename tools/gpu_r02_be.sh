#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/tp_parity_diag.py"
WF_TP_PARITY_BATCHES=16 WF_DECODE_STREAM=0 timeout 300 $T 2>&1 | grep DIAG >> gpurun_out/summary.txt
WF_TP_PARITY_BATCHES=9,12,16,16,24 timeout 300 $T 2>&1 | grep DIAG >> gpurun_out/summary.txt
WF_TP_PARITY_BATCHES=16 WF_PDL=0 timeout 300 $T 2>&1 | grep DIAG >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
