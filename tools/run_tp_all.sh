#!/bin/bash
# usage: tools/run_tp_all.sh <n_gpus>  — TP parity check + the three TP workloads at TP = n_gpus (one JSON line each)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$1
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/tp_gpus_$N.txt
if [ "$N" != "1" ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tools/tp_check.py > gpurun_out/tp_check_$N.log 2>&1
  echo "tp_check n=$N exit=$?"; tail -3 gpurun_out/tp_check_$N.log
fi
for W in 8b_bs1 8b_bs32 70b_bs16; do
  bash tools/run_tp_bench.sh $N $W 30
done
