#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for lam in 1 0; do
echo "WF_TAIL_LAM=$lam" >> gpurun_out/summary.txt
WF_TAIL_LAM=$lam timeout 120 python tools/profile_step.py --shape 2b --batch 1 --ctx 64 --layers 1 --steps 3 --no-graph 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
WF_TAIL_LAM=$lam timeout 120 python tools/profile_step.py --shape 2b --batch 1 --ctx 512 --steps 30 2>&1 | tail -1 | cut -c1-120 >> gpurun_out/summary.txt
done
timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -q --no-header -rf -x -p no:cacheprovider -k "not config0" > gpurun_out/test_gpu_engine.log 2>&1
echo "test_gpu_engine (no config0) exit=$?" >> gpurun_out/summary.txt
tail -4 gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
