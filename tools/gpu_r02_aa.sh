#!/bin/bash
# round-2 call AA: tcgen05 LM head (parity, kernel timing at several batches, step timing), new attention chunk model re-check
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_gpu_attention.py -m gpu -q --no-header -rf -p no:cacheprovider -k "lm_head" > gpurun_out/test_lm.log 2>&1
echo "lm head test exit=$?" >> gpurun_out/summary.txt
tail -5 gpurun_out/test_lm.log >> gpurun_out/summary.txt
for tc in 1 0; do
  echo "WF_LM_TC=$tc" >> gpurun_out/summary.txt
  WF_LM_TC=$tc timeout 300 python tools/bench_lm_head.py 2>&1 | tail -4 >> gpurun_out/summary.txt
done
P="python tools/profile_step.py"
for cfg in "2b 64 4096" "2b 16 2048" "2b 1 16384" "8b 32 2048"; do set -- $cfg
  timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 20 2>&1 | tail -1 | cut -c1-150 >> gpurun_out/summary.txt
done
timeout 900 python -m pytest tests/test_gpu_engine.py tests/test_gpu_sglang_model.py -m gpu -q --no-header -rf -p no:cacheprovider -k "not config0" > gpurun_out/test_eng.log 2>&1
echo "engine tests exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_eng.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
