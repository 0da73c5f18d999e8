#!/bin/bash
# round-2 call Z: cost-model chunking of the attention grid vs the old rule (WF_ATTN_WANT=4)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
P="python tools/profile_step.py"
for mode in "new8" "new14" "old"; do
  case $mode in new8) export WF_ATTN_WANT=0 WF_ATTN_C0=8;; new14) export WF_ATTN_WANT=0 WF_ATTN_C0=14;; old) export WF_ATTN_WANT=4 WF_ATTN_C0=8;; esac
  echo "mode=$mode" >> gpurun_out/summary.txt
  for cfg in "2b 64 4096" "2b 32 2048" "2b 64 1024" "2b 8 8192" "2b 1 16384" "8b 32 2048" "8b 16 2048" "8b 64 2048" "70b 16 8192"; do set -- $cfg
    timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 15 2>&1 | tail -1 | cut -c1-150 >> gpurun_out/summary.txt
  done
done
timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_kv_int8.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/test_attn.log 2>&1
echo "attention tests exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_attn.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
