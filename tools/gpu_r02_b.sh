#!/bin/bash
# round-2 call B: the tests that changed, the layer-tail kernel A/B (bs1), config-0 diagnostics, compute-sanitizer memcheck
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in tests/test_gpu_engine.py tests/test_gpu_linear.py tests/test_gpu_attention.py tests/test_gpu_sglang_model.py tests/test_gpu_kv_int8.py; do
  b=$(basename $f .py)
  timeout 1500 python -m pytest $f -m gpu -q --no-header -rf -s -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$b.log >> gpurun_out/summary.txt
done
grep -h "config0" gpurun_out/test_gpu_engine.log >> gpurun_out/summary.txt
for t in 1 0; do
  echo "WF_TAIL=$t" >> gpurun_out/summary.txt
  WF_TAIL=$t timeout 300 python tools/profile_step.py --batch 1 --ctx 512 --steps 50 >> gpurun_out/summary.txt 2>&1
  WF_TAIL=$t timeout 300 python tools/profile_step.py --batch 4 --ctx 512 --steps 50 >> gpurun_out/summary.txt 2>&1
  WF_TAIL=$t timeout 300 python tools/profile_step.py --batch 8 --ctx 512 --steps 50 >> gpurun_out/summary.txt 2>&1
  WF_TAIL=$t timeout 300 python tools/profile_step.py --batch 16 --ctx 2048 --steps 20 >> gpurun_out/summary.txt 2>&1
  WF_TAIL=$t timeout 300 python tools/profile_step.py --batch 64 --ctx 4096 --steps 20 >> gpurun_out/summary.txt 2>&1
done
timeout 600 python tools/cfg0_diag.py --layers 4,30 --steps 24 > gpurun_out/cfg0_diag.log 2>&1
echo "cfg0_diag exit=$?" >> gpurun_out/summary.txt
grep summary gpurun_out/cfg0_diag.log >> gpurun_out/summary.txt
timeout 300 python tools/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 &&
timeout 1200 compute-sanitizer --tool memcheck --log-file gpurun_out/r02_sanitizer_memcheck.log python tools/sanitize_target.py > gpurun_out/sanitize_memcheck.out 2>&1
echo "memcheck exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/r02_sanitizer_memcheck.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
