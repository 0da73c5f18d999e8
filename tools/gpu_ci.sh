#!/bin/bash
# Runs the GPU test files one process each (a hung or faulted kernel only loses its own file),
# then a short bench.  Everything lands in gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
for f in tests/test_gpu_pack_quant.py tests/test_gpu_linear.py tests/test_gpu_attention.py tests/test_gpu_engine.py tests/test_gpu_gemm_tc.py tests/test_gpu_loader.py tests/test_gpu_sglang_model.py tests/test_gpu_kv_int8.py tests/test_zz_no_alloc.py; do
  b=$(basename $f .py)
  timeout ${TEST_TIMEOUT:-600} python -m pytest $f -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/$b.log 2>&1
  echo "$b exit=$?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$b.log >> gpurun_out/summary.txt
done
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit=$?" >> gpurun_out/summary.txt
tail -1 gpurun_out/smoke.log >> gpurun_out/summary.txt
if [ "${RUN_BENCH:-1}" = "1" ]; then
  timeout 900 python bench.py --steps ${BENCH_STEPS:-10} --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
  echo "bench exit=$?" >> gpurun_out/summary.txt
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
  echo "bench_ref exit=$?" >> gpurun_out/summary.txt
fi
cat gpurun_out/summary.txt
