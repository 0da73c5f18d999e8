#!/bin/bash
# round-2 multi-GPU call (N = 2): TP parity tests, the bench exactly as the driver launches it (aux.tp), tp_check
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
N=${N:-2}
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -q --no-header -rf -p no:cacheprovider > gpurun_out/test_gpu_tp.log 2>&1
echo "test_gpu_tp exit=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/test_gpu_tp.log >> gpurun_out/summary.txt
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
echo "bench n=$N exit=$?" >> gpurun_out/summary.txt
python - <<PY >> gpurun_out/summary.txt
import json
try:
    d = json.load(open("gpurun_out/r02_bench_n$N.json"))
    print("value", d["value"], "ms", d["ms_per_step"])
    for k, v in d["aux"]["tp"].items():
        if isinstance(v, dict) and "ms_per_step" in v:
            print(k, {x: v[x] for x in ("tp", "ms_per_step", "tok_s", "speedup_vs_tp1", "hbm_frac_per_gpu", "us_per_exchange", "parity_bit_identical") if x in v})
    print("parity", d["aux"]["tp"].get("parity"))
except Exception as ex:
    print("bench json error", ex)
PY
cat gpurun_out/summary.txt
