#!/bin/bash
# round-2 multi-GPU call (N = 8 by default): the bench exactly as the driver launches it (replica headline + aux.tp strong-scaling points and TP parity)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
N=${N:-8}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8 >> gpurun_out/summary.txt
timeout 480 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
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
tail -5 gpurun_out/r02_bench_n$N.err >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
