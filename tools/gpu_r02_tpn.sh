#!/bin/bash
# round-2 multi-GPU call: TP parity (tools/tp_check.py: batches 1 / 3 / 12, bf16 and int8 KV) and the driver's bench line at N GPUs, on the tree with
# the streaming batched GEMM (row-parallel linears of up to 16 rows RED through it into the owner ranks)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
N=${N:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tools/tp_check.py > gpurun_out/tp_check_$N.log 2>&1
echo "tp_check n=$N exit=$?" >> gpurun_out/summary.txt; grep "tp_check" gpurun_out/tp_check_$N.log | tail -8 | cut -c1-200 >> gpurun_out/summary.txt
timeout 480 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02b_bench_n$N.json 2> gpurun_out/r02b_bench_n$N.err
echo "bench n=$N exit=$?" >> gpurun_out/summary.txt
python - <<PY >> gpurun_out/summary.txt
import json
try:
    d = json.load(open("gpurun_out/r02b_bench_n$N.json"))
    print("value", d["value"], "ms", d["ms_per_step"])
    for k, v in d["aux"]["tp"].items():
        if isinstance(v, dict) and "ms_per_step" in v:
            print(k, {x: v[x] for x in ("tp", "ms_per_step", "tok_s", "speedup_vs_tp1", "us_per_exchange", "parity_bit_identical") if x in v})
    print("parity", d["aux"]["tp"].get("parity", {}).get("bit_identical"))
except Exception as ex:
    print("bench json error", ex)
PY
cat gpurun_out/summary.txt
