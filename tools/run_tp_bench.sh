#!/bin/bash
# usage: tools/run_tp_bench.sh <n_gpus> <workload> [steps]   -> one JSON line appended to gpurun_out/tp_bench.jsonl
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$1; W=$2; S=${3:-30}
if [ "$N" = "1" ]; then
  timeout 600 python bench.py --gpus 1 --steps $S --warmup 3 --workload $W --no-cpu --no-aux 2> gpurun_out/tp_bench_${W}_$N.err | tee -a gpurun_out/tp_bench.jsonl | cut -c1-400
else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps $S --warmup 3 --workload $W --no-cpu --no-aux 2> gpurun_out/tp_bench_${W}_$N.err | tee -a gpurun_out/tp_bench.jsonl | cut -c1-400
fi
echo "exit=$? n=$N workload=$W"
