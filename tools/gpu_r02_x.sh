#!/bin/bash
# round-2 call X: hand-off tail at the 8B shape (register glue with 5 granules for the ffn row): parity vs phase kernels + timing
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
python - <<'PY' >> gpurun_out/summary.txt 2>&1
import os, sys
sys.path.insert(0, "."); sys.path.insert(0, "wrinklefree-inference-engine_b200")
import dataclasses
import numpy as np, torch
from wf_bitnet.engine import Engine, EngineConfig
outs = {}
for name, env in (("lam", {"WF_TAIL": "1", "WF_TAIL_LAM": "1"}), ("bar", {"WF_TAIL": "1", "WF_TAIL_LAM": "0"}), ("phase", {"WF_TAIL": "0"})):
    os.environ.update(env)
    cfg = dataclasses.replace(EngineConfig.llama3_8b(max_seq=256, max_batch=1), layers=3, vocab=4096)
    eng = Engine(cfg); eng.init_synthetic(7)
    eng.set_context(0, 100, fill_random=True)
    toks = [5]; res = []
    for _ in range(6):
        toks = eng.decode(toks)
        res.append(eng.buffer("logits", (1, cfg.vocab), torch.float32).cpu())
    outs[name] = (res, eng.launches_per_step()); eng.close()
for a in ("lam", "bar"):
    same = all(torch.equal(x, y) for x, y in zip(outs[a][0], outs["phase"][0]))
    print("8B-shape 3 layers, batch 1:", a, "launches", outs[a][1], "bit-identical to phase kernels:", same)
PY
P="python tools/profile_step.py"
for lam in 1 0; do
  echo "WF_TAIL_LAM=$lam" >> gpurun_out/summary.txt
  for cfg in "8b 1 512" "8b 1 4096" "2b 1 512"; do set -- $cfg
    WF_TAIL_LAM=$lam timeout 300 $P --shape $1 --batch $2 --ctx $3 --steps 30 2>&1 | tail -1 | cut -c1-150 >> gpurun_out/summary.txt
  done
done
cat gpurun_out/summary.txt
