#!/usr/bin/env python
"""torchrun diagnostic: bench._tp_parity at the batch sizes of WF_TP_PARITY_BATCHES"""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    sys.path.insert(0, p)
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
import bench
ok, detail = bench._tp_parity(rank, world, dist)
if rank == 0:
    print("DIAG", os.environ.get("WF_DECODE_STREAM", "default"), os.environ.get("WF_TP_PARITY_BATCHES"), ok, detail, flush=True)
dist.destroy_process_group()
