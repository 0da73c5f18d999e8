#!/usr/bin/env python
"""Tensor-parallel parity check, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/tp_check.py

Every rank builds the SAME seeded synthetic model, keeps its shard (wf_bitnet.tp.shard_weights), and the TP engine
decodes a few steps.  Rank 0 also runs the unsharded engine on its own GPU: hidden states, normalised final
activations and greedy tokens must be BIT-IDENTICAL (exact integer all-reduce, full-vector quantisation), for the
small-batch (GEMV) and the batched (one-shot GEMM) phase kernels.  Exit code 0 = parity.
"""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    from wf_bitnet import _lib, tp
    from wf_bitnet.engine import Engine, EngineConfig
    from wf_bitnet.synth import make_weights

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.require_gpu()
    ok = True
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    # last case: the TP engine is filled by the checkpoint loader from a FULL HF-packed state dict (every rank shards on the fly)
    for batch, steps, ctx, kvd, via_loader in ((1, 6, 0, "bf16", False), (3, 4, 37, "bf16", False), (12, 4, 21, "bf16", False),
                                               (3, 6, 0, "int8", False), (12, 20, 0, "int8", False), (3, 4, 0, "bf16", True)):
        cfg = EngineConfig(hidden=512, layers=3, heads=4 * world, kv_heads=world, head_dim=128, ffn=256 * world, vocab=1000, max_seq=128,
                           page_size=16, max_batch=batch, kv_dtype=kvd)
        weights = make_weights(cfg, seed=5 + batch, keep_ternary=via_loader)
        eng = Engine(cfg, tp_rank=rank, tp_size=world)
        if via_loader:
            from test_gpu_loader import _state_dict
            from wf_bitnet import loader
            loader.load_state_dict(eng, _state_dict(cfg, weights, "hf"))
        else:
            eng.load_weights(tp.shard_weights(weights, rank, world))
        tp.attach(eng)
        ref = None
        if rank == 0:
            ref = Engine(cfg)
            ref.load_weights(weights)
        rng = np.random.default_rng(batch)
        if ctx:   # a synthetic random context: the TP engine holds its KV heads of the same content
            for e in ([eng, ref] if ref else [eng]):
                for slot in range(batch):
                    e.set_context(slot, ctx + slot, fill_random=False)
        tokens = rng.integers(0, cfg.vocab, size=batch).astype(np.int32)
        t_tp, t_ref = tokens.copy(), tokens.copy()
        for step in range(steps):
            t_tp = eng.decode(t_tp)
            if ref is not None:
                t_ref = ref.decode(t_ref)
                h_tp = eng.buffer("normed", (batch, cfg.hidden), torch.bfloat16)
                h_ref = ref.buffer("normed", (batch, cfg.hidden), torch.bfloat16)
                same_h = torch.equal(h_tp, h_ref)
                same_t = np.array_equal(t_tp, t_ref)
                if not (same_h and same_t):
                    ok = False
                    print(f"[tp_check] MISMATCH batch={batch} step={step}: hidden equal={same_h} tokens {t_tp} vs {t_ref}", flush=True)
        # device-resident graph replay path
        for _ in range(3):
            eng.decode_device(batch)
            if ref is not None:
                ref.decode_device(batch)
        eng.sync()
        if ref is not None:
            ref.sync()
            n_tp = eng.buffer("next", (batch,), torch.int32)
            n_ref = ref.buffer("next", (batch,), torch.int32)
            if not torch.equal(n_tp, n_ref):
                ok = False
                print(f"[tp_check] MISMATCH after graph replay batch={batch}: {n_tp} vs {n_ref}", flush=True)
            print(f"[tp_check] batch={batch} kv={kvd} loader={via_loader}: tp{world} == tp1 over {steps}+3 steps: {ok}; launches/step tp={eng.launches_per_step()} ref={ref.launches_per_step()}", flush=True)
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item())
        dist.barrier()
        eng.close()
        if ref is not None:
            ref.close()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
