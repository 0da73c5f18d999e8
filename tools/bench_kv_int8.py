#!/usr/bin/env python
"""bf16 vs int8 KV on the headline shape (2B, batch 64, 4K context): device-timed decode step and the attention kernel alone."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "wrinklefree-inference-engine_b200"))
from wf_bitnet.engine import Engine, EngineConfig  # noqa: E402

batch, ctx, steps = int(os.environ.get("B", 64)), int(os.environ.get("CTX", 4096)), 30
for kvd in os.environ.get("KVS", "bf16,int8").split(","):
    cfg = EngineConfig.bitnet_2b(max_batch=batch, max_seq=ctx + 64, kv_dtype=kvd)
    eng = Engine(cfg)
    eng.init_synthetic(1234)
    for s in range(batch):
        eng.set_context(s, ctx, fill_random=True)
    for _ in range(5):
        eng.decode_device(batch)
    eng.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = torch.cuda.ExternalStream(eng.stream)
    with torch.cuda.stream(st):
        e0.record()
        for _ in range(steps):
            eng.decode_device(batch)
        e1.record()
    eng.sync()
    ms = e0.elapsed_time(e1) / steps
    attn_ms = eng.time_kernel("attn", batch, 60)
    elt = 2 if kvd == "bf16" else 1
    abytes = batch * (ctx + 6) * 2 * cfg.kv_heads * cfg.head_dim * elt + (0 if kvd == "bf16" else batch * (ctx + 6) * 2 * cfg.kv_heads * 4)
    print(json.dumps({"kv": kvd, "batch": batch, "ctx": ctx, "ms_per_step": round(ms, 4), "tok_s": round(batch / ms * 1e3, 1),
                      "attn_us": round(attn_ms * 1e3, 2), "attn_GBs": round(abytes / attn_ms / 1e6, 1), "attn_MB": round(abytes / 1e6, 1),
                      "pool_GB": round(eng.device_bytes() / 1e9, 2)}))
    eng.close()
    del eng
    torch.cuda.empty_cache()
