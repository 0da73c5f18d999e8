#!/usr/bin/env python
"""Golden vectors for the int8 KV quantiser, produced by the reference's own KVCache (KVCacheDtype.INT8).

Runs only in the build container (imports /root/reference/src/wrinklefree_inference/kv_cache/kv_cache.py); writes
tests/golden/kv_int8.npz.  The reference quantises whatever tensor one update() call hands it with ONE scale
(kv_cache.py:198-218); feeding it one (token, head) row at a time yields the per-row arithmetic the GPU path and
oracle.model.kv_int8_quant use: codes, the stored dequant scale, and get()'s dequantised bf16 row (:220-228)."""
import importlib.util
import sys
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent


def main():
    spec = importlib.util.spec_from_file_location("ref_kv_cache", REF / "src/wrinklefree_inference/kv_cache/kv_cache.py")
    kvmod = importlib.util.module_from_spec(spec)
    sys.modules["ref_kv_cache"] = kvmod
    spec.loader.exec_module(kvmod)
    D, R = 128, 96
    g = torch.Generator().manual_seed(2024)
    x = torch.randn(R, D, generator=g) * torch.logspace(-3, 3, R).unsqueeze(1)
    x[0] = 0.0                      # absmax < 1e-6 -> scale 1
    x[1] = 1e-8
    x[2, 5] = 300.0                 # one outlier
    x = x.to(torch.bfloat16).float()   # values a bf16 K/V row can hold, handed over as fp32
    cfg = kvmod.KVCacheConfig(max_seq_len=4, num_layers=1, num_heads=1, head_dim=D, dtype=kvmod.KVCacheDtype.INT8)
    cache = kvmod.KVCache(cfg)
    codes, scales, deq = [], [], []
    for r in range(R):
        row = x[r].view(1, 1, 1, D)
        cache.update(0, row, row, 0)
        codes.append(cache.cache[0, 0, 0].to(torch.int8).flatten().numpy().copy())
        scales.append(float(cache.k_scales[0]))
        k, _ = cache.get(0, 1)
        deq.append(k.float().flatten().numpy().copy())
    np.savez_compressed(OUT / "kv_int8.npz", x=x.numpy(), codes=np.stack(codes), scales=np.array(scales, np.float64), deq=np.stack(deq))
    print("wrote", OUT / "kv_int8.npz")


if __name__ == "__main__":
    main()
