#!/usr/bin/env python
"""Golden vectors for the remaining cache dtypes of the reference's KVCache: the two "FP8" modes (int8 storage with a
per-update absmax scale, kv_cache/kv_cache.py:150-196) and FP16 (plain cast, :262-264), produced by the reference's own
class.  Runs only in the build container (imports /root/reference/src/wrinklefree_inference/kv_cache/kv_cache.py);
writes tests/golden/kv_modes.npz.  As for INT8 (make_golden_kv_int8.py) the class is fed one (token, head) row per
update() call, which is the granularity the GPU pool quantises at."""
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
    g = torch.Generator().manual_seed(2024)                     # the same rows as kv_int8.npz
    x = torch.randn(R, D, generator=g) * torch.logspace(-3, 3, R).unsqueeze(1)
    x[0] = 0.0
    x[1] = 1e-8
    x[2, 5] = 300.0
    x = x.to(torch.bfloat16).float()
    out = {"x": x.numpy()}
    for name in ("FP8_E4M3", "FP8_E5M2"):
        cfg = kvmod.KVCacheConfig(max_seq_len=4, num_layers=1, num_heads=1, head_dim=D, dtype=getattr(kvmod.KVCacheDtype, name))
        cache = kvmod.KVCache(cfg)
        codes, scales, deq = [], [], []
        for r in range(R):
            row = x[r].view(1, 1, 1, D)
            cache.update(0, row, row, 0)
            codes.append(cache.cache[0, 0, 0].view(torch.int8).flatten().numpy().copy())
            scales.append(float(cache.k_scales[0]))
            k, _ = cache.get(0, 1)
            deq.append(k.float().flatten().numpy().copy())
        out[f"{name}_codes"] = np.stack(codes)
        out[f"{name}_scales"] = np.array(scales, np.float64)
        out[f"{name}_deq"] = np.stack(deq)
    cfg = kvmod.KVCacheConfig(max_seq_len=R, num_layers=1, num_heads=1, head_dim=D, dtype=kvmod.KVCacheDtype.FP16)
    cache = kvmod.KVCache(cfg)
    for r in range(R):
        row = x[r].view(1, 1, 1, D)
        cache.update(0, row, row * 0.5, r)
    k, v = cache.get(0)
    out["FP16_k"] = k.float().reshape(R, D).numpy()
    out["FP16_v"] = v.float().reshape(R, D).numpy()
    # attention_with_kv_cache of the reference itself over an FP16 cache (bf16 query), 24 tokens of 2 heads
    H, S = 2, 24
    cfg = kvmod.KVCacheConfig(max_seq_len=32, num_layers=1, num_heads=H, head_dim=D, dtype=kvmod.KVCacheDtype.FP16)
    cache = kvmod.KVCache(cfg)
    g2 = torch.Generator().manual_seed(77)
    ks = torch.randn(S, H, D, generator=g2).to(torch.bfloat16)
    vs = torch.randn(S, H, D, generator=g2).to(torch.bfloat16)
    q = torch.randn(1, H, 1, D, generator=g2).to(torch.bfloat16)
    for pos in range(S - 1):
        cache.update(0, ks[pos].view(1, 1, H, D), vs[pos].view(1, 1, H, D), pos)
    o = kvmod.attention_with_kv_cache(q, cache, 0, ks[S - 1].view(1, 1, H, D), vs[S - 1].view(1, 1, H, D), S - 1)
    out["attn_k"] = ks.float().numpy(); out["attn_v"] = vs.float().numpy(); out["attn_q"] = q.float().numpy(); out["attn_out"] = o.float().numpy()
    np.savez_compressed(OUT / "kv_modes.npz", **out)
    i8 = np.load(OUT / "kv_int8.npz")
    for name in ("FP8_E4M3", "FP8_E5M2"):
        print(name, "codes == INT8 codes:", np.array_equal(out[f"{name}_codes"], i8["codes"]),
              " scales == INT8 scales:", np.array_equal(np.float32(out[f"{name}_scales"]), np.float32(i8["scales"])),
              " deq == INT8 deq:", np.array_equal(out[f"{name}_deq"], i8["deq"]))
    print("wrote", OUT / "kv_modes.npz")


if __name__ == "__main__":
    main()
