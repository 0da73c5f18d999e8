"""KV cache + decode attention on the GPU, with the reference's names and call shapes
(src/wrinklefree_inference/kv_cache/kv_cache.py: KVCacheDtype :30-37, KVCacheConfig :40-93,
KVCache :96-313, attention_with_kv_cache :316-369, compute_kv_cache_memory :372-404).

Storage is the paged pool of the CUDA library ([page][layer][K|V][head][slot][D], page_size from the
config — the reference declares `page_size=16` but never uses it); a sequence owns pages 0..n in order.
`attention_with_kv_cache` = append + split-KV flash-decode kernel (wf_paged_decode_attn).

Every cache dtype of the reference is supported:
  BF16               paged bf16 pool, attention straight from the pool.
  INT8 (the reference's default, kv_cache.py:56), FP8_E4M3, FP8_E5M2
                     paged int8 pool, attention straight from the pool.  The reference's two "FP8" modes are int8 storage with
                     an absmax / 127 scale (kv_cache.py:150-196) — the same arithmetic as INT8 (:198-228); tests/golden/kv_modes.npz,
                     written by the reference's own class, shows identical codes, scales and dequantised rows for all three.
  FP16, FP32         16- / 32-bit paged pool (update / get are exact copies); attention casts the cached rows to bf16 into a
                     scratch pool first (the reference casts them to the query dtype, :344-346) — O(context) per call, kept for
                     API completeness, not a fast path.
INT8 / FP8 use the reference's arithmetic (scale = 127/absmax, round, clamp +-127, dequantise to bf16) with one scale per
(token, head) row.  The reference keeps ONE scale per layer, re-derived from the most recent update() call and applied by get()
to everything already cached - with token-by-token decode that rescales old entries by the newest token's absmax; a row-wise
scale is what that code does when it is handed one row, without the defect (tests/golden/kv_int8.npz pins the row arithmetic to
the reference's own KVCache).  `k_scales` / `v_scales` are kept for API compatibility and hold the scale of the most recent row.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from enum import Enum
from typing import Optional, Tuple

import torch


class KVCacheDtype(Enum):
    BF16 = "bfloat16"
    FP16 = "float16"
    FP32 = "float32"
    FP8_E4M3 = "fp8_e4m3"
    FP8_E5M2 = "fp8_e5m2"
    INT8 = "int8"


@dataclass
class KVCacheConfig:
    max_seq_len: int = 4096
    num_layers: int = 32
    num_heads: int = 32
    head_dim: int = 128
    dtype: KVCacheDtype = KVCacheDtype.INT8   # the reference's default (kv_cache.py:56-57)
    page_size: int = 16

    @property
    def hidden_dim(self) -> int:
        return self.num_heads * self.head_dim

    @property
    def bytes_per_token_per_layer(self) -> int:
        if self.dtype in (KVCacheDtype.BF16, KVCacheDtype.FP16):
            b = 2
        elif self.dtype == KVCacheDtype.FP32:
            b = 4
        elif self.dtype in (KVCacheDtype.FP8_E4M3, KVCacheDtype.FP8_E5M2, KVCacheDtype.INT8):
            b = 1
        else:
            b = 2
        return 2 * self.hidden_dim * b

    @property
    def total_bytes(self) -> int:
        return self.max_seq_len * self.num_layers * self.bytes_per_token_per_layer

    def __repr__(self) -> str:
        return (f"KVCacheConfig(max_seq_len={self.max_seq_len}, layers={self.num_layers}, heads={self.num_heads}, "
                f"head_dim={self.head_dim}, dtype={self.dtype.value}, total={self.total_bytes / (1024 * 1024):.1f}MB)")


class KVCache:
    """Single-sequence KV cache for all layers (reference :96-313) held in the paged device pool."""

    def __init__(self, config: KVCacheConfig, device: str = "cuda"):
        from wf_bitnet import ops
        self.config = config
        self.current_seq_len = 0
        self._q8 = config.dtype in (KVCacheDtype.INT8, KVCacheDtype.FP8_E4M3, KVCacheDtype.FP8_E5M2)
        if config.dtype == KVCacheDtype.BF16:
            self.storage_dtype = torch.bfloat16
        elif config.dtype == KVCacheDtype.FP16:
            self.storage_dtype = torch.float16
        elif config.dtype == KVCacheDtype.FP32:
            self.storage_dtype = torch.float32
        elif self._q8:
            self.storage_dtype = torch.bfloat16     # dtype of the rows handed to / returned by the int8 pool
        else:
            raise ValueError(f"unknown KV cache dtype {config.dtype!r}")
        # fp16 rows travel through the 16-bit pool bit for bit (scatter / gather are copies)
        self._pool_dtype = torch.int8 if self._q8 else (torch.bfloat16 if self.storage_dtype == torch.float16 else self.storage_dtype)
        self._scratch = None
        ps = config.page_size
        if ps % 16 != 0:
            raise ValueError("page_size must be a multiple of 16")
        self.num_pages = (config.max_seq_len + ps - 1) // ps
        self.kv = ops.PagedKV(self.num_pages, config.num_layers, config.num_heads, ps, config.head_dim,
                              dtype=self._pool_dtype, device=device)
        self.block_table = torch.arange(self.num_pages, dtype=torch.int32, device=device).reshape(1, -1)
        sdev = device if self._q8 else "cpu"
        self.k_scales = torch.ones(config.num_layers, dtype=torch.float32, device=sdev)
        self.v_scales = torch.ones(config.num_layers, dtype=torch.float32, device=sdev)

    def _pages_slots(self, start: int, n: int):
        pos = torch.arange(start, start + n, dtype=torch.int32, device=self.kv.pool.device)
        ps = self.config.page_size
        return (pos // ps).to(torch.int32), (pos % ps).to(torch.int32)

    def update(self, layer_idx: int, key: torch.Tensor, value: torch.Tensor, seq_pos: int) -> None:
        """key/value: [batch(=1), new_tokens, num_heads, head_dim]."""
        from wf_bitnet import ops
        if key.dim() != 4 or key.shape[0] != 1:
            raise ValueError("KVCache.update expects [1, new_tokens, num_heads, head_dim]")
        new_tokens = key.shape[1]
        end_pos = seq_pos + new_tokens
        if end_pos > self.config.max_seq_len:
            raise ValueError(f"Sequence position {end_pos} exceeds max {self.config.max_seq_len}")
        dev = self.kv.pool.device
        k = key.to(dev, self.storage_dtype).reshape(new_tokens, -1)
        v = value.to(dev, self.storage_dtype).reshape(new_tokens, -1)
        if self.storage_dtype == torch.float16:
            k, v = k.view(torch.bfloat16), v.view(torch.bfloat16)
        pages, slots = self._pages_slots(seq_pos, new_tokens)
        ops.kv_scatter(self.kv, k, v, pages, slots, layer_idx)
        if self._q8:
            pg, sl, hh = (end_pos - 1) // self.config.page_size, (end_pos - 1) % self.config.page_size, self.config.num_heads - 1
            self.k_scales[layer_idx] = self.kv.scales[pg, layer_idx, 0, hh, sl]      # device-side copies, no sync
            self.v_scales[layer_idx] = self.kv.scales[pg, layer_idx, 1, hh, sl]
        self.current_seq_len = max(self.current_seq_len, end_pos)

    def get(self, layer_idx: int, end_pos: Optional[int] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """Returns (key, value) [end_pos, num_heads, head_dim] (device tensors)."""
        from wf_bitnet import ops
        if end_pos is None:
            end_pos = self.current_seq_len
        c = self.config
        f16 = self.storage_dtype == torch.float16
        k = torch.empty(end_pos, c.num_heads * c.head_dim, dtype=torch.bfloat16 if f16 else self.storage_dtype, device=self.kv.pool.device)
        v = torch.empty_like(k)
        if end_pos > 0:
            pages, slots = self._pages_slots(0, end_pos)
            ops.kv_gather(self.kv, k, v, pages, slots, layer_idx)
        if f16:
            k, v = k.view(torch.float16), v.view(torch.float16)
        return k.view(end_pos, c.num_heads, c.head_dim), v.view(end_pos, c.num_heads, c.head_dim)

    def clear(self) -> None:
        (self.kv.raw if self._q8 else self.kv.pool).zero_()
        self.k_scales.fill_(1.0)
        self.v_scales.fill_(1.0)
        self.current_seq_len = 0

    def memory_usage_mb(self) -> float:
        t = self.kv.raw if self._q8 else self.kv.pool
        return t.numel() * t.element_size() / (1024 * 1024)


def attention_with_kv_cache(query: torch.Tensor, kv_cache: KVCache, layer_idx: int, new_key: torch.Tensor,
                            new_value: torch.Tensor, seq_pos: int, scale: Optional[float] = None) -> torch.Tensor:
    """query [1, H, 1, D]; new_key/new_value [1, 1, H, D] -> output [1, H, 1, D] (reference :316-369)."""
    from wf_bitnet import ops
    if query.dim() != 4 or query.shape[0] != 1 or query.shape[2] != 1:
        raise ValueError("query must be [1, num_heads, 1, head_dim]")
    kv_cache.update(layer_idx, new_key, new_value, seq_pos)
    if scale is None:
        scale = 1.0 / math.sqrt(query.shape[-1])
    dev = kv_cache.kv.pool.device
    q = query.to(dev, torch.bfloat16).reshape(1, query.shape[1], query.shape[3])
    seq_lens = torch.tensor([seq_pos + 1], dtype=torch.int32, device=dev)
    kv, layer = kv_cache.kv, layer_idx
    if kv_cache.storage_dtype != torch.bfloat16:
        # FP16 / FP32 caches: the reference casts what get() returns to the query dtype (:344-346); here the rows are cast to
        # bf16 into a one-layer scratch pool the attention kernel can stream.  O(context) per call — completeness, not speed.
        c = kv_cache.config
        if kv_cache._scratch is None:
            kv_cache._scratch = ops.PagedKV(kv_cache.num_pages, 1, c.num_heads, c.page_size, c.head_dim, dtype=torch.bfloat16, device=dev)
        k_all, v_all = kv_cache.get(layer_idx, seq_pos + 1)
        pages, slots = kv_cache._pages_slots(0, seq_pos + 1)
        ops.kv_scatter(kv_cache._scratch, k_all.to(torch.bfloat16).reshape(seq_pos + 1, -1), v_all.to(torch.bfloat16).reshape(seq_pos + 1, -1), pages, slots, 0)
        kv, layer = kv_cache._scratch, 0
    out = ops.paged_decode_attn(q, kv, layer, kv_cache.block_table, seq_lens, scale=scale,
                                out_dtype=torch.bfloat16 if query.dtype == torch.bfloat16 else torch.float32)
    return out.to(query.dtype).reshape(1, query.shape[1], 1, query.shape[3])


def compute_kv_cache_memory(model_config: dict, context_length: int, dtype: KVCacheDtype = KVCacheDtype.BF16) -> dict:
    config = KVCacheConfig(max_seq_len=context_length, num_layers=model_config.get("num_layers", 32),
                           num_heads=model_config.get("num_heads", 32), head_dim=model_config.get("head_dim", 128), dtype=dtype)
    return {"config": str(config), "total_mb": config.total_bytes / (1024 * 1024),
            "per_token_kb": config.bytes_per_token_per_layer * config.num_layers / 1024,
            "context_length": context_length, "dtype": dtype.value}
