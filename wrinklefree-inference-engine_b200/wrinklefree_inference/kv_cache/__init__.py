"""KV cache utilities (reference: src/wrinklefree_inference/kv_cache/__init__.py:8-24; the HTTP
validator is out of scope — it needs a running server)."""
from .kv_cache import (
    KVCache,
    KVCacheConfig,
    KVCacheDtype,
    attention_with_kv_cache,
    compute_kv_cache_memory,
)

__all__ = ["KVCache", "KVCacheConfig", "KVCacheDtype", "attention_with_kv_cache", "compute_kv_cache_memory"]
