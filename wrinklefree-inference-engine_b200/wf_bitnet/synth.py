"""Seeded synthetic BitNet models (random-init ternary weights), SURVEY.md §8(d).

weights: ternary uniform over {-1,0,1} (as scripts/benchmark_kernels.py:35, extern/BitNet/gpu/test.py:43),
weight_scale ~ U(0.5, 2.5), norm gammas 1 + 0.1 N(0,1), embeddings N(0, 0.02) bf16.
The packer below is a vectorised numpy statement of the blocked I2_S layout
(cache/bitnet_converter.py:52-71) used to PREPARE inputs; it is host-side data generation, not a compute path.
"""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch


def pack_i2s_numpy(w: np.ndarray) -> np.ndarray:
    """ternary int8/float [N,K] (K % 128 == 0) -> blocked I2_S u8 [N,K/4]."""
    N, K = w.shape
    assert K % 128 == 0
    if w.dtype != np.int8:
        w = np.rint(w).clip(-1, 1).astype(np.int8)
    c = (w.clip(-1, 1) + 1).view(np.uint8).reshape(N, K // 128, 4, 32)   # int8 all the way: 30x faster than the int16 detour on a 2B model
    return ((c[:, :, 0] << 6) | (c[:, :, 1] << 4) | (c[:, :, 2] << 2) | c[:, :, 3]).reshape(N, K // 4)


def rope_tables(head_dim: int, max_seq: int, theta: float):
    """RotaryEmbedding.__init__ (direct_inference_full.py:31-45), computed with torch CPU like the reference."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    freqs = torch.outer(torch.arange(max_seq, dtype=torch.float32), inv_freq)
    return torch.cos(freqs).to(torch.bfloat16), torch.sin(freqs).to(torch.bfloat16)


def make_weights(cfg, seed: int = 1234, keep_ternary: bool = False) -> Dict:
    """cfg needs hidden, layers, heads, kv_heads, head_dim, ffn, vocab, rope_theta, max_seq."""
    shapes = {
        "q": (cfg.heads * cfg.head_dim, cfg.hidden), "k": (cfg.kv_heads * cfg.head_dim, cfg.hidden),
        "v": (cfg.kv_heads * cfg.head_dim, cfg.hidden), "o": (cfg.hidden, cfg.heads * cfg.head_dim),
        "gate": (cfg.ffn, cfg.hidden), "up": (cfg.ffn, cfg.hidden), "down": (cfg.hidden, cfg.ffn),
    }
    layers = []
    for li in range(cfg.layers):
        layer = {}
        for pi, (name, (n, k)) in enumerate(shapes.items()):
            rng = np.random.default_rng(seed + li * 7 + pi)
            w = rng.integers(-1, 2, size=(n, k), dtype=np.int8)
            layer[f"{name}_proj"] = pack_i2s_numpy(w)
            layer[f"{name}_scale"] = float(np.float32(rng.uniform(0.5, 2.5)))
            if keep_ternary:
                layer[f"{name}_ternary"] = w
        g = torch.Generator().manual_seed(seed * 31 + li)
        for nm, dim in (("input_layernorm", cfg.hidden), ("post_attention_layernorm", cfg.hidden),
                        ("attn_sub_norm", cfg.heads * cfg.head_dim), ("ffn_sub_norm", cfg.ffn)):
            layer[nm] = (1.0 + 0.1 * torch.randn(dim, generator=g)).to(torch.bfloat16)
        layers.append(layer)
    g = torch.Generator().manual_seed(seed * 131 + 7)
    embed = (0.02 * torch.randn(cfg.vocab, cfg.hidden, generator=g)).to(torch.bfloat16)
    final_norm = (1.0 + 0.1 * torch.randn(cfg.hidden, generator=g)).to(torch.bfloat16)
    cos, sin = rope_tables(cfg.head_dim, cfg.max_seq, cfg.rope_theta)
    return {"layers": layers, "embed": embed, "final_norm": final_norm, "rope_cos": cos, "rope_sin": sin}
