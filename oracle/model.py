"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference decode forward.

Follows ``scripts/direct_inference_full.py`` of the reference line by line (torch CPU ops
for the floating-point glue, exactly as the reference uses them; the integer GEMV comes
from ``oracle/wf_oracle.c`` or, when built, the reference's own SIMD kernel in
``oracle/_ref``).  Also restates ``kv_cache/kv_cache.py::attention_with_kv_cache`` and the
paged KV manager semantics of ``tests/test_kv_cache_cpp.py:20-127``.

Attention parity note: the reference holds no test that pins attention VALUES
 — the function itself, run in fp32/bf16 on CPU, is the oracle.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

import oracle as O


@dataclass
class ModelConfig:
    hidden: int = 2560
    layers: int = 30
    heads: int = 20
    kv_heads: int = 5
    head_dim: int = 128
    ffn: int = 6912
    vocab: int = 128256
    rope_theta: float = 500000.0
    eps: float = 1e-5
    max_seq: int = 4096


class RotaryEmbedding:
    """scripts/direct_inference_full.py:28-70 (half-split RoPE, bf16 sin/cos cache)."""

    def __init__(self, head_dim: int, max_seq_len: int, base: float):
        self.head_dim = head_dim
        inv_freq = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
        positions = torch.arange(max_seq_len, dtype=torch.float32)
        freqs = torch.outer(positions, inv_freq)
        self.cos_cache = torch.cos(freqs).to(torch.bfloat16)
        self.sin_cache = torch.sin(freqs).to(torch.bfloat16)

    def rotate(self, x: torch.Tensor, pos: int) -> torch.Tensor:
        cos = self.cos_cache[pos]
        sin = self.sin_cache[pos]
        half = self.head_dim // 2
        x1 = x[..., :half]
        x2 = x[..., half:]
        return torch.cat([x1 * cos - x2 * sin, x1 * sin + x2 * cos], dim=-1)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """direct_inference_full.py:181-184."""
    variance = x.float().pow(2).mean(-1, keepdim=True)
    return (x * torch.rsqrt(variance + eps).to(x.dtype)) * weight


def bitnet_linear(x: torch.Tensor, packed: np.ndarray, scale: float, use_ref: bool = False) -> torch.Tensor:
    """direct_inference_full.py:186-208. x: bf16/fp32 [K] -> bf16 [N]."""
    xf = x.float().reshape(-1).numpy()
    q, s_a = O.quantize_act_i8(xf)
    if use_ref and O.ref_available():
        out = torch.from_numpy(O.ref_gemv(packed, q, float(scale)))
    else:
        out = torch.from_numpy(O.gemv(packed, q, float(scale)))
    act_sum = float(q.astype(np.float32).sum())  # x_int8.float().sum().item()
    out = (out - scale * act_sum) * s_a
    return out.to(torch.bfloat16)


def kv_int8_quant(x: torch.Tensor):
    """KVCache._quantize_to_int8 (kv_cache/kv_cache.py:198-218) applied per (token, head) row instead of once per update call:
    scale = 127 / absmax (absmax < 1e-6 -> 1), codes = clamp(round(x * scale), -127, 127), stored scale = absmax / 127.
    x: [..., D] (bf16 or fp32).  Returns (int8 codes, fp32 dequant scales [..., 1]).  fp32 arithmetic, round-half-even."""
    xf = x.float()
    amax = xf.abs().amax(dim=-1, keepdim=True)
    amax = torch.where(amax < 1e-6, torch.ones_like(amax), amax)
    inv = torch.tensor(127.0, dtype=torch.float32) / amax
    codes = (xf * inv).round().clamp(-127, 127).to(torch.int8)
    return codes, amax / torch.tensor(127.0, dtype=torch.float32)


def kv_int8_roundtrip(x: torch.Tensor) -> torch.Tensor:
    """What KVCache.get returns for an INT8 cache: (codes.float() * scale).to(bf16)  (kv_cache.py:220-228)."""
    codes, sc = kv_int8_quant(x)
    return (codes.float() * sc).to(torch.bfloat16)


class OracleModel:
    """DirectBitNetInference (direct_inference_full.py:73-302) on caller-supplied weights.

    ``weights``: {"embed": bf16 [V,H], "final_norm": bf16 [H], "layers": [ {
        "input_layernorm","post_attention_layernorm","attn_sub_norm","ffn_sub_norm": bf16,
        "<p>_proj": u8 I2_S [N,K/4], "<p>_scale": float  for p in q,k,v,o,gate,up,down } ]}
    """

    def __init__(self, cfg: ModelConfig, weights: Dict, use_ref: bool = True, kv_int8: bool = False, attn_fp32: bool = False):
        self.cfg = cfg
        self.w = weights
        self.use_ref = use_ref
        self.kv_int8 = kv_int8   # KVCacheDtype.INT8: what the cache returns is dequant(quant(x)), see kv_int8_roundtrip
        # attn_fp32: the same attention evaluated in exact fp32 math and rounded to bf16 once, instead of torch's bf16 SDPA kernel.
        # Both are "the reference within bf16 tolerance"; running one against the other measures how far a bf16-level change of
        # the attention arithmetic moves the logits / greedy tokens of a deep random-init model (test sensitivity budget).
        self.attn_fp32 = attn_fp32
        self.taps: Dict = {}     # ("q"|"k"|"v"|"attn", layer) of the last forward: q, k (after RoPE) / v before the cache, attention output
        self.rope = RotaryEmbedding(cfg.head_dim, cfg.max_seq, cfg.rope_theta)
        self.k_cache = torch.zeros(cfg.layers, cfg.max_seq, cfg.kv_heads, cfg.head_dim, dtype=torch.bfloat16)
        self.v_cache = torch.zeros_like(self.k_cache)

    def _lin(self, x, layer, name):
        return bitnet_linear(x, layer[f"{name}_proj"], layer[f"{name}_scale"], self.use_ref)

    def attention(self, hidden, li, pos, layer):
        c = self.cfg
        q = self._lin(hidden, layer, "q").view(c.heads, c.head_dim)
        k = self._lin(hidden, layer, "k").view(c.kv_heads, c.head_dim)
        v = self._lin(hidden, layer, "v").view(c.kv_heads, c.head_dim)
        q = self.rope.rotate(q, pos)
        k = self.rope.rotate(k, pos)
        self.taps[("q", li)], self.taps[("k", li)], self.taps[("v", li)] = q, k, v
        if self.kv_int8:
            k, v = kv_int8_roundtrip(k), kv_int8_roundtrip(v)
        self.k_cache[li, pos] = k
        self.v_cache[li, pos] = v
        k_seq = self.k_cache[li, : pos + 1]
        v_seq = self.v_cache[li, : pos + 1]
        q_attn = q.unsqueeze(0).unsqueeze(2)
        k_attn = k_seq.transpose(0, 1).unsqueeze(0)
        v_attn = v_seq.transpose(0, 1).unsqueeze(0)
        if self.attn_fp32:
            attn_out = attention_gqa_fp32(q, k_seq, v_seq).to(torch.bfloat16).flatten()
        else:
            attn_out = F.scaled_dot_product_attention(q_attn, k_attn, v_attn, enable_gqa=True)
            attn_out = attn_out.squeeze(0).squeeze(1).flatten()
        self.taps[("attn", li)] = attn_out
        attn_out = rms_norm(attn_out, layer["attn_sub_norm"], c.eps)
        return self._lin(attn_out, layer, "o")

    def mlp(self, hidden, layer):
        gate = self._lin(hidden, layer, "gate")
        up = self._lin(hidden, layer, "up")
        act = torch.relu(gate).square() * up
        act = rms_norm(act, layer["ffn_sub_norm"], self.cfg.eps)
        return self._lin(act, layer, "down")

    def forward_layer(self, hidden, li, pos):
        layer = self.w["layers"][li]
        normed = rms_norm(hidden, layer["input_layernorm"], self.cfg.eps)
        hidden = hidden + self.attention(normed, li, pos, layer)
        normed = rms_norm(hidden, layer["post_attention_layernorm"], self.cfg.eps)
        hidden = hidden + self.mlp(normed, layer)
        return hidden

    def forward(self, token_id: int, pos: int, return_hidden: bool = False):
        hidden = self.w["embed"][token_id]
        for li in range(self.cfg.layers):
            hidden = self.forward_layer(hidden, li, pos)
        hidden = rms_norm(hidden, self.w["final_norm"], self.cfg.eps)
        logits = F.linear(hidden, self.w["embed"])
        if return_hidden:
            return logits, hidden
        return logits

    def generate_greedy(self, prompt: List[int], n_new: int):
        """generate() (:328-371) with temperature 0: token-by-token prefill then argmax decode.
        Note the reference feeds the LAST PROMPT TOKEN again at pos=len(prompt) (:349-353)."""
        for pos, t in enumerate(prompt):
            self.forward(int(t), pos)
        out, gaps = [], []
        pos = len(prompt)
        nxt = int(prompt[-1])
        for _ in range(n_new):
            logits = self.forward(nxt, pos).float()
            top2 = torch.topk(logits, 2)
            nxt = int(top2.indices[0])
            gaps.append(float(top2.values[0] - top2.values[1]))
            out.append(nxt)
            pos += 1
        return out, gaps


# ------------------------------------------------------------------------------ attention oracles
def attention_mha_reference(query, keys, values, scale: Optional[float] = None):
    """kv_cache/kv_cache.py:316-369 arithmetic after the cache update/get:
    query [B,H,1,D]; keys/values [B,S,H,D] (already in query dtype). Returns [B,H,1,D]."""
    key = keys.transpose(1, 2)
    value = values.transpose(1, 2)
    if scale is None:
        scale = 1.0 / math.sqrt(query.shape[-1])
    w = torch.matmul(query, key.transpose(-2, -1)) * scale
    w = torch.softmax(w, dim=-1)
    return torch.matmul(w, value)


def attention_gqa_reference(q, k_seq, v_seq):
    """direct_inference_full.py:229-244. q [H,D]; k_seq,v_seq [S,Hkv,D]. Returns [H,D]."""
    q_attn = q.unsqueeze(0).unsqueeze(2)
    k_attn = k_seq.transpose(0, 1).unsqueeze(0)
    v_attn = v_seq.transpose(0, 1).unsqueeze(0)
    out = F.scaled_dot_product_attention(q_attn, k_attn, v_attn, enable_gqa=True)
    return out.squeeze(0).squeeze(1)


def attention_gqa_fp32(q, k_seq, v_seq, scale: Optional[float] = None):
    """Exact-math (fp32) form of the same attention, used for tolerance budgeting."""
    H, D = q.shape
    Hkv = k_seq.shape[1]
    g = H // Hkv
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    qf = q.float().view(Hkv, g, D)
    kf = k_seq.float().permute(1, 0, 2)  # [Hkv,S,D]
    vf = v_seq.float().permute(1, 0, 2)
    s = torch.einsum("hgd,hsd->hgs", qf, kf) * scale
    p = torch.softmax(s, dim=-1)
    return torch.einsum("hgs,hsd->hgd", p, vf).reshape(H, D)


class PagedKVReference:
    """PythonKVCacheReference semantics (tests/test_kv_cache_cpp.py:20-127), extended with real
    slots: storage [max_pages, L, 2, page_size, H, D].  With all slots == 0 it behaves exactly
    like the reference's (slot-ignoring) Python class."""

    def __init__(self, num_layers, num_heads, head_dim, page_size, max_pages, dtype=torch.float32):
        self.num_layers, self.num_heads, self.head_dim = num_layers, num_heads, head_dim
        self.page_size, self.max_pages = page_size, max_pages
        self.store = torch.zeros(max_pages, num_layers, 2, page_size, num_heads, head_dim, dtype=dtype)
        self._free = list(range(max_pages - 1, -1, -1))

    def num_free_pages(self):
        return len(self._free)

    def allocate_pages(self, n):
        if len(self._free) < n:
            return torch.empty(0, dtype=torch.int32)
        return torch.tensor([self._free.pop() for _ in range(n)], dtype=torch.int32)

    def free_pages(self, pages):
        for p in pages.tolist():
            if 0 <= p < self.max_pages:
                self._free.append(p)

    def scatter_kv(self, k_in, v_in, pages, slots, layer_id):
        T = pages.shape[0]
        k_in = k_in.reshape(T, self.num_heads, self.head_dim)
        v_in = v_in.reshape(T, self.num_heads, self.head_dim)
        for t in range(T):
            self.store[int(pages[t]), layer_id, 0, int(slots[t])] = k_in[t]
            self.store[int(pages[t]), layer_id, 1, int(slots[t])] = v_in[t]

    def gather_kv(self, k_out, v_out, pages, slots, layer_id):
        T = pages.shape[0]
        ko = k_out.view(T, self.num_heads, self.head_dim)
        vo = v_out.view(T, self.num_heads, self.head_dim)
        for t in range(T):
            ko[t] = self.store[int(pages[t]), layer_id, 0, int(slots[t])]
            vo[t] = self.store[int(pages[t]), layer_id, 1, int(slots[t])]
