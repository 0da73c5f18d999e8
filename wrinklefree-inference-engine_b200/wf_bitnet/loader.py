"""Checkpoint loader for the decode engine.

The reference loads one `model.safetensors` and repacks every projection with Python loops
(`DirectBitNetInference._preprocess_weights`, scripts/direct_inference_full.py:116-163: HF `[N/4, K]` u8 ->
unpack -> blocked I2_S `[N, K/4]`); its cache layer stores the repacked form as `packed_bitnet_v1`
(`{name}.qweight` + `{name}.weight_scale`, cache/bitnet_converter.py:165-185).  Here the three on-disk forms

  * HF checkpoint      `{p}.weight`  u8 `[N/4, K]` (blocked along N) + `{p}.weight_scale`
  * packed_bitnet_v1   `{p}.qweight` u8 `[N, K/4]` (blocked I2_S)    + `{p}.weight_scale`
  * ternary float      `{p}.weight`  fp `[N, K]` in {-1, 0, +1}      (+ optional `{p}.weight_scale`)

go through one device kernel each (`wf_hf_to_i2s` / `wf_pack_ternary`) and land in the engine, which fuses q|k|v and
gate|up and repacks to `b200_tile_v1` in `wf_engine_finalize_weights`.  No CPU compute path: conversion needs the GPU.
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Dict, Optional

import torch

from . import ops
from ._lib import check, lib

_ATTN = ("q", "k", "v", "o")
_MLP = ("gate", "up", "down")
_NORMS = {
    "input_layernorm": "input_layernorm.weight",
    "post_attention_layernorm": "post_attention_layernorm.weight",
    "attn_sub_norm": "self_attn.attn_sub_norm.weight",
    "ffn_sub_norm": "mlp.ffn_sub_norm.weight",
}


def _proj_key(layer: int, name: str) -> str:
    block = "self_attn" if name in _ATTN else "mlp"
    return f"model.layers.{layer}.{block}.{name}_proj"


def _i2s_and_scale(sd: Dict[str, torch.Tensor], base: str, N: int, K: int):
    """-> (device u8 [N, K/4] blocked I2_S, python float weight_scale) whatever the stored form."""
    scale_t = sd.get(f"{base}.weight_scale")
    if f"{base}.qweight" in sd:                                   # packed_bitnet_v1
        q = sd[f"{base}.qweight"]
        if tuple(q.shape) != (N, K // 4) or q.dtype != torch.uint8:
            raise ValueError(f"{base}.qweight: expected uint8 [{N}, {K // 4}], got {q.dtype} {tuple(q.shape)}")
        packed = q.to("cuda").contiguous()
    else:
        w = sd.get(f"{base}.weight")
        if w is None:
            raise KeyError(f"{base}.weight / .qweight missing from the checkpoint")
        if w.dtype == torch.uint8:                                # HF checkpoint [N/4, K]
            if tuple(w.shape) != (N // 4, K):
                raise ValueError(f"{base}.weight: expected uint8 [{N // 4}, {K}], got {tuple(w.shape)}")
            packed = ops.hf_to_i2s(w.to("cuda"))
        else:                                                     # ternary float [N, K]
            if tuple(w.shape) != (N, K):
                raise ValueError(f"{base}.weight: expected [{N}, {K}], got {tuple(w.shape)}")
            wd = w.to("cuda").float()
            # the reference converter only packs tensors that ARE ternary (_is_ternary_float, cache/bitnet_converter.py:80-103);
            # a bf16 master-weight checkpoint would otherwise be clipped to {-1,0,+1} silently and decode garbage
            if not bool(((wd == -1) | (wd == 0) | (wd == 1)).all().item()):
                raise ValueError(f"{base}.weight: float tensor is not ternary {{-1, 0, +1}} — quantise the checkpoint first "
                                 f"(this loader packs, it does not quantise; see cache/bitnet_converter.py:80-103)")
            packed, s = ops.pack_ternary(wd)
            if scale_t is None:
                scale_t = s
    if scale_t is None:
        raise KeyError(f"{base}.weight_scale missing from the checkpoint")
    return packed, float(scale_t.reshape(-1)[0].item())


def _shard(packed: torch.Tensor, name: str, rank: int, size: int) -> torch.Tensor:
    """This rank's piece of a full I2_S matrix (wf_bitnet/tp.py: q/k/v/gate/up by output rows, o/down by input columns in
    whole 128-blocks = 32 packed bytes)."""
    if size == 1:
        return packed
    if name in ("o", "down"):
        k4 = packed.shape[1]
        if k4 % (32 * size):
            raise ValueError(f"{name}_proj: in_features ({k4 * 4}) must be divisible by 128 * tp_size ({128 * size})")
        s = k4 // size
        return packed[:, rank * s:(rank + 1) * s].contiguous()
    n = packed.shape[0]
    if n % size:
        raise ValueError(f"{name}_proj: rows ({n}) not divisible by tp_size {size}")
    s = n // size
    return packed[rank * s:(rank + 1) * s].contiguous()


def load_state_dict(engine, sd: Dict[str, torch.Tensor]) -> None:
    """Fill an engine from an HF-named state dict (any of the three forms above, mixed freely).  A tensor-parallel engine
    (tp_size > 1) takes the FULL state dict on every rank and keeps its own shard; call wf_bitnet.tp.attach afterwards."""
    cfg = engine.cfg
    shapes = {
        "q": (cfg.heads * cfg.head_dim, cfg.hidden), "k": (cfg.kv_heads * cfg.head_dim, cfg.hidden),
        "v": (cfg.kv_heads * cfg.head_dim, cfg.hidden), "o": (cfg.hidden, cfg.heads * cfg.head_dim),
        "gate": (cfg.ffn, cfg.hidden), "up": (cfg.ffn, cfg.hidden), "down": (cfg.hidden, cfg.ffn),
    }
    for li in range(cfg.layers):
        for name, (N, K) in shapes.items():
            packed, scale = _i2s_and_scale(sd, _proj_key(li, name), N, K)
            packed = _shard(packed, name, engine.tp_rank, engine.tp_size)
            check(lib().wf_engine_set_linear(engine._h, li, name.encode(), packed.data_ptr(), scale, 0))
            torch.cuda.synchronize()   # `packed` may be freed by the next iteration; the engine staged its own copy
        for ours, theirs in _NORMS.items():
            engine._set_tensor(li, ours, sd[f"model.layers.{li}.{theirs}"])
    engine._set_tensor(-1, "embed", sd["model.embed_tokens.weight"])      # tied LM head (direct_inference_full.py:124-126)
    engine._set_tensor(-1, "final_norm", sd["model.norm.weight"])
    check(lib().wf_engine_finalize_weights(engine._h))


def load_safetensors_dir(engine, path, expect_format: Optional[str] = None) -> Dict:
    """Load every `*.safetensors` shard under `path` (an HF snapshot or a packed_bitnet_v1 cache entry written by
    wrinklefree_inference.cache.bitnet_converter.convert_and_save_bitnet).  Returns the cache metadata ({} if none)."""
    from safetensors.torch import load_file
    path = Path(path)
    shards = sorted(path.glob("*.safetensors"))
    if not shards:
        raise FileNotFoundError(f"no *.safetensors under {path}")
    meta = {}
    mf = path / "cache_metadata.json"
    if mf.exists():
        meta = json.loads(mf.read_text())
    if expect_format is not None and meta.get("format_version") != expect_format:
        raise ValueError(f"{path}: cache format {meta.get('format_version')!r}, expected {expect_format!r}")
    sd: Dict[str, torch.Tensor] = {}
    for s in shards:
        sd.update(load_file(str(s)))
    load_state_dict(engine, sd)
    return meta
