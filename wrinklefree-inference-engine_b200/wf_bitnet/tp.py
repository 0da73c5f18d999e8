"""Tensor-parallel host layer (north-star item 6): weight sharding and the one-off IPC handle exchange.

One process per GPU (torchrun).  torch.distributed is plumbing only: it carries the 64-byte CUDA IPC handles of the
ranks' exchange buffers once; the per-step all-gathers / all-reduces are peer-memory stores and REDs issued by the
attention and linear kernels themselves (csrc/tp_comm.cuh), not NCCL calls.

Sharding (Megatron pattern on the reference's per-layer matrices, scripts/direct_inference_full.py:95-140):
  q, k, v   : output rows, whole heads           -> rank r owns Q heads [r*H/T, (r+1)*H/T) and KV heads [r*Hkv/T, ...)
  gate, up  : output rows                        -> ffn rows [r*ffn/T, (r+1)*ffn/T)
  o, down   : input columns (I2_S bytes [k0/4, k1/4) of every row; 128-blocks never straddle a shard)
  norms, embedding / tied LM head, rope tables   : replicated (the LM head READS only its vocab slice)
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np


def shard_rows(packed: np.ndarray, rank: int, size: int) -> np.ndarray:
    n = packed.shape[0]
    if n % size:
        raise ValueError(f"rows ({n}) not divisible by tp_size {size}")
    s = n // size
    return np.ascontiguousarray(packed[rank * s:(rank + 1) * s])


def shard_cols(packed: np.ndarray, rank: int, size: int) -> np.ndarray:
    """packed: I2_S u8 [N, K/4]; the K axis is cut into `size` contiguous pieces of whole 128-blocks (32 bytes)."""
    k4 = packed.shape[1]
    if k4 % (32 * size):
        raise ValueError(f"in_features ({k4 * 4}) must be divisible by 128 * tp_size ({128 * size})")
    s = k4 // size
    return np.ascontiguousarray(packed[:, rank * s:(rank + 1) * s])


def shard_weights(weights: Dict, rank: int, size: int) -> Dict:
    """Shard a full-model weight dict (the layout of wf_bitnet.synth.make_weights / oracle.model) for one rank."""
    if size == 1:
        return weights
    out = {k: v for k, v in weights.items() if k != "layers"}
    out["layers"] = []
    for layer in weights["layers"]:
        sl = dict(layer)
        for name in ("q", "k", "v", "gate", "up"):
            sl[f"{name}_proj"] = shard_rows(np.asarray(layer[f"{name}_proj"]), rank, size)
        for name in ("o", "down"):
            sl[f"{name}_proj"] = shard_cols(np.asarray(layer[f"{name}_proj"]), rank, size)
        out["layers"].append(sl)
    return out


def vocab_slice(vocab: int, rank: int, size: int):
    """The rows of the tied LM head rank `rank` scores (whole 128-row CTAs per rank; mirrors engine.cu)."""
    s = (((vocab + size - 1) // size) + 127) // 128 * 128
    lo = min(rank * s, vocab)
    return lo, min(s, vocab - lo)


def exchange_handles(local: bytes, group=None) -> List[bytes]:
    """all-gather of the ranks' IPC handles, in rank order (works on gloo and nccl groups)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    out: List[Optional[bytes]] = [None] * world
    dist.all_gather_object(out, local, group=group)
    return [bytes(h) for h in out]


def attach(engine, group=None) -> None:
    """Export this rank's exchange buffer, gather everybody's handle, open the peers'.  Collective."""
    import torch.distributed as dist
    if engine.tp_size == 1:
        return
    if dist.get_world_size(group) != engine.tp_size or dist.get_rank(group) != engine.tp_rank:
        raise ValueError("process group does not match the engine's tp_rank / tp_size")
    engine.tp_attach(exchange_handles(engine.tp_export(), group))
    dist.barrier(group)
