"""packed_bitnet_v1 converter on the GPU (reference: src/wrinklefree_inference/cache/bitnet_converter.py).

The reference rewrites safetensors shards: ternary float weights -> `{name}.qweight` u8 [N,K/4] (blocked
I2_S) + `{name}.weight_scale` f32 [1] (:165-176), with a Python double loop over blocks.  Here the same
mapping is a state-dict transform whose packing runs in one kernel (wf_pack_ternary / wf_hf_to_i2s), plus
the second cached artefact the B200 path wants: the b200_tile_v1 repack (`pack_format: "b200_tile_v1"`).
File IO / HF download / GCS stay with the reference's own code.
"""
from __future__ import annotations

from typing import Dict

import torch

BITNET_WEIGHT_PATTERNS = ["q_proj.weight", "k_proj.weight", "v_proj.weight", "o_proj.weight",
                          "gate_proj.weight", "up_proj.weight", "down_proj.weight"]

CACHE_FORMAT_VERSION = "packed_bitnet_v1"
B200_PACK_FORMAT = "b200_tile_v1"


def _is_ternary_float(weight: torch.Tensor) -> bool:
    """Reference :80-100."""
    if weight.ndim != 2 or weight.dtype not in (torch.float16, torch.bfloat16, torch.float32):
        return False
    unique = torch.unique(torch.round(weight.float()))
    if len(unique) > 3 or unique.abs().max() > 1.0:
        return False
    return weight.shape[1] % 4 == 0


def convert_state_dict(weights: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """The per-tensor loop of convert_and_save_bitnet (:150-163): ternary float BitNet linears become
    `.qweight` + `.weight_scale`; HF pre-packed uint8 [N/4,K] linears become `.qweight` (+ their stored
    `.weight_scale`); everything else passes through."""
    from wf_bitnet import ops
    out: Dict[str, torch.Tensor] = {}
    for name, tensor in weights.items():
        is_bitnet = any(p in name for p in BITNET_WEIGHT_PATTERNS)
        base = name.replace(".weight", "")
        if is_bitnet and _is_ternary_float(tensor):
            packed, scale = ops.pack_ternary(tensor.to("cuda").float())
            out[f"{base}.qweight"] = packed
            out[f"{base}.weight_scale"] = scale
        elif is_bitnet and tensor.dtype == torch.uint8 and tensor.ndim == 2:
            out[f"{base}.qweight"] = ops.hf_to_i2s(tensor.to("cuda"))
        else:
            out[name] = tensor
    return out


def pack_state_for_b200(converted: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """packed_bitnet_v1 state dict -> adds `{name}.tile` (b200_tile_v1) next to every `.qweight`."""
    from wf_bitnet import ops
    out = dict(converted)
    for name, t in converted.items():
        if name.endswith(".qweight") and (t.shape[1] * 4) % 128 == 0:
            out[name[: -len(".qweight")] + ".tile"] = ops.i2s_to_tile(t.to("cuda"))
    return out
