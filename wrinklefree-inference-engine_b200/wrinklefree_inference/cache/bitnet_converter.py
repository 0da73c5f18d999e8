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


def convert_state_dict(weights: Dict[str, torch.Tensor], all_keys=None) -> Dict[str, torch.Tensor]:
    """The per-tensor loop of convert_and_save_bitnet (:150-163): ternary float BitNet linears become
    `.qweight` + `.weight_scale`; HF pre-packed uint8 [N/4,K] linears become `.qweight` (+ their stored
    `.weight_scale`); everything else passes through.  A `.weight_scale` the checkpoint already stores (in this shard or, via
    `all_keys`, in another one) wins over the max|w| = 1.0 the packer computes for an exactly ternary tensor."""
    from wf_bitnet import ops
    known = set(weights) if all_keys is None else set(all_keys)
    out: Dict[str, torch.Tensor] = {}
    for name, tensor in weights.items():
        is_bitnet = any(name.endswith(p) for p in BITNET_WEIGHT_PATTERNS)
        base = name[: -len(".weight")] if name.endswith(".weight") else name
        if is_bitnet and _is_ternary_float(tensor):
            packed, scale = ops.pack_ternary(tensor.to("cuda").float())
            out[f"{base}.qweight"] = packed
            if f"{base}.weight_scale" not in known:
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


def convert_and_save_bitnet(source_model_path, output_path, revision=None, pack_format: str = "blocked_uint8_v1"):
    """convert_and_save_bitnet (reference :103-187) for a LOCAL model directory: every `*.safetensors` shard is rewritten
    with the BitNet linears as `.qweight` + `.weight_scale` (packing on the GPU), config / tokenizer files are copied and
    `cache_metadata.json` is written with the reference's keys plus the `pack_format` tag of cache/cache_key.py:24-28.
    `pack_format="b200_tile_v1"` additionally stores `{name}.tile`.  Downloading from the Hub stays with the reference."""
    import json
    import shutil
    from pathlib import Path

    from safetensors.torch import load_file, save_file

    model_dir = Path(source_model_path)
    if not model_dir.exists():
        raise FileNotFoundError(f"{source_model_path}: only local model directories are converted here (no network)")
    if pack_format not in ("blocked_uint8_v1", B200_PACK_FORMAT):
        raise ValueError(f"unknown pack_format {pack_format!r}")
    output_path = Path(output_path)
    output_path.mkdir(parents=True, exist_ok=True)
    for config_file in ["config.json", "tokenizer.json", "tokenizer_config.json", "special_tokens_map.json", "tokenizer.model",
                        "generation_config.json"]:
        src = model_dir / config_file
        if src.exists():
            shutil.copy(src, output_path / config_file)
    from safetensors import safe_open
    all_keys = set()
    for wf in sorted(model_dir.glob("*.safetensors")):
        with safe_open(str(wf), framework="pt") as f:
            all_keys.update(f.keys())
    for wf in sorted(model_dir.glob("*.safetensors")):
        converted = convert_state_dict(load_file(str(wf)), all_keys)
        if pack_format == B200_PACK_FORMAT:
            converted = pack_state_for_b200(converted)
        save_file({k: v.detach().cpu().contiguous() for k, v in converted.items()}, str(output_path / wf.name))
    meta = {"format_version": CACHE_FORMAT_VERSION, "source_model": str(source_model_path), "revision": revision, "pack_format": pack_format}
    (output_path / "cache_metadata.json").write_text(json.dumps(meta, indent=2))
    return output_path
