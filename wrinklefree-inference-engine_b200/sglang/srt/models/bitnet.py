"""`sglang.srt.models.bitnet` of the reference's fork on the B200 kernels.

The fork's file is not in the scrape; its surface is pinned by the reference's own tests and scripts:
  * `_pack_ternary_weights(w) -> (u8 [N, K/4] blocked I2_S, scale[1])`      tests/test_vllm_cpu_stub.py:234-247
  * `_unpack_i2_to_ternary(packed, N, K) -> f32 [N, K]`                      tests/test_bitnet_comparison.py:14-36
  * `_unpack_ternary_weights(hf u8 [N/4, K]) -> f32 [N, K]`                  tests/test_bitnet_weight_loading.py:20-58
  * `_is_packed_weight`, `_is_ternary_float`                                 tests/test_vllm_cpu_stub.py:269-295
  * `BitNetLinear(in_features, out_features, bias=False)` with `.qweight` / `.weight_scale`, forward = per-tensor int8
    absmax quant -> ternary GEMV/GEMM -> (y - s_w * sum a) * s_a             tests/test_bitnet_weight_loading.py:91-135,
                                                                             scripts/direct_inference_full.py:186-208
  * `BitNetForCausalLM(config)` + `load_weights(iterable of (name, tensor))`  scripts/direct_bitnet_inference.py:25
Every function runs on the GPU through the C ABI; CPU tensors are moved to the device and the result comes back on the
caller's device (the reference's tests hand in CPU tensors).  There is no CPU compute path: without the CUDA library the
first call raises.
"""
from __future__ import annotations

from typing import Iterable, Optional, Tuple

import torch
from torch import nn

BITNET_BLOCK_SIZE = 128


def _ops():
    from wf_bitnet import ops
    return ops


def _dev(t: torch.Tensor) -> torch.Tensor:
    return t if t.is_cuda else t.to("cuda")


def _pack_ternary_weights(weight: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Ternary {-1,0,+1} float -> blocked I2_S u8 [N, K/4] + scale[1] (= max|w|, 1.0 for a ternary tensor), the arithmetic of
    quantize_i2_s (extern/BitNet/src/ggml-bitnet-mad.cpp:46-92) / _pack_ternary_weights (tests/test_bitnet_core.py:56-106):
    code = round(w) clamped to {-1,0,+1}, + 1.  It packs; it does not quantise master weights."""
    packed, scale = _ops().pack_ternary(_dev(weight.detach()))
    return packed.to(weight.device), scale.to(weight.device)


def _unpack_i2_to_ternary(packed: torch.Tensor, out_features: int, in_features: int) -> torch.Tensor:
    if tuple(packed.shape) != (out_features, in_features // 4):
        raise ValueError(f"packed must be [{out_features}, {in_features // 4}], got {tuple(packed.shape)}")
    return _ops().unpack_i2s(_dev(packed)).to(packed.device)


def _unpack_ternary_weights(packed_hf: torch.Tensor) -> torch.Tensor:
    """HF checkpoint u8 [N/4, K] (4 row blocks in the bit pairs of one byte) -> f32 [N, K]."""
    o = _ops()
    return o.unpack_i2s(o.hf_to_i2s(_dev(packed_hf))).to(packed_hf.device)


def _is_packed_weight(t: torch.Tensor) -> bool:
    return t.dtype == torch.uint8


def _is_ternary_float(t: torch.Tensor) -> bool:
    if not t.is_floating_point():
        return False
    return bool(((t == -1) | (t == 0) | (t == 1)).all().item())


class BitNetLinear(nn.Module):
    """Ternary linear. `qweight` (u8 [out, in/4], blocked I2_S) and `weight_scale` ([1]) are filled after construction,
    the way the reference's tests and loader do (`linear.qweight.data.copy_(packed)`), so forward reads `qweight` in place;
    `finalize()` snapshots it into the `b200_tile_v1` layout for the streaming GEMV (call again after changing qweight)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False, device=None):
        super().__init__()
        if in_features % BITNET_BLOCK_SIZE != 0:
            raise ValueError(f"in_features ({in_features}) must be multiple of {BITNET_BLOCK_SIZE}")
        dev = torch.device(device) if device is not None else torch.device("cuda")
        self.in_features, self.out_features = in_features, out_features
        self.qweight = nn.Parameter(torch.full((out_features, in_features // 4), 0x55, dtype=torch.uint8, device=dev),
                                    requires_grad=False)           # 0x55 = all zeros (tests/test_bitnet_core.py:186-203)
        self.weight_scale = nn.Parameter(torch.ones(1, dtype=torch.float32, device=dev), requires_grad=False)
        self.bias = nn.Parameter(torch.zeros(out_features, dtype=torch.float32, device=dev), requires_grad=False) if bias else None
        self._tile: Optional[torch.Tensor] = None

    def finalize(self) -> "BitNetLinear":
        self._tile = _ops().i2s_to_tile(_dev(self.qweight.data))
        return self

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        o = _ops()
        src = x.device
        xf = _dev(x).reshape(-1, self.in_features)
        q, act_scale, rowsum = o.quantize_activations(xf, per_token=False)
        if self._tile is not None:
            w, layout = self._tile, "tile"
        else:
            w, layout = _dev(self.qweight.data).contiguous(), "i2s"
        out = o.ternary_linear(w, self.out_features, self.in_features, q, layout=layout, mode="linear", rowsum=rowsum,
                               act_scale=act_scale, w_scale=_dev(self.weight_scale.data).float().reshape(-1)[:1],
                               out_dtype=torch.float32)
        if self.bias is not None:
            out = out + _dev(self.bias.data)
        return out.view(*x.shape[:-1], self.out_features).to(src)


def _cfg_get(config, name, default=None):
    if isinstance(config, dict):
        return config.get(name, default)
    return getattr(config, name, default)


class BitNetForCausalLM(nn.Module):
    """The fork's model class on the CUDA-graphed decode engine. `config` is the HF config (object or dict:
    hidden_size, num_hidden_layers, num_attention_heads, num_key_value_heads, intermediate_size, vocab_size, rope_theta,
    rms_norm_eps, max_position_embeddings).  The engine owns KV paging and positions, so `forward(input_ids)` is one decode
    step for `len(input_ids)` running sequences (slot i = sequence i) and returns their greedy next ids; `logits()` exposes
    the fp32 logits of that step."""

    def __init__(self, config, quant_config=None, prefix: str = "", *, max_batch: int = 1, max_seq: Optional[int] = None,
                 page_size: int = 16):
        super().__init__()
        from wf_bitnet.engine import Engine, EngineConfig
        hidden = int(_cfg_get(config, "hidden_size"))
        heads = int(_cfg_get(config, "num_attention_heads"))
        self.config = config
        self.engine_config = EngineConfig(
            hidden=hidden, layers=int(_cfg_get(config, "num_hidden_layers")), heads=heads,
            kv_heads=int(_cfg_get(config, "num_key_value_heads", heads)), head_dim=int(_cfg_get(config, "head_dim", None) or hidden // heads),
            ffn=int(_cfg_get(config, "intermediate_size")), vocab=int(_cfg_get(config, "vocab_size")),
            rope_theta=float(_cfg_get(config, "rope_theta", 500000.0)), eps=float(_cfg_get(config, "rms_norm_eps", 1e-5)),
            max_seq=int(max_seq or _cfg_get(config, "max_position_embeddings", 4096)), page_size=page_size, max_batch=max_batch)
        self.engine = Engine(self.engine_config)
        self._last_batch = 0

    def load_weights(self, weights: Iterable[Tuple[str, torch.Tensor]]) -> None:
        from wf_bitnet import loader
        sd = dict(weights.items()) if isinstance(weights, dict) else dict(weights)
        loader.load_state_dict(self.engine, sd)

    def reset(self) -> None:
        self.engine.reset()

    @torch.no_grad()
    def forward(self, input_ids, positions=None, forward_batch=None) -> torch.Tensor:
        ids = input_ids.detach().cpu().tolist() if isinstance(input_ids, torch.Tensor) else list(input_ids)
        self._last_batch = len(ids)
        return torch.from_numpy(self.engine.decode(ids)).long()

    def logits(self) -> torch.Tensor:
        return self.engine.buffer("logits", (self._last_batch, self.engine_config.vocab), torch.float32)

    def generate(self, prompt_ids, max_new_tokens: int):
        return self.engine.generate_greedy(list(prompt_ids), max_new_tokens)


EntryClass = BitNetForCausalLM
