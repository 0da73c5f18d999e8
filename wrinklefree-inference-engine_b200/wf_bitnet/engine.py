"""Python handle on the native decode engine (wf_engine_* in include/wf_bitnet.h).

Mirrors the reference's DirectBitNetInference (scripts/direct_inference_full.py:73-371): load weights,
keep a KV cache, `forward`/`generate` token by token — but batched, paged and CUDA-graph captured.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import ModelConfig as _CConfig, check, lib


@dataclass
class EngineConfig:
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
    page_size: int = 16
    max_batch: int = 1
    max_pages: int = 0   # 0 -> max_batch * max_seq / page_size
    kv_dtype: str = "bf16"   # "bf16" | "int8" (the reference's KVCacheDtype.INT8 at (token, head) granularity)

    @staticmethod
    def bitnet_2b(**kw) -> "EngineConfig":
        return EngineConfig(**kw)

    @staticmethod
    def llama3_8b(**kw) -> "EngineConfig":
        return EngineConfig(hidden=4096, layers=32, heads=32, kv_heads=8, ffn=14336, **kw)

    @staticmethod
    def llama3_70b(**kw) -> "EngineConfig":
        return EngineConfig(hidden=8192, layers=80, heads=64, kv_heads=8, ffn=28672, **kw)


LINEARS = ("q", "k", "v", "o", "gate", "up", "down")
NORMS = ("input_layernorm", "post_attention_layernorm", "attn_sub_norm", "ffn_sub_norm")


class Engine:
    def __init__(self, cfg: EngineConfig, tp_rank: int = 0, tp_size: int = 1):
        """tp_size > 1: this process holds shard tp_rank of a tensor-parallel engine; weights passed to load_weights
        must already be this rank's shard (wf_bitnet.tp.shard_weights) and wf_bitnet.tp.attach() must run before decode."""
        _lib.require_gpu()
        self.cfg = cfg
        self.tp_rank, self.tp_size = tp_rank, tp_size
        max_pages = cfg.max_pages or cfg.max_batch * (cfg.max_seq // cfg.page_size)
        c = _CConfig(cfg.hidden, cfg.layers, cfg.heads, cfg.kv_heads, cfg.head_dim, cfg.ffn, cfg.vocab, cfg.rope_theta, cfg.eps,
                     cfg.max_seq, cfg.page_size, cfg.max_batch, max_pages, tp_rank, tp_size,
                     {"bf16": _lib.WF_BF16, "int8": _lib.WF_I8}[cfg.kv_dtype])
        h = C.c_void_p()
        check(lib().wf_engine_create(C.byref(c), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().wf_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ weights
    def load_weights(self, weights: Dict) -> None:
        """weights: the dict layout of oracle.model.OracleModel / wf_bitnet.synth.make_weights
        (I2_S numpy matrices + python-float scales + bf16 torch tensors)."""
        for li, layer in enumerate(weights["layers"]):
            for name in LINEARS:
                m = np.ascontiguousarray(layer[f"{name}_proj"], dtype=np.uint8)
                check(lib().wf_engine_set_linear(self._h, li, name.encode(), m.ctypes.data, float(layer[f"{name}_scale"]), 1))
            for name in NORMS:
                self._set_tensor(li, name, layer[name])
        self._set_tensor(-1, "embed", weights["embed"])
        self._set_tensor(-1, "final_norm", weights["final_norm"])
        if "rope_cos" in weights:
            self._set_tensor(-1, "rope_cos", weights["rope_cos"])
            self._set_tensor(-1, "rope_sin", weights["rope_sin"])
        check(lib().wf_engine_finalize_weights(self._h))

    def _set_tensor(self, layer: int, name: str, t: torch.Tensor) -> None:
        t = t.detach().to(torch.bfloat16).contiguous()
        if t.is_cuda:
            check(lib().wf_engine_set_tensor(self._h, layer, name.encode(), t.data_ptr(), t.numel(), 0))
        else:
            check(lib().wf_engine_set_tensor(self._h, layer, name.encode(), t.data_ptr(), t.numel(), 1))

    # ------------------------------------------------------------------ tensor parallel
    def tp_export(self) -> bytes:
        n = int(lib().wf_engine_tp_handle_bytes())
        buf = C.create_string_buffer(n)
        check(lib().wf_engine_tp_export(self._h, buf))
        return buf.raw

    def tp_attach(self, handles: Sequence[bytes]) -> None:
        blob = b"".join(handles)
        check(lib().wf_engine_tp_attach(self._h, C.c_char_p(blob)))

    def init_synthetic(self, seed: int = 1234) -> None:
        check(lib().wf_engine_init_synthetic(self._h, seed))

    # ------------------------------------------------------------------ state
    def reset(self) -> None:
        check(lib().wf_engine_reset(self._h))

    def set_context(self, slot: int, length: int, fill_random: bool = False) -> None:
        check(lib().wf_engine_set_context(self._h, slot, length, 1 if fill_random else 0))

    def device_bytes(self) -> int:
        return int(lib().wf_engine_device_bytes(self._h))

    def launches_per_step(self) -> int:
        return int(lib().wf_engine_launches_per_step(self._h))

    def sync(self) -> None:
        check(lib().wf_engine_sync(self._h))

    @property
    def stream(self) -> int:
        return int(lib().wf_engine_stream(self._h) or 0)

    # ------------------------------------------------------------------ decode
    def decode(self, tokens: Sequence[int]) -> np.ndarray:
        """One step for len(tokens) sequences; HOST ids in, HOST greedy next ids out (the e2e call)."""
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.empty_like(t)
        check(lib().wf_engine_decode(self._h, t.ctypes.data, t.size, out.ctypes.data))
        return out

    def decode_device(self, batch: int, use_graph: bool = True) -> None:
        """Same step with device-resident tokens (sampled token fed back); asynchronous."""
        check(lib().wf_engine_decode_device(self._h, batch, 1 if use_graph else 0))

    def generate_greedy(self, prompt: Sequence[int], n_new: int) -> List[int]:
        """DirectBitNetInference.generate at temperature 0 (direct_inference_full.py:328-371): the prompt is
        processed token by token, then the LAST prompt token is fed again at pos=len(prompt)."""
        self.reset()
        for t in prompt:
            self.decode([int(t)])
        out = []
        nxt = int(prompt[-1])
        for _ in range(n_new):
            nxt = int(self.decode([nxt])[0])
            out.append(nxt)
        return out

    # ------------------------------------------------------------------ taps
    def buffer(self, name: str, shape, dtype: torch.dtype) -> torch.Tensor:
        """Copy of an engine-resident buffer (parity taps)."""
        p = lib().wf_engine_buffer(self._h, name.encode())
        if not p:
            raise KeyError(name)
        self.sync()
        n = int(np.prod(shape))
        out = torch.empty(n, dtype=dtype, device="cuda")
        check(lib().wf_copy_d2d(out.data_ptr(), p, n * out.element_size(), None))
        torch.cuda.synchronize()
        return out.reshape(shape)

    def time_kernel(self, which: str, batch: int, iters: int) -> float:
        ms = C.c_float()
        check(lib().wf_engine_time_kernel(self._h, which.encode(), batch, iters, C.byref(ms)))
        return float(ms.value)
