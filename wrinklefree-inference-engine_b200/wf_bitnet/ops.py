"""Tensor-level host API over the C ABI (torch is used for device memory and streams only).

Every function takes CUDA tensors, launches on the current torch CUDA stream and returns CUDA
tensors.  Nothing here computes on the CPU and nothing falls back: a missing library or device
raises `WfError`.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import (WF_BF16, WF_EPI_RELU2_MUL, WF_EPI_RESIDUAL, WF_F16, WF_F32, WF_I8, WF_OUT_COMPAT_F32, WF_OUT_I32,
                   WF_OUT_I32_PLUS1, WF_OUT_LINEAR, WF_W_I2S, WF_W_TILE, KvLayout, LinearArgs, check, lib)

_DT = {torch.float32: WF_F32, torch.bfloat16: WF_BF16, torch.float16: WF_F16}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts: torch.Tensor) -> None:
    _lib.require_gpu()
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("expected a CUDA tensor (the B200 BitNet ops have no CPU path)")


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise ValueError(f"unsupported dtype {t.dtype}") from None


_workspaces = {}


def workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Zero-initialised, self-cleaning split-K workspace (kernels leave it zero). One per device+stream."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), _stream())
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


# ----------------------------------------------------------------------------------------- weight formats
def pack_ternary(weight: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """_pack_ternary_weights (reference: cache/bitnet_converter.py:26-77) on the GPU."""
    _need_cuda(weight)
    if weight.dim() != 2:
        raise ValueError("weight must be 2-D [out_features, in_features]")
    w = weight.contiguous()
    N, K = w.shape
    if K % 4 != 0:
        raise ValueError(f"in_features ({K}) must be divisible by 4")
    packed = torch.empty(N, K // 4, dtype=torch.uint8, device=w.device)
    scale = torch.empty(1, dtype=torch.float32, device=w.device)
    check(lib().wf_pack_ternary(w.data_ptr(), _dt(w), N, K, packed.data_ptr(), scale.data_ptr(), _stream()))
    return packed, scale


def unpack_i2s(packed: torch.Tensor) -> torch.Tensor:
    _need_cuda(packed)
    p = packed.contiguous()
    N, K4 = p.shape
    out = torch.empty(N, K4 * 4, dtype=torch.float32, device=p.device)
    check(lib().wf_unpack_i2s(p.data_ptr(), N, K4 * 4, out.data_ptr(), _stream()))
    return out


def hf_to_i2s(hf_packed: torch.Tensor) -> torch.Tensor:
    """HF checkpoint layout u8 [N/4, K] -> kernel layout u8 [N, K/4] (reference: direct_inference_full.py:113-117)."""
    _need_cuda(hf_packed)
    h = hf_packed.contiguous()
    if h.dtype != torch.uint8 or h.dim() != 2:
        raise ValueError("hf_packed must be a 2-D uint8 tensor")
    N4, K = h.shape
    out = torch.empty(N4 * 4, K // 4, dtype=torch.uint8, device=h.device)
    check(lib().wf_hf_to_i2s(h.data_ptr(), N4, K, out.data_ptr(), _stream()))
    return out


def tile_bytes(N: int, K: int) -> int:
    return int(lib().wf_tile_bytes(N, K))


def i2s_to_tile(packed: torch.Tensor) -> torch.Tensor:
    """I2_S [N,K/4] -> b200_tile_v1 (flat uint8)."""
    _need_cuda(packed)
    p = packed.contiguous()
    N, K4 = p.shape
    K = K4 * 4
    if K % 128 != 0:
        raise ValueError(f"in_features ({K}) must be multiple of 128")
    out = torch.empty(tile_bytes(N, K), dtype=torch.uint8, device=p.device)
    check(lib().wf_repack_i2s_to_tile(p.data_ptr(), N, K, out.data_ptr(), _stream()))
    return out


def tile_to_i2s(tile: torch.Tensor, N: int, K: int) -> torch.Tensor:
    _need_cuda(tile)
    out = torch.empty(N, K // 4, dtype=torch.uint8, device=tile.device)
    check(lib().wf_repack_tile_to_i2s(tile.data_ptr(), N, K, out.data_ptr(), _stream()))
    return out


def interleave_rows16(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    _need_cuda(a, b)
    a, b = a.contiguous(), b.contiguous()
    if a.shape != b.shape:
        raise ValueError("gate/up shapes differ")
    N, K4 = a.shape
    out = torch.empty(2 * N, K4, dtype=torch.uint8, device=a.device)
    check(lib().wf_interleave_rows16(a.data_ptr(), b.data_ptr(), N, K4, out.data_ptr(), _stream()))
    return out


# ----------------------------------------------------------------------------------------- activation quant
def quantize_activations(x: torch.Tensor, per_token: bool = False):
    """int8 absmax quantisation. Returns (q int8 like x, scale f32[1] or f32[M], rowsum i32[M])."""
    _need_cuda(x)
    shape = x.shape
    x2 = x.reshape(-1, shape[-1]) if x.dim() != 1 else x.reshape(1, -1)
    x2 = x2.contiguous()
    M, K = x2.shape
    q = torch.empty(M, K, dtype=torch.int8, device=x.device)
    scale = torch.empty(M if per_token else 1, dtype=torch.float32, device=x.device)
    rowsum = torch.empty(max(M, 1), dtype=torch.int32, device=x.device)
    cell = torch.empty(1, dtype=torch.int32, device=x.device)   # per-tensor absmax cell (the C side zeroes it)
    check(lib().wf_act_quant_i8(x2.data_ptr(), _dt(x2), M, K, K, 1 if per_token else 0, q.data_ptr(), scale.data_ptr(),
                                rowsum.data_ptr(), cell.data_ptr(), _stream()))
    return q.reshape(shape), scale, rowsum


def rmsnorm_quant(x: torch.Tensor, gamma: torch.Tensor, eps: float, want_y: bool = False):
    _need_cuda(x, gamma)
    if x.dtype != torch.bfloat16 or gamma.dtype != torch.bfloat16:
        raise ValueError("rmsnorm_quant expects bf16 activations and gamma")
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    M, K = x2.shape
    q = torch.empty(M, K, dtype=torch.int8, device=x.device)
    scale = torch.empty(M, dtype=torch.float32, device=x.device)
    rowsum = torch.empty(M, dtype=torch.int32, device=x.device)
    y = torch.empty_like(x2) if want_y else None
    check(lib().wf_rmsnorm_quant(x2.data_ptr(), gamma.contiguous().data_ptr(), float(eps), M, K,
                                 y.data_ptr() if want_y else None, q.data_ptr(), scale.data_ptr(), rowsum.data_ptr(), _stream()))
    return q, scale, rowsum, y


# ----------------------------------------------------------------------------------------- ternary linear
def ternary_linear(w: torch.Tensor, N: int, K: int, q: torch.Tensor, *, layout: str = "tile", mode: str = "i32_plus1",
                   rowsum: Optional[torch.Tensor] = None, act_scale: Optional[torch.Tensor] = None,
                   w_scale: Optional[torch.Tensor] = None, seg_end: Optional[Sequence[int]] = None, host_scale: float = 1.0,
                   residual: Optional[torch.Tensor] = None, relu2_mul: bool = False, out_dtype: torch.dtype = torch.float32,
                   algo: int = 0) -> torch.Tensor:
    """General entry point (wf_ternary_linear). q: int8 [M,K]."""
    _need_cuda(w, q)
    if q.dtype != torch.int8:
        raise ValueError("activations must be int8")
    q2 = q.reshape(1, -1) if q.dim() == 1 else q
    q2 = q2.contiguous()
    M = q2.shape[0]
    if q2.shape[1] != K:
        raise ValueError("activations.shape[-1] != in_features")
    a = LinearArgs()
    a.w = w.data_ptr()
    a.w_layout = WF_W_TILE if layout == "tile" else WF_W_I2S
    a.N, a.K = N, K
    a.q, a.M, a.ldq = q2.data_ptr(), M, K
    modes = {"i32_plus1": WF_OUT_I32_PLUS1, "i32": WF_OUT_I32, "compat": WF_OUT_COMPAT_F32, "linear": WF_OUT_LINEAR}
    a.out_mode = modes[mode]
    out_cols = N // 2 if relu2_mul else N
    if mode in ("i32_plus1", "i32"):
        out = torch.empty(M, N, dtype=torch.int32, device=q.device)
    elif mode == "compat":
        out = torch.empty(M, N, dtype=torch.float32, device=q.device)
    else:
        out = torch.empty(M, out_cols, dtype=out_dtype, device=q.device)
        a.out_dtype = _DT[out_dtype]
    if rowsum is not None:
        a.rowsum = rowsum.data_ptr()
    if act_scale is not None:
        a.act_scale = act_scale.data_ptr()
        a.act_scale_per_token = 1 if act_scale.numel() > 1 or M == 1 else 0
    if w_scale is not None:
        a.w_scale = w_scale.data_ptr()
        a.nseg = w_scale.numel()
        se = list(seg_end) if seg_end is not None else [N]
        for i, v in enumerate(se[:4]):
            a.seg_end[i] = int(v)
    a.host_scale = float(host_scale)
    a.epi_flags = (WF_EPI_RESIDUAL if residual is not None else 0) | (WF_EPI_RELU2_MUL if relu2_mul else 0)
    if residual is not None:
        residual = residual.contiguous()
        a.residual = residual.data_ptr()
    a.out, a.ldo = out.data_ptr(), out.shape[1]
    nbytes = int(lib().wf_linear_workspace_bytes(M, N, K)) + 4 * ((N + 15) // 16) + 256
    ws = workspace(q.device, nbytes)
    a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
    a.algo = algo
    check(lib().wf_ternary_linear(C.byref(a), _stream()))
    return out


def bitnet_gemv(packed: torch.Tensor, x: torch.Tensor, scale: float) -> torch.Tensor:
    """torch.ops.sgl_kernel.bitnet_gemv_cpu semantics: f32[N] = scale * sum (w+1)*x, packed = I2_S [N,K/4]."""
    _need_cuda(packed, x)
    p = packed.contiguous()
    xx = x.contiguous()
    N, K4 = p.shape
    y = torch.empty(N, dtype=torch.float32, device=p.device)
    ws = workspace(p.device, int(lib().wf_linear_workspace_bytes(1, N, K4 * 4)))
    check(lib().wf_bitnet_gemv(p.data_ptr(), xx.data_ptr(), float(scale), N, K4 * 4, y.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return y


def bitnet_gemm(packed: torch.Tensor, x: torch.Tensor, scale: float) -> torch.Tensor:
    _need_cuda(packed, x)
    p = packed.contiguous()
    xx = x.contiguous()
    N, K4 = p.shape
    M = xx.shape[0]
    y = torch.empty(M, N, dtype=torch.float32, device=p.device)
    ws = workspace(p.device, int(lib().wf_linear_workspace_bytes(M, N, K4 * 4)))
    check(lib().wf_bitnet_gemm(p.data_ptr(), xx.data_ptr(), float(scale), M, N, K4 * 4, y.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return y


# ----------------------------------------------------------------------------------------- KV + attention
class PagedKV:
    """Device pool [page][layer][K|V][kv_head][slot][D] + the KvLayout struct the C ABI wants."""

    def __init__(self, max_pages: int, layers: int, kv_heads: int, page_size: int, head_dim: int,
                 dtype: torch.dtype = torch.bfloat16, device="cuda"):
        _lib.require_gpu()
        self.max_pages, self.layers, self.kv_heads, self.page_size, self.head_dim = max_pages, layers, kv_heads, page_size, head_dim
        self.dtype = dtype
        if dtype == torch.int8:
            # codes [page][layer][K|V][head][slot][D], stored offset-binary (uint8 = code + 128), followed by the fp32 scales
            # [page][layer][K|V][head][slot] in ONE allocation
            elems = max_pages * layers * 2 * kv_heads * page_size * head_dim
            nbytes = int(lib().wf_kv_pool_bytes(max_pages, layers, kv_heads, page_size, head_dim, WF_I8))
            self.raw = torch.zeros(nbytes, dtype=torch.uint8, device=device)
            self.pool = self.raw[:elems].view(max_pages, layers, 2, kv_heads, page_size, head_dim)
            self.scales = self.raw[elems:].view(torch.float32).view(max_pages, layers, 2, kv_heads, page_size)
            self.c = KvLayout(self.raw.data_ptr(), max_pages, layers, kv_heads, page_size, head_dim, WF_I8)
            return
        self.pool = torch.zeros(max_pages, layers, 2, kv_heads, page_size, head_dim, dtype=dtype, device=device)
        self.c = KvLayout(self.pool.data_ptr(), max_pages, layers, kv_heads, page_size, head_dim, _DT[dtype])

    def ref(self):
        return C.byref(self.c)

    def codes(self) -> torch.Tensor:
        """int8 pools: the signed codes (pool bytes are offset-binary)."""
        return (self.pool.to(torch.int16) - 128).to(torch.int8)

    def set_codes(self, codes: torch.Tensor) -> None:
        self.pool.copy_((codes.to(torch.int16) + 128).to(torch.uint8))


def kv_scatter(kv: PagedKV, k: torch.Tensor, v: torch.Tensor, pages: torch.Tensor, slots: torch.Tensor, layer: int) -> None:
    _need_cuda(k, v, pages, slots)
    T = pages.shape[0]
    check(lib().wf_kv_scatter(kv.ref(), k.contiguous().data_ptr(), v.contiguous().data_ptr(), pages.data_ptr(), slots.data_ptr(), T, layer, _stream()))


def kv_gather(kv: PagedKV, k_out: torch.Tensor, v_out: torch.Tensor, pages: torch.Tensor, slots: torch.Tensor, layer: int) -> None:
    _need_cuda(k_out, v_out, pages, slots)
    if not (k_out.is_contiguous() and v_out.is_contiguous()):
        raise ValueError("gather outputs must be contiguous")
    T = pages.shape[0]
    check(lib().wf_kv_gather(kv.ref(), k_out.data_ptr(), v_out.data_ptr(), pages.data_ptr(), slots.data_ptr(), T, layer, _stream()))


def rope_kv_append(qkv: torch.Tensor, H: int, kv: PagedKV, layer: int, positions: torch.Tensor, block_table: torch.Tensor,
                   cos_tab: torch.Tensor, sin_tab: torch.Tensor) -> torch.Tensor:
    _need_cuda(qkv, positions, block_table, cos_tab, sin_tab)
    B = qkv.shape[0]
    q_out = torch.empty(B, H, kv.head_dim, dtype=torch.bfloat16, device=qkv.device)
    check(lib().wf_rope_kv_append(qkv.contiguous().data_ptr(), B, H, kv.ref(), layer, positions.data_ptr(), block_table.data_ptr(),
                                  block_table.shape[1], cos_tab.data_ptr(), sin_tab.data_ptr(), cos_tab.shape[0], q_out.data_ptr(), _stream()))
    return q_out


_attn_ws = {}


def paged_decode_attn(q: torch.Tensor, kv: PagedKV, layer: int, block_table: torch.Tensor, seq_lens: torch.Tensor,
                      scale: Optional[float] = None, out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    _need_cuda(q, block_table, seq_lens)
    B, H, D = q.shape
    if scale is None:
        scale = 1.0 / (D ** 0.5)
    out = torch.empty(B, H, D, dtype=out_dtype, device=q.device)
    max_seq = block_table.shape[1] * kv.page_size
    nbytes = int(lib().wf_attn_workspace_bytes(B, H, D, max_seq))
    key = (q.device.index, _stream())
    ws = _attn_ws.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(nbytes, dtype=torch.uint8, device=q.device)
        _attn_ws[key] = ws
    check(lib().wf_paged_decode_attn(q.contiguous().data_ptr(), kv.ref(), layer, block_table.data_ptr(), block_table.shape[1],
                                     seq_lens.data_ptr(), B, H, float(scale), out.data_ptr(), _DT[out_dtype], ws.data_ptr(), ws.numel(), _stream()))
    return out


# ----------------------------------------------------------------------------------------- glue
def embed(table: torch.Tensor, tokens: torch.Tensor) -> torch.Tensor:
    _need_cuda(table, tokens)
    B = tokens.shape[0]
    out = torch.empty(B, table.shape[1], dtype=torch.bfloat16, device=table.device)
    check(lib().wf_embed(table.data_ptr(), tokens.data_ptr(), B, table.shape[1], out.data_ptr(), _stream()))
    return out


def lm_head_argmax(h: torch.Tensor, E: torch.Tensor, want_logits: bool = True):
    _need_cuda(h, E)
    B, K = h.shape
    V = E.shape[0]
    logits = torch.empty(B, V, dtype=torch.float32, device=h.device) if want_logits else None
    nxt = torch.empty(B, dtype=torch.int32, device=h.device)
    ws = torch.empty(int(lib().wf_lm_head_workspace_bytes(B, V)), dtype=torch.uint8, device=h.device)
    check(lib().wf_lm_head_argmax(h.contiguous().data_ptr(), E.data_ptr(), B, V, K, logits.data_ptr() if want_logits else None,
                                  nxt.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return nxt, logits
