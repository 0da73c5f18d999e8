"""torch.library registration of the BitNet ops under the reference's op names.

Reference op schemas (docs/research/sglang_optimization_plan.md:59-75):
    sgl_kernel::bitnet_gemv_cpu(packed_weights, activations, scale) -> f32[N]
    sgl_kernel::bitnet_gemm_cpu(packed_weights, activations, scale) -> f32[M,N]
    sgl_kernel::bitnet_quantize_activations_cpu(activations) -> (int8 like x, f32[1])
Optional fused ops the reference experimented with (sglang_optimization_plan.md:326-331; the fork's source is absent, so the
argument lists below are this library's):
    sgl_kernel::bitnet_qkv_forward_cpu(x, q_w, k_w, v_w, q_scale, k_scale, v_scale) -> (q, k, v)
    sgl_kernel::bitnet_mlp_forward_cpu(x, gate_w, up_w, down_w, gate_scale, up_scale, down_scale, ffn_sub_norm, eps) -> y
Both follow scripts/direct_inference_full.py (:186-208 per linear, :253-264 the MLP): per-token int8 absmax quantisation, exact
int32 contraction, Sigma-a correction + dequant to bf16, relu(gate)^2 * up, ffn_sub_norm, re-quantisation, down projection.
The names keep their `_cpu` suffix so existing `torch.ops.sgl_kernel.*` call sites and existence
checks resolve unchanged; the implementations are registered for the CUDA dispatch key only —
calling them with CPU tensors raises (no CPU fallback).  `*_cuda`-free aliases are added too.
"""
import torch

from wf_bitnet import ops as _wf

_LIB = torch.library.Library("sgl_kernel", "FRAGMENT")
_DEFS = {
    "bitnet_gemv_cpu": "(Tensor packed_weights, Tensor activations, float scale) -> Tensor",
    "bitnet_gemm_cpu": "(Tensor packed_weights, Tensor activations, float scale) -> Tensor",
    "bitnet_quantize_activations_cpu": "(Tensor activations) -> (Tensor, Tensor)",
    "bitnet_gemv": "(Tensor packed_weights, Tensor activations, float scale) -> Tensor",
    "bitnet_gemm": "(Tensor packed_weights, Tensor activations, float scale) -> Tensor",
    "bitnet_quantize_activations": "(Tensor activations) -> (Tensor, Tensor)",
    "bitnet_qkv_forward_cpu": "(Tensor x, Tensor q_w, Tensor k_w, Tensor v_w, float q_scale, float k_scale, float v_scale) -> (Tensor, Tensor, Tensor)",
    "bitnet_mlp_forward_cpu": "(Tensor x, Tensor gate_w, Tensor up_w, Tensor down_w, float gate_scale, float up_scale, float down_scale, Tensor ffn_sub_norm, float eps) -> Tensor",
    "bitnet_qkv_forward": "(Tensor x, Tensor q_w, Tensor k_w, Tensor v_w, float q_scale, float k_scale, float v_scale) -> (Tensor, Tensor, Tensor)",
    "bitnet_mlp_forward": "(Tensor x, Tensor gate_w, Tensor up_w, Tensor down_w, float gate_scale, float up_scale, float down_scale, Tensor ffn_sub_norm, float eps) -> Tensor",
}


def _gemv(packed_weights, activations, scale):
    return _wf.bitnet_gemv(packed_weights, activations, scale)


def _gemm(packed_weights, activations, scale):
    return _wf.bitnet_gemm(packed_weights, activations, scale)


def _quant(activations):
    q, scale, _ = _wf.quantize_activations(activations, per_token=False)
    return q, scale


def _fake_gemv(packed_weights, activations, scale):
    return torch.empty(packed_weights.shape[0], dtype=torch.float32, device=packed_weights.device)


def _fake_gemm(packed_weights, activations, scale):
    return torch.empty(activations.shape[0], packed_weights.shape[0], dtype=torch.float32, device=packed_weights.device)


def _fake_quant(activations):
    return torch.empty_like(activations, dtype=torch.int8), torch.empty(1, dtype=torch.float32, device=activations.device)


# ---- fused forms: one fused q|k|v matrix / one interleaved gate|up matrix in the resident tile layout, built once per weight set
_fused_cache = {}


def _fused_tile(kind, ws):
    key = (kind,) + tuple((w.data_ptr(), tuple(w.shape), w._version) for w in ws)
    hit = _fused_cache.get(key)
    if hit is not None and not all(a is b for a, b in zip(hit[3], ws)):
        hit = None          # same address, other tensor object: never trust a data_ptr whose owner may have been freed
    if hit is None:
        if len(_fused_cache) > 64:
            _fused_cache.clear()
        for w in ws:
            if w.dim() != 2 or w.dtype != torch.uint8 or (w.shape[1] * 4) % 128:
                raise ValueError("packed weights must be uint8 [out_features, in_features/4] with in_features % 128 == 0")
        if kind == "qkv":
            cat = torch.cat([w.contiguous() for w in ws], dim=0)
        else:
            cat = _wf.interleave_rows16(ws[0], ws[1])       # gate|up in 16-row blocks: relu^2*up pairs sit in one CTA
        hit = (_wf.i2s_to_tile(cat), cat.shape[0], cat.shape[1] * 4, tuple(ws))   # the entry keeps its source tensors alive
        _fused_cache[key] = hit
    return hit[:3]


def _rows(x):
    if not x.is_floating_point():
        raise ValueError("x must be a floating-point activation tensor")
    return x.reshape(-1, x.shape[-1]).contiguous()


def _qkv_forward(x, q_w, k_w, v_w, q_scale, k_scale, v_scale):
    x2 = _rows(x)
    tile, N, K = _fused_tile("qkv", (q_w, k_w, v_w))
    if x2.shape[1] != K:
        raise ValueError("x.shape[-1] != in_features")
    q8, s, rs = _wf.quantize_activations(x2, per_token=True)
    nq, nk = q_w.shape[0], k_w.shape[0]
    ws = torch.tensor([q_scale, k_scale, v_scale], dtype=torch.float32, device=x.device)
    out = _wf.ternary_linear(tile, N, K, q8, mode="linear", rowsum=rs, act_scale=s, w_scale=ws, seg_end=[nq, nq + nk, N], out_dtype=torch.bfloat16)
    lead = x.shape[:-1]
    return (out[:, :nq].reshape(*lead, nq).to(x.dtype), out[:, nq:nq + nk].reshape(*lead, nk).to(x.dtype), out[:, nq + nk:].reshape(*lead, N - nq - nk).to(x.dtype))


def _mlp_forward(x, gate_w, up_w, down_w, gate_scale, up_scale, down_scale, ffn_sub_norm, eps):
    x2 = _rows(x)
    tile, N2, K = _fused_tile("gu", (gate_w, up_w))
    if x2.shape[1] != K:
        raise ValueError("x.shape[-1] != in_features")
    q8, s, rs = _wf.quantize_activations(x2, per_token=True)
    ws = torch.tensor([gate_scale, up_scale], dtype=torch.float32, device=x.device)
    h = _wf.ternary_linear(tile, N2, K, q8, mode="linear", rowsum=rs, act_scale=s, w_scale=ws, relu2_mul=True, out_dtype=torch.bfloat16)
    q2, s2, rs2, _ = _wf.rmsnorm_quant(h, ffn_sub_norm.to(torch.bfloat16), eps)
    wd = torch.tensor([down_scale], dtype=torch.float32, device=x.device)
    y = _wf.ternary_linear(down_w.contiguous(), down_w.shape[0], down_w.shape[1] * 4, q2, layout="i2s", mode="linear", rowsum=rs2, act_scale=s2,
                           w_scale=wd, out_dtype=torch.bfloat16)
    return y.reshape(*x.shape[:-1], down_w.shape[0]).to(x.dtype)


def _fake_qkv(x, q_w, k_w, v_w, q_scale, k_scale, v_scale):
    lead = x.shape[:-1]
    return tuple(torch.empty(*lead, w.shape[0], dtype=x.dtype, device=x.device) for w in (q_w, k_w, v_w))


def _fake_mlp(x, gate_w, up_w, down_w, gate_scale, up_scale, down_scale, ffn_sub_norm, eps):
    return torch.empty(*x.shape[:-1], down_w.shape[0], dtype=x.dtype, device=x.device)


_IMPL = {"bitnet_gemv": (_gemv, _fake_gemv), "bitnet_gemm": (_gemm, _fake_gemm), "bitnet_quantize_activations": (_quant, _fake_quant),
         "bitnet_qkv_forward": (_qkv_forward, _fake_qkv), "bitnet_mlp_forward": (_mlp_forward, _fake_mlp)}

_registered = False


def register():
    global _registered
    if _registered:
        return
    for name, schema in _DEFS.items():
        _LIB.define(name + schema)
        impl, fake = _IMPL[name.replace("_cpu", "")]
        _LIB.impl(name, impl, "CUDA")
        torch.library.register_fake(f"sgl_kernel::{name}", fake, lib=_LIB)
    _registered = True


register()
