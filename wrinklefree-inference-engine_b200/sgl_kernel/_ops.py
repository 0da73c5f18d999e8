"""torch.library registration of the BitNet ops under the reference's op names.

Reference op schemas (docs/research/sglang_optimization_plan.md:59-75):
    sgl_kernel::bitnet_gemv_cpu(packed_weights, activations, scale) -> f32[N]
    sgl_kernel::bitnet_gemm_cpu(packed_weights, activations, scale) -> f32[M,N]
    sgl_kernel::bitnet_quantize_activations_cpu(activations) -> (int8 like x, f32[1])
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


_IMPL = {"bitnet_gemv": (_gemv, _fake_gemv), "bitnet_gemm": (_gemm, _fake_gemm), "bitnet_quantize_activations": (_quant, _fake_quant)}

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
