"""sgl_kernel.quantization.bitnet — same names, argument meaning and error behaviour as the missing
reference module (its Python is preserved verbatim in the reference's tests/test_bitnet_core.py:139-167
and legacy/benchmark/test_sglang_bitnet_kernel.py:44-101), executed by the sm_100a kernels.

Semantics kept from the SIMD kernel the server uses:
  * packed weights are the BLOCKED I2_S layout u8 [N, K/4];
  * bitnet_gemv / bitnet_gemm return  scale * sum_k (w+1) * a  in fp32 — the caller subtracts
    scale * sum(a) (scripts/direct_inference_full.py:204-206);
  * bitnet_quantize_activations uses ONE absmax for the whole tensor and returns the scale as a Python
    float (bitnet_patch.py:62-65) — this forces a device->host sync; graph-capturable code should use
    wf_bitnet.ops.quantize_activations / the engine instead.
"""
from typing import Tuple

import torch

import wf_bitnet
from wf_bitnet import ops as _ops

BITNET_BLOCK_SIZE = wf_bitnet.BITNET_BLOCK_SIZE  # 128 == QK_I2_S (extern/BitNet/src/ggml-bitnet-mad.cpp:9)
QK_I2_S = BITNET_BLOCK_SIZE


def bitnet_check_kernel_available() -> bool:
    """True when the CUDA library is loadable and an sm_100 device is present."""
    return wf_bitnet.available()


check_kernel_available = bitnet_check_kernel_available


def auto_tune_tiles(*args, **kwargs) -> None:
    """The CPU kernel tunes cache tiles per shape (docs/research/notebook.md:92); nothing to tune here —
    tile shapes are fixed by the tensor-core fragment layout."""
    return None


def bitnet_quantize_activations(activations: torch.Tensor) -> Tuple[torch.Tensor, float]:
    """Quantize FP32/BF16 activations to INT8 (per-tensor absmax). Returns (int8 like x, float scale)."""
    q, scale = torch.ops.sgl_kernel.bitnet_quantize_activations_cpu(activations)
    return q, float(scale.item())


quantize_activations_i8 = bitnet_quantize_activations


def _check(packed_weights: torch.Tensor, in_features_of_x: int) -> int:
    if packed_weights.dim() != 2 or packed_weights.dtype != torch.uint8:
        raise ValueError("packed_weights must be a 2-D uint8 tensor [out_features, in_features/4]")
    in_features = packed_weights.shape[1] * 4
    if in_features % QK_I2_S != 0:
        raise ValueError(f"in_features ({in_features}) must be multiple of {QK_I2_S}")
    if in_features_of_x != in_features:
        raise ValueError("activations.shape[0] != in_features" if in_features_of_x >= 0 else "bad activations")
    return in_features


def bitnet_gemv(packed_weights: torch.Tensor, activations: torch.Tensor, scale: float) -> torch.Tensor:
    """y[N] = scale * sum_k (w[n,k]+1) * x[k]   (fp32)."""
    if activations.dim() != 1:
        raise ValueError("bitnet_gemv expects 1-D int8 activations [in_features]")
    if activations.dtype != torch.int8:
        raise ValueError("activations must be int8")
    _check(packed_weights, activations.shape[0])
    return torch.ops.sgl_kernel.bitnet_gemv_cpu(packed_weights, activations, float(scale))


def bitnet_gemm(packed_weights: torch.Tensor, activations: torch.Tensor, scale: float) -> torch.Tensor:
    """Y[M,N] = scale * X[M,K] @ (W+1)^T   (fp32)."""
    if activations.dim() != 2:
        raise ValueError("bitnet_gemm expects 2-D int8 activations [batch, in_features]")
    if activations.dtype != torch.int8:
        raise ValueError("activations must be int8")
    _check(packed_weights, activations.shape[1])
    return torch.ops.sgl_kernel.bitnet_gemm_cpu(packed_weights, activations, float(scale))
