"""sgl_kernel.quantization — BitNet names (reference import list: kernels/bitnet_patch.py:14-21,
scripts/benchmark_kernels.py:11-17; legacy alias scripts/test_fallback.py:20)."""
from .bitnet import (
    BITNET_BLOCK_SIZE,
    QK_I2_S,
    auto_tune_tiles,
    bitnet_check_kernel_available,
    bitnet_gemm,
    bitnet_gemv,
    bitnet_quantize_activations,
    check_kernel_available,
    quantize_activations_i8,
)

__all__ = [
    "BITNET_BLOCK_SIZE", "QK_I2_S", "auto_tune_tiles", "bitnet_check_kernel_available", "bitnet_gemm", "bitnet_gemv",
    "bitnet_quantize_activations", "check_kernel_available", "quantize_activations_i8",
]
