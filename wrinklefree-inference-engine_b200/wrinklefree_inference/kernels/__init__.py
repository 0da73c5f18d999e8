# BitNet native kernels (reference: src/wrinklefree_inference/kernels/__init__.py:3-15)
from .bitnet_patch import (
    NATIVE_KERNELS_AVAILABLE,
    BitNetLinearNative,
    patch_model_with_native_kernels,
    benchmark_kernels,
)

__all__ = [
    "NATIVE_KERNELS_AVAILABLE",
    "BitNetLinearNative",
    "patch_model_with_native_kernels",
    "benchmark_kernels",
]
