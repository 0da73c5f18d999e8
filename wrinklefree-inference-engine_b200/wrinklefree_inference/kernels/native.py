"""Mirror of the reference's src/wrinklefree_inference/kernels/native.py API
(`build_kernel`, `get_kernel`, `pack_weights`, `unpack_weights`, `repack_hf_weights`,
`quantize_activations`) with the scalar OpenMP kernel replaced by the sm_100a library.

The reference module's kernel works on the LSB-SEQUENTIAL layout (byte j bits 2i+1..2i = w[4j+i],
native.py:181-198) and returns scale * sum (code-1)*a (:27-41).  `get_kernel()` here returns an object
with the same `gemv / gemm / num_threads` methods on numpy arrays in that layout: the bytes are
re-blocked to I2_S on the host (a pure permutation of 2-bit fields) and the integer work runs on the GPU.
"""
from __future__ import annotations

import numpy as np


def pack_weights(weights: np.ndarray) -> np.ndarray:
    """float32 ternary [N,K] -> LSB-sequential u8 [N,K/4] (reference :181-198), vectorised."""
    out_f, in_f = weights.shape
    assert in_f % 4 == 0, f"in_features must be divisible by 4, got {in_f}"
    c = (weights.astype(np.int32) + 1).clip(0, 2).astype(np.uint8).reshape(out_f, in_f // 4, 4)
    return (c[:, :, 0] | (c[:, :, 1] << 2) | (c[:, :, 2] << 4) | (c[:, :, 3] << 6)).astype(np.uint8)


def unpack_weights(packed: np.ndarray) -> np.ndarray:
    """LSB-sequential u8 [N,K/4] -> float32 ternary [N,K] (reference :201-216)."""
    out_f = packed.shape[0]
    w = np.stack([((packed >> (2 * j)) & 3) for j in range(4)], axis=-1).reshape(out_f, -1)
    return w.astype(np.float32) - 1.0


def repack_hf_weights(packed_hf: np.ndarray) -> np.ndarray:
    """Reference :218-241 (interleaved-along-N reading of the HF bytes, pinned by its own round trip)."""
    out_packed, in_features = packed_hf.shape
    unpacked = np.zeros((out_packed * 4, in_features), dtype=np.float32)
    p = packed_hf.astype(np.int32)
    for i in range(4):
        unpacked[i::4, :] = ((p >> (i * 2)) & 3) - 1.0
    return pack_weights(unpacked)


def quantize_activations(x: np.ndarray):
    """Reference :243-256 — note: TRUNCATES toward zero (astype) instead of rounding."""
    max_val = np.abs(x).max()
    scale = max_val / 127.0 if max_val > 1e-6 else 1.0
    return np.clip(x / scale, -128, 127).astype(np.int8), scale


def _lsb_to_i2s(packed: np.ndarray) -> np.ndarray:
    n, k4 = packed.shape
    k = k4 * 4
    if k % 128 != 0:
        raise ValueError(f"in_features ({k}) must be multiple of 128")
    codes = np.stack([((packed >> (2 * j)) & 3) for j in range(4)], axis=-1).reshape(n, k // 128, 4, 32)
    return ((codes[:, :, 0] << 6) | (codes[:, :, 1] << 4) | (codes[:, :, 2] << 2) | codes[:, :, 3]).reshape(n, k4).astype(np.uint8)


class _CudaKernel:
    """`bitnet_native` look-alike (reference PYBIND11_MODULE, native.py:104-108)."""

    def gemv(self, packed_weights: np.ndarray, activations: np.ndarray, scale: float) -> np.ndarray:
        return self.gemm(packed_weights, activations[None, :], scale)[0]

    def gemm(self, packed_weights: np.ndarray, activations: np.ndarray, scale: float) -> np.ndarray:
        import torch
        from wf_bitnet import ops
        w = torch.from_numpy(_lsb_to_i2s(np.ascontiguousarray(packed_weights, np.uint8))).cuda()
        a = torch.from_numpy(np.ascontiguousarray(activations, np.int8)).cuda()
        # Sigma-a of the caller's (host) activations: input preparation, done where the data already is
        rowsum = torch.from_numpy(np.ascontiguousarray(activations, np.int8).astype(np.int32).sum(axis=1).astype(np.int32)).cuda()
        acc = ops.ternary_linear(w, w.shape[0], w.shape[1] * 4, a, layout="i2s", mode="i32", rowsum=rowsum)
        return (acc.to(torch.float32) * float(scale)).cpu().numpy()

    def num_threads(self) -> int:
        from wf_bitnet import _lib
        return int(_lib.lib().wf_sm_count())


_kernel = None


def build_kernel(force: bool = False):
    """Load (not JIT-build) the native library; raises when it is missing (no fallback)."""
    global _kernel
    from wf_bitnet import _lib
    _lib.require_gpu()
    if _kernel is None or force:
        _kernel = _CudaKernel()
    return _kernel


def get_kernel():
    return build_kernel() if _kernel is None else _kernel
