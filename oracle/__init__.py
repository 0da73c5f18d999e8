"""TEST INFRASTRUCTURE ONLY — CPU oracle for the BitNet b1.58 decode hot path.

numpy/ctypes front-end of ``oracle/wf_oracle.c`` (our C restatement of the reference
algorithm) and of ``oracle/_ref/libwf_ref.so`` (the reference's own, unmodified SIMD
kernel ``extern/BitNet/src/ggml-bitnet-mad.cpp`` compiled by ``oracle/Makefile``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker / the timed
CPU baseline.  The product (``wrinklefree-inference-engine_b200/``) never does; it fails
loudly when its CUDA library is missing instead of falling back to anything here.

Parity status: pinned (see ``tests/test_oracle_golden.py``).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_ORACLE_SO = _HERE / "libwf_oracle.so"
_REF_SO = _HERE / "_ref" / "libwf_ref.so"

_u8p = ctypes.POINTER(ctypes.c_uint8)
_i8p = ctypes.POINTER(ctypes.c_int8)
_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int32)
_f64p = ctypes.POINTER(ctypes.c_double)


def build(force: bool = False) -> None:
    """Compile the C restatement (and oracle/_ref when /root/reference is present)."""
    if force or not _ORACLE_SO.exists() or (_ORACLE_SO.stat().st_mtime < (_HERE / "wf_oracle.c").stat().st_mtime):
        subprocess.run(["make", "-C", str(_HERE), "all"], check=True, capture_output=True)
    elif not _REF_SO.exists() and Path("/root/reference/extern/BitNet/src/ggml-bitnet-mad.cpp").exists():
        subprocess.run(["make", "-C", str(_HERE), "ref"], check=True, capture_output=True)


_lib = None
_ref = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(str(_ORACLE_SO))
        L.wfo_pack_ternary.argtypes = [_f32p, ctypes.c_int, ctypes.c_int, _u8p, _f32p]
        L.wfo_pack_ternary.restype = ctypes.c_int
        L.wfo_quantize_i2_s.argtypes = [_f32p, ctypes.c_int, ctypes.c_int, _u8p]
        L.wfo_quantize_i2_s.restype = ctypes.c_float
        L.wfo_unpack_i2s.argtypes = [_u8p, ctypes.c_int, ctypes.c_int, _f32p]
        L.wfo_unpack_hf.argtypes = [_u8p, ctypes.c_int, ctypes.c_int, _f32p]
        L.wfo_hf_to_i2s.argtypes = [_u8p, ctypes.c_int, ctypes.c_int, _u8p]
        L.wfo_native_pack.argtypes = [_f32p, ctypes.c_int, ctypes.c_int, _u8p]
        L.wfo_quantize_act_i8.argtypes = [_f32p, ctypes.c_size_t, _i8p, _f64p]
        L.wfo_quantize_act_i8.restype = ctypes.c_float
        L.wfo_quantize_act_i8_rows.argtypes = [_f32p, ctypes.c_int, ctypes.c_int, _i8p, _f32p, _i32p]
        L.wfo_vec_dot_i2_i8.argtypes = [ctypes.c_int, _u8p, _i8p]
        L.wfo_vec_dot_i2_i8.restype = ctypes.c_int64
        L.wfo_vec_dot_i2_i8_avx2model.argtypes = [ctypes.c_int, _u8p, _i8p]
        L.wfo_vec_dot_i2_i8_avx2model.restype = ctypes.c_int32
        L.wfo_gemv.argtypes = [_u8p, _i8p, ctypes.c_float, ctypes.c_int, ctypes.c_int, _f32p]
        L.wfo_gemm.argtypes = [_u8p, _i8p, ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        L.wfo_gemm_i32.argtypes = [_u8p, _i8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _i32p]
        L.wfo_native_dot.argtypes = [ctypes.c_int, _u8p, _i8p]
        L.wfo_native_dot.restype = ctypes.c_int32
        L.wfo_bitnet_linear.argtypes = [_f32p, _u8p, ctypes.c_float, ctypes.c_int, ctypes.c_int, _f32p]
        _lib = L
    return _lib


def ref_available() -> bool:
    if not _REF_SO.exists():
        try:
            build()
        except Exception:
            return False
    return _REF_SO.exists()


def ref() -> ctypes.CDLL:
    """The unmodified reference kernel (oracle/_ref). Raises if it was never built."""
    global _ref
    if _ref is None:
        if not ref_available():
            raise RuntimeError("oracle/_ref/libwf_ref.so missing (run `make -C oracle` where /root/reference exists)")
        R = ctypes.CDLL(str(_REF_SO))
        R.wfref_num_threads.restype = ctypes.c_int
        R.wfref_set_threads.argtypes = [ctypes.c_int]
        R.wfref_gemv.argtypes = [_u8p, _i8p, ctypes.c_float, ctypes.c_int, ctypes.c_int, _f32p]
        R.wfref_gemm.argtypes = [_u8p, _i8p, ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        R.wfref_gemm_raw.argtypes = [_u8p, _i8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        R.wfref_quantize_i2_s.argtypes = [_f32p, _u8p, ctypes.c_int, ctypes.c_int]
        R.wfref_quantize_i2_s.restype = ctypes.c_float
        _ref = R
    return _ref


def _c(a: np.ndarray, dtype) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=dtype)


def _p(a: np.ndarray, typ):
    return a.ctypes.data_as(typ)


# ----------------------------------------------------------------------------- packing
def pack_ternary(w: np.ndarray):
    """_pack_ternary_weights (cache/bitnet_converter.py:26-77). Returns (u8 [N,K/4], scale)."""
    w = _c(w, np.float32)
    N, K = w.shape
    if K % 4 != 0:
        raise ValueError(f"in_features ({K}) must be divisible by 4")
    packed = np.empty((N, K // 4), np.uint8)
    s = ctypes.c_float()
    lib().wfo_pack_ternary(_p(w, _f32p), N, K, _p(packed, _u8p), ctypes.byref(s))
    return packed, np.float32(s.value)


def quantize_i2_s(w: np.ndarray):
    """quantize_i2_s restated (ggml-bitnet-mad.cpp:46-92)."""
    w = _c(w, np.float32)
    N, K = w.shape
    packed = np.empty((N, K // 4), np.uint8)
    s = lib().wfo_quantize_i2_s(_p(w, _f32p), N, K, _p(packed, _u8p))
    return packed, np.float32(s)


def unpack_i2s(packed: np.ndarray) -> np.ndarray:
    packed = _c(packed, np.uint8)
    N, K4 = packed.shape
    w = np.empty((N, K4 * 4), np.float32)
    lib().wfo_unpack_i2s(_p(packed, _u8p), N, K4 * 4, _p(w, _f32p))
    return w


def unpack_hf(hf: np.ndarray) -> np.ndarray:
    hf = _c(hf, np.uint8)
    N4, K = hf.shape
    w = np.empty((N4 * 4, K), np.float32)
    lib().wfo_unpack_hf(_p(hf, _u8p), N4, K, _p(w, _f32p))
    return w


def hf_to_i2s(hf: np.ndarray) -> np.ndarray:
    hf = _c(hf, np.uint8)
    N4, K = hf.shape
    out = np.empty((N4 * 4, K // 4), np.uint8)
    lib().wfo_hf_to_i2s(_p(hf, _u8p), N4, K, _p(out, _u8p))
    return out


def native_pack(w: np.ndarray) -> np.ndarray:
    w = _c(w, np.float32)
    N, K = w.shape
    out = np.empty((N, K // 4), np.uint8)
    lib().wfo_native_pack(_p(w, _f32p), N, K, _p(out, _u8p))
    return out


# ----------------------------------------------------------------------------- act quant
def quantize_act_i8(x: np.ndarray):
    """quantize_activations_i8 (tests/test_bitnet_core.py:139-146): per-TENSOR absmax.
    Returns (int8 like x, python-float scale as the reference returns it)."""
    x = _c(x, np.float32)
    q = np.empty(x.shape, np.int8)
    sd = ctypes.c_double()
    lib().wfo_quantize_act_i8(_p(x, _f32p), x.size, _p(q, _i8p), ctypes.byref(sd))
    return q, float(sd.value)


def quantize_act_i8_rows(x: np.ndarray):
    """Per-token: the reference function applied row by row. Returns (q, scales f32[M], rowsum i32[M])."""
    x = _c(x, np.float32)
    M, K = x.shape
    q = np.empty((M, K), np.int8)
    s = np.empty(M, np.float32)
    rs = np.empty(M, np.int32)
    lib().wfo_quantize_act_i8_rows(_p(x, _f32p), M, K, _p(q, _i8p), _p(s, _f32p), _p(rs, _i32p))
    return q, s, rs


# ----------------------------------------------------------------------------- gemv / gemm
def gemv(packed: np.ndarray, act: np.ndarray, scale: float) -> np.ndarray:
    packed = _c(packed, np.uint8)
    act = _c(act, np.int8)
    N, K4 = packed.shape
    K = K4 * 4
    if K % 128 != 0:
        raise ValueError(f"in_features ({K}) must be multiple of 128")
    if act.shape != (K,):
        raise ValueError("activations.shape[0] != in_features")
    out = np.empty(N, np.float32)
    lib().wfo_gemv(_p(packed, _u8p), _p(act, _i8p), scale, N, K, _p(out, _f32p))
    return out


def gemm(packed: np.ndarray, act: np.ndarray, scale: float) -> np.ndarray:
    packed = _c(packed, np.uint8)
    act = _c(act, np.int8)
    N, K4 = packed.shape
    K = K4 * 4
    if K % 128 != 0:
        raise ValueError(f"in_features ({K}) must be multiple of 128")
    M = act.shape[0]
    assert act.shape == (M, K)
    out = np.empty((M, N), np.float32)
    lib().wfo_gemm(_p(packed, _u8p), _p(act, _i8p), scale, M, N, K, _p(out, _f32p))
    return out


def gemm_i32(packed: np.ndarray, act: np.ndarray, mode: int = 0) -> np.ndarray:
    """Raw accumulators. mode 0: sum (w+1)a (kernel), mode 1: sum w*a."""
    packed = _c(packed, np.uint8)
    act = _c(act, np.int8)
    if act.ndim == 1:
        act = act[None]
    N, K4 = packed.shape
    K = K4 * 4
    M = act.shape[0]
    out = np.empty((M, N), np.int32)
    lib().wfo_gemm_i32(_p(packed, _u8p), _p(act, _i8p), M, N, K, mode, _p(out, _i32p))
    return out


def gemm_i32_fast(packed: np.ndarray, act: np.ndarray, mode: int = 0) -> np.ndarray:
    """Same integers as gemm_i32, for BASELINE-size matrices: unpack with the C oracle, then an exact
    float64 BLAS product (|sum| <= 2*128*28672 < 2^53, so every partial sum is an exact integer)."""
    act = _c(act, np.int8)
    if act.ndim == 1:
        act = act[None]
    w = unpack_i2s(packed).astype(np.float64)
    if mode == 0:
        w += 1.0
    out = act.astype(np.float64) @ w.T
    return np.rint(out).astype(np.int64).astype(np.int32)


def vec_dot(row: np.ndarray, act: np.ndarray) -> int:
    row = _c(row, np.uint8)
    act = _c(act, np.int8)
    return int(lib().wfo_vec_dot_i2_i8(act.size, _p(row, _u8p), _p(act, _i8p)))


def vec_dot_avx2model(row: np.ndarray, act: np.ndarray) -> int:
    row = _c(row, np.uint8)
    act = _c(act, np.int8)
    return int(lib().wfo_vec_dot_i2_i8_avx2model(act.size, _p(row, _u8p), _p(act, _i8p)))


def native_dot(row: np.ndarray, act: np.ndarray) -> int:
    row = _c(row, np.uint8)
    act = _c(act, np.int8)
    return int(lib().wfo_native_dot(act.size, _p(row, _u8p), _p(act, _i8p)))


def bitnet_linear(x: np.ndarray, packed: np.ndarray, w_scale: float) -> np.ndarray:
    """scripts/direct_inference_full.py:186-208 for one row, fp32 result (before the bf16 cast)."""
    x = _c(x, np.float32).reshape(-1)
    packed = _c(packed, np.uint8)
    N, K4 = packed.shape
    out = np.empty(N, np.float32)
    lib().wfo_bitnet_linear(_p(x, _f32p), _p(packed, _u8p), w_scale, N, K4 * 4, _p(out, _f32p))
    return out


# ----------------------------------------------------------------------------- reference (_ref) wrappers
def ref_gemv(packed: np.ndarray, act: np.ndarray, scale: float, threads: int | None = None) -> np.ndarray:
    packed = _c(packed, np.uint8)
    act = _c(act, np.int8)
    N, K4 = packed.shape
    out = np.empty(N, np.float32)
    R = ref()
    if threads:
        R.wfref_set_threads(threads)
    R.wfref_gemv(_p(packed, _u8p), _p(act, _i8p), scale, N, K4 * 4, _p(out, _f32p))
    return out


def ref_gemm(packed: np.ndarray, act: np.ndarray, scale: float, threads: int | None = None) -> np.ndarray:
    packed = _c(packed, np.uint8)
    act = _c(act, np.int8)
    N, K4 = packed.shape
    M = act.shape[0]
    out = np.empty((M, N), np.float32)
    R = ref()
    if threads:
        R.wfref_set_threads(threads)
    R.wfref_gemm(_p(packed, _u8p), _p(act, _i8p), scale, M, N, K4 * 4, _p(out, _f32p))
    return out


def ref_quantize_i2_s(w: np.ndarray):
    w = _c(w, np.float32)
    N, K = w.shape
    buf = np.zeros(N * K // 4 + 64, np.uint8)
    s = ref().wfref_quantize_i2_s(_p(w, _f32p), _p(buf, _u8p), N, K)
    return buf[: N * K // 4].reshape(N, K // 4).copy(), np.float32(s)


def ref_threads() -> int:
    return int(ref().wfref_num_threads())
