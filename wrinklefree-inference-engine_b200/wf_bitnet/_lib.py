"""ctypes binding of libwf_bitnet.so (C ABI: include/wf_bitnet.h).

The library is the product: there is NO fallback.  If it is missing, or no sm_100 device is
present when a compute entry point is called, we raise — loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG_ROOT = Path(__file__).resolve().parent.parent
LIB_PATH = Path(os.environ.get("WF_BITNET_LIB", _PKG_ROOT / "libwf_bitnet.so"))

WF_OK, WF_ERR_INVALID, WF_ERR_UNSUPPORTED, WF_ERR_CUDA = 0, -1, -2, -3
WF_F32, WF_BF16, WF_F16, WF_I8 = 0, 1, 2, 3
WF_W_I2S, WF_W_TILE = 0, 1
WF_OUT_I32_PLUS1, WF_OUT_I32, WF_OUT_COMPAT_F32, WF_OUT_LINEAR = 0, 1, 2, 3
WF_EPI_RESIDUAL, WF_EPI_RELU2_MUL = 1, 2
BITNET_BLOCK_SIZE = 128


class WfError(RuntimeError):
    pass


class LinearArgs(C.Structure):
    _fields_ = [
        ("w", C.c_void_p), ("w_layout", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("q", C.c_void_p), ("M", C.c_int), ("ldq", C.c_int),
        ("rowsum", C.c_void_p), ("act_scale", C.c_void_p), ("act_scale_per_token", C.c_int),
        ("w_scale", C.c_void_p), ("nseg", C.c_int), ("seg_end", C.c_int * 4),
        ("host_scale", C.c_float), ("out_mode", C.c_int), ("epi_flags", C.c_int), ("out_dtype", C.c_int),
        ("out", C.c_void_p), ("ldo", C.c_int), ("residual", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("algo", C.c_int),
    ]


class KvLayout(C.Structure):
    _fields_ = [
        ("pool", C.c_void_p), ("max_pages", C.c_int), ("layers", C.c_int), ("kv_heads", C.c_int),
        ("page_size", C.c_int), ("head_dim", C.c_int), ("dtype", C.c_int),
    ]


class ModelConfig(C.Structure):
    _fields_ = [
        ("hidden", C.c_int), ("layers", C.c_int), ("heads", C.c_int), ("kv_heads", C.c_int), ("head_dim", C.c_int),
        ("ffn", C.c_int), ("vocab", C.c_int), ("rope_theta", C.c_float), ("eps", C.c_float),
        ("max_seq", C.c_int), ("page_size", C.c_int), ("max_batch", C.c_int), ("max_pages", C.c_int),
        ("tp_rank", C.c_int), ("tp_size", C.c_int), ("kv_dtype", C.c_int),
    ]


# every symbol include/wf_bitnet.h declares: name -> (restype, argtypes)
_vp, _i, _f, _sz, _u64 = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_uint64
SYMBOLS = {
    "wf_bitnet_available": (_i, []),
    "wf_last_error": (C.c_char_p, []),
    "wf_sm_count": (_i, []),
    "wf_version": (C.c_char_p, []),
    "wf_copy_d2d": (_i, [_vp, _vp, _sz, _vp]),
    "wf_pack_ternary": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp]),
    "wf_unpack_i2s": (_i, [_vp, _i, _i, _vp, _vp]),
    "wf_hf_to_i2s": (_i, [_vp, _i, _i, _vp, _vp]),
    "wf_tile_bytes": (_sz, [_i, _i]),
    "wf_repack_i2s_to_tile": (_i, [_vp, _i, _i, _vp, _vp]),
    "wf_repack_tile_to_i2s": (_i, [_vp, _i, _i, _vp, _vp]),
    "wf_interleave_rows16": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "wf_act_quant_i8": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "wf_linear_workspace_bytes": (_sz, [_i, _i, _i]),
    "wf_ternary_linear": (_i, [C.POINTER(LinearArgs), _vp]),
    "wf_bitnet_gemv": (_i, [_vp, _vp, _f, _i, _i, _vp, _vp, _sz, _vp]),
    "wf_bitnet_gemm": (_i, [_vp, _vp, _f, _i, _i, _i, _vp, _vp, _sz, _vp]),
    "wf_static_workspace_max_bytes": (_sz, []),
    "wf_bitnet_gemm_static": (_i, [_vp, _vp, _f, _i, _i, _i, _vp, _vp]),
    "wf_kv_pool_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "wf_kv_scatter": (_i, [C.POINTER(KvLayout), _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "wf_kv_gather": (_i, [C.POINTER(KvLayout), _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "wf_rope_kv_append": (_i, [_vp, _i, _i, C.POINTER(KvLayout), _i, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp]),
    "wf_attn_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "wf_paged_decode_attn": (_i, [_vp, C.POINTER(KvLayout), _i, _vp, _i, _vp, _i, _i, _f, _vp, _i, _vp, _sz, _vp]),
    "wf_rmsnorm_quant": (_i, [_vp, _vp, _f, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "wf_embed": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "wf_lm_head_workspace_bytes": (_sz, [_i, _i]),
    "wf_lm_head_argmax": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "wf_engine_create": (_i, [C.POINTER(ModelConfig), C.POINTER(_vp)]),
    "wf_engine_destroy": (None, [_vp]),
    "wf_engine_device_bytes": (_sz, [_vp]),
    "wf_engine_set_linear": (_i, [_vp, _i, C.c_char_p, _vp, _f, _i]),
    "wf_engine_set_tensor": (_i, [_vp, _i, C.c_char_p, _vp, _sz, _i]),
    "wf_engine_init_synthetic": (_i, [_vp, _u64]),
    "wf_engine_finalize_weights": (_i, [_vp]),
    "wf_engine_reset": (_i, [_vp]),
    "wf_engine_set_context": (_i, [_vp, _i, _i, _i]),
    "wf_engine_decode": (_i, [_vp, _vp, _i, _vp]),
    "wf_engine_decode_device": (_i, [_vp, _i, _i]),
    "wf_engine_sync": (_i, [_vp]),
    "wf_engine_stream": (_vp, [_vp]),
    "wf_engine_tp_handle_bytes": (_sz, []),
    "wf_engine_tp_export": (_i, [_vp, _vp]),
    "wf_engine_tp_attach": (_i, [_vp, _vp]),
    "wf_engine_buffer": (_vp, [_vp, C.c_char_p]),
    "wf_engine_launches_per_step": (_i, [_vp]),
    "wf_engine_time_kernel": (_i, [_vp, C.c_char_p, _i, _i, C.POINTER(_f)]),
    "wf_bench_gemv": (_i, [_i, _i, _i, _i, _i, C.POINTER(_f)]),
    "wf_gemm_stream_acc": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _i, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libwf_bitnet.so and bind every declared symbol.  Raises if the library is absent."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise WfError(
                f"{LIB_PATH} not found: build it with `make -C {_PKG_ROOT / 'csrc'}` (or __graft_entry__.build()). "
                "There is no CPU fallback."
            )
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)   # AttributeError here == header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    return lib().wf_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    """Map a C status to the reference's Python error behaviour (ValueError for bad arguments)."""
    if rc == WF_OK:
        return
    msg = last_error()
    if rc == WF_ERR_INVALID:
        raise ValueError(msg)
    if rc == WF_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise WfError(msg)


def available() -> bool:
    try:
        return bool(lib().wf_bitnet_available())
    except WfError:
        return False


def require_gpu() -> None:
    if not lib().wf_bitnet_available():
        raise WfError("no compute-capability-10.x CUDA device visible; the BitNet kernels are sm_100a-only and there is no CPU fallback")
