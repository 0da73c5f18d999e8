"""C-ABI checks that need no GPU: the library loads, exports every symbol include/wf_bitnet.h declares,
argument validation returns the documented statuses, and the product refuses to run without a device
instead of falling back to anything."""
import ctypes as C
import re
from pathlib import Path

import pytest
import torch

from wf_bitnet import _lib

ROOT = Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "wf_bitnet.h").read_text()


def declared_functions():
    # prototypes look like:  <type> wf_name(...);
    names = set(re.findall(r"\b(wf_[a-z0-9_]+)\s*\(", HEADER))
    return sorted(n for n in names if n not in ("wf_linear_args", "wf_kv_layout", "wf_model_config", "wf_engine"))


def test_library_builds_and_loads():
    L = _lib.lib()
    assert L.wf_version().startswith(b"wf_bitnet")


def test_every_declared_symbol_is_exported_and_bound():
    L = C.CDLL(str(_lib.LIB_PATH))
    decl = declared_functions()
    assert len(decl) >= 40
    for name in decl:
        assert hasattr(L, name), f"{name} declared in wf_bitnet.h but not exported"
        assert name in _lib.SYMBOLS, f"{name} has no ctypes signature in wf_bitnet/_lib.py"
    for name in _lib.SYMBOLS:
        assert name in decl, f"{name} bound in _lib.py but not declared in the header"


def test_struct_sizes_match_header_layout():
    # sizes follow the field lists of the header (LP64): catches drift between header and ctypes mirror
    assert C.sizeof(_lib.KvLayout) == 8 + 6 * 4
    assert C.sizeof(_lib.ModelConfig) == 16 * 4   # + kv_dtype
    assert C.sizeof(_lib.LinearArgs) % 8 == 0 and C.sizeof(_lib.LinearArgs) >= 120


def test_pure_size_helpers():
    L = _lib.lib()
    assert L.wf_tile_bytes(2560, 2560) == 2560 * 2560 // 4
    assert L.wf_tile_bytes(6912, 2560) == 6912 * 2560 // 4
    assert L.wf_tile_bytes(17, 128) == 2 * 512            # rows padded to 16
    assert L.wf_tile_bytes(16, 100) == 0                  # K % 128 != 0
    assert L.wf_kv_pool_bytes(4, 30, 5, 16, 128, _lib.WF_BF16) == 4 * 30 * 2 * 5 * 16 * 128 * 2
    assert L.wf_linear_workspace_bytes(8, 3840, 2560) >= 8 * 3840 * 4
    assert L.wf_attn_workspace_bytes(64, 20, 128, 4096) >= 64 * 20 * 64 * 130 * 4
    assert L.wf_lm_head_workspace_bytes(64, 128256) >= 64 * 1002 * 8


def test_argument_validation_without_gpu():
    L = _lib.lib()
    a = _lib.LinearArgs()
    assert L.wf_ternary_linear(C.byref(a), None) == _lib.WF_ERR_INVALID
    assert b"null" in L.wf_last_error()
    assert L.wf_ternary_linear(None, None) == _lib.WF_ERR_INVALID
    # K not a multiple of 128 -> the reference's ValueError text
    buf = (C.c_uint8 * 64)()
    a.w = C.addressof(buf); a.q = C.addressof(buf); a.out = C.addressof(buf)
    a.N, a.K, a.M, a.ldq, a.ldo = 4, 64, 1, 64, 4
    assert L.wf_ternary_linear(C.byref(a), None) == _lib.WF_ERR_INVALID
    assert b"must be multiple of 128" in L.wf_last_error()
    assert L.wf_repack_i2s_to_tile(C.addressof(buf), 16, 100, C.addressof(buf), None) == _lib.WF_ERR_INVALID
    assert L.wf_pack_ternary(C.addressof(buf), 0, 2, 6, C.addressof(buf), C.addressof(buf), None) == _lib.WF_ERR_INVALID
    assert b"divisible by 4" in L.wf_last_error()
    with pytest.raises(ValueError):
        _lib.check(_lib.WF_ERR_INVALID)
    with pytest.raises(NotImplementedError):
        _lib.check(_lib.WF_ERR_UNSUPPORTED)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device behaviour")
def test_no_device_means_loud_failure_not_fallback():
    from wf_bitnet import ops
    from sgl_kernel.quantization import bitnet_check_kernel_available, bitnet_gemv, bitnet_quantize_activations
    assert _lib.available() is False and bitnet_check_kernel_available() is False
    with pytest.raises(_lib.WfError):
        _lib.require_gpu()
    with pytest.raises(_lib.WfError):
        ops.pack_ternary(torch.zeros(4, 128))
    with pytest.raises(Exception):
        bitnet_quantize_activations(torch.zeros(128))           # CPU tensor: no kernel registered for CPU
    with pytest.raises(Exception):
        bitnet_gemv(torch.zeros(4, 32, dtype=torch.uint8), torch.zeros(128, dtype=torch.int8), 1.0)
    from wrinklefree_inference.kernels import NATIVE_KERNELS_AVAILABLE, patch_model_with_native_kernels
    assert NATIVE_KERNELS_AVAILABLE is False
    assert patch_model_with_native_kernels(torch.nn.Linear(128, 4)) == 0   # warning + no-op, like the reference


def test_product_never_imports_oracle():
    pkg = ROOT / "wrinklefree-inference-engine_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")):
        txt = f.read_text()
        assert "import oracle" not in txt and "from oracle" not in txt and "libwf_oracle" not in txt and "libwf_ref" not in txt, f
