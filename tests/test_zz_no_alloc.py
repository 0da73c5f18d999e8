"""wf_bitnet_gemm_static: the allocation-free twin of wf_bitnet_gemm (scratch in a static __device__ array, csrc/no_alloc.cu).

No counterpart in the reference; the checker is the existing entry point, which the
other GPU tests pin against the oracle.  Compat outputs are fp32 `scale * (float)sum` of exact integers, so the bar is bit-equality.
Collected last on purpose (pytest -x must not lose the other tests to a failure here); passes alone and after the rest.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(m, n, k, seed):
    g = torch.Generator().manual_seed(seed)
    packed = torch.randint(0, 256, (n, k // 4), dtype=torch.uint8, generator=g)
    # I2_S codes are 0..2; remove the 3s (keep the byte layout arbitrary otherwise)
    b = packed.to(torch.int32)
    for sh in (0, 2, 4, 6):
        f = (b >> sh) & 3
        b = b - ((f == 3).to(torch.int32) << sh)
    packed = b.to(torch.uint8)
    x = torch.randint(-127, 128, (m, k), dtype=torch.int8, generator=g)
    return packed.cuda(), x.cuda()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(L, packed, x, scale, m, n, k):
    from wf_bitnet import ops
    y = torch.full((m, n), float("nan"), dtype=torch.float32, device="cuda")
    ws = ops.workspace(packed.device, int(L.wf_linear_workspace_bytes(m, n, k)))
    rc = L.wf_bitnet_gemm(packed.data_ptr(), x.data_ptr(), scale, m, n, k, y.data_ptr(), ws.data_ptr(), ws.numel(), _stream())
    assert rc == 0, L.wf_last_error().decode()
    return y


def _static(L, packed, x, scale, m, n, k):
    y = torch.full((m, n), float("nan"), dtype=torch.float32, device="cuda")
    rc = L.wf_bitnet_gemm_static(packed.data_ptr(), x.data_ptr(), scale, m, n, k, y.data_ptr(), _stream())
    assert rc == 0, L.wf_last_error().decode()
    return y


@pytest.mark.parametrize("m,n,k", [(1, 2560, 2560), (1, 6912, 2560), (8, 2560, 6912), (64, 3840, 2560), (3, 17, 256), (1, 28672, 4096)])
def test_static_scratch_entry_point_equals_workspace_entry_point(m, n, k):
    from wf_bitnet._lib import lib
    L = lib()
    assert int(L.wf_linear_workspace_bytes(m, n, k)) <= int(L.wf_static_workspace_max_bytes())
    packed, x = _case(m, n, k, 1000 * m + n + k)
    want = _existing(L, packed, x, 0.0123, m, n, k)
    torch.cuda.synchronize()
    got = torch.empty_like(want)
    for call in range(2):                       # two calls in a row share the array: the second must see clean scratch
        got.fill_(float("nan"))
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        rc = L.wf_bitnet_gemm_static(packed.data_ptr(), x.data_ptr(), 0.0123, m, n, k, got.data_ptr(), _stream())
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
        assert rc == 0, L.wf_last_error().decode()
        assert torch.equal(got, want), f"call {call}: differs from wf_bitnet_gemm"
        if call:                                # the first call may page kernel code in (lazy module loading); the second may not move
            assert free1 == free0, "device memory in use changed across an allocation-free call"
    # a different shape, then the first one again: stale contents of a larger / smaller previous call must not leak in
    p2, x2 = _case(5, 1024, 512, 7)
    assert torch.equal(_static(L, p2, x2, 1.5, 5, 1024, 512), _existing(L, p2, x2, 1.5, 5, 1024, 512))
    assert torch.equal(_static(L, packed, x, 0.0123, m, n, k), want)


def test_static_scratch_rejects_sizes_beyond_the_array():
    from wf_bitnet._lib import lib
    L = lib()
    cap = int(L.wf_static_workspace_max_bytes())
    assert cap == 4 << 20
    m, n, k = 64, 32768, 128                    # 8 MiB of accumulators
    assert int(L.wf_linear_workspace_bytes(m, n, k)) > cap
    y = torch.zeros(1, device="cuda")
    rc = L.wf_bitnet_gemm_static(y.data_ptr(), y.data_ptr(), 1.0, m, n, k, y.data_ptr(), _stream())   # rejected before any launch
    assert rc == -2
    msg = L.wf_last_error().decode()
    assert "static array holds" in msg and str(cap) in msg
    assert L.wf_bitnet_gemm_static(y.data_ptr(), y.data_ptr(), 1.0, 1, 0, 128, y.data_ptr(), _stream()) == -1
    torch.cuda.synchronize()
