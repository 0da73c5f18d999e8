"""GPU parity: the tcgen05 (kind::i8) prefill GEMM.  int32 accumulators bit-exact vs the oracle on sizes it
finishes in seconds, fused epilogues identical to the mma.sync path, and at the BASELINE prefill size
(8 x 2048 tokens) size-independent properties: equality with the already-pinned mma.sync kernel on the same
inputs, linearity in the activations and an oracle check of a random sample of rows."""
import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def make(n, k, m, seed, raw=False):
    rng = np.random.default_rng(seed)
    w = rng.integers(-1, 2, size=(n, k)).astype(np.float32)
    packed = O.pack_ternary(w)[0]
    if raw:
        q = rng.integers(-128, 128, size=(m, k)).astype(np.int8)
        return packed, q, None, q.astype(np.int32).sum(1).astype(np.int32)
    x = (rng.standard_normal((m, k)) * 3).astype(np.float32)
    q, s, rs = O.quantize_act_i8_rows(x)
    return packed, q, s, rs


@pytest.mark.parametrize("nk", [(2560, 2560), (3840, 2560), (2560, 6912), (200, 640), (128, 128), (1000, 384)])
@pytest.mark.parametrize("m", [128, 130, 333, 512])
def test_tc_accumulators_bit_exact(nk, m):
    from wf_bitnet import ops
    n, k = nk
    packed, q, _, rs = make(n, k, m, n * 13 + k + m, raw=(m == 333))
    want = O.gemm_i32_fast(packed, q, 0)
    pi = cu(packed)
    tile = ops.i2s_to_tile(pi)
    for layout, wt in (("i2s", pi), ("tile", tile)):
        got = ops.ternary_linear(wt, n, k, cu(q), layout=layout, mode="i32_plus1", algo=3).cpu().numpy()
        assert np.array_equal(got, want), layout
    got = ops.ternary_linear(tile, n, k, cu(q), mode="i32", rowsum=cu(rs), algo=3).cpu().numpy()
    assert np.array_equal(got, O.gemm_i32_fast(packed, q, 1))


def test_tc_small_m_and_extremes():
    """M below one tile (TMA zero-fills the missing tokens) and the adversarial +-max inputs (|acc| = 2*128*K)."""
    from wf_bitnet import ops
    packed, q, _, _ = make(384, 512, 5, 3, raw=True)
    got = ops.ternary_linear(cu(packed), 384, 512, cu(q), layout="i2s", mode="i32_plus1", algo=3).cpu().numpy()
    assert np.array_equal(got, O.gemm_i32_fast(packed, q, 0))
    ones = cu(O.pack_ternary(np.ones((256, 28672), np.float32))[0])
    for val in (127, -128):
        a = torch.full((130, 28672), val, dtype=torch.int8, device="cuda")
        out = ops.ternary_linear(ones, 256, 28672, a, layout="i2s", mode="i32_plus1", algo=3)
        assert torch.all(out == 2 * val * 28672)


def test_tc_fused_epilogues_match_mma_path():
    from wf_bitnet import ops
    rng = np.random.default_rng(11)
    m, k, n, ffn = 300, 512, 640, 256
    packed, q, s, rs = make(n, k, m, 5)
    tile = ops.i2s_to_tile(cu(packed))
    ws = torch.tensor([0.7, 1.3, 2.1]).cuda()
    kw = dict(mode="linear", rowsum=cu(rs), act_scale=cu(s))
    a = ops.ternary_linear(tile, n, k, cu(q), w_scale=ws, seg_end=[384, 512, 640], out_dtype=torch.float32, algo=3, **kw)
    b = ops.ternary_linear(tile, n, k, cu(q), w_scale=ws, seg_end=[384, 512, 640], out_dtype=torch.float32, algo=2, **kw)
    assert torch.equal(a, b)
    from test_gpu_linear import ref_dequant          # the reference's fp32 rounding sequence (direct_inference_full.py:202-206)
    scale_n = np.concatenate([np.full(384, 0.7), np.full(128, 1.3), np.full(128, 2.1)]).astype(np.float32)
    assert np.array_equal(a.cpu().numpy(), ref_dequant(packed, q, rs, scale_n, s))
    res = torch.from_numpy(rng.standard_normal((m, n)).astype(np.float32) * 50).to(torch.bfloat16).cuda()
    w1 = torch.tensor([1.1]).cuda()
    a = ops.ternary_linear(tile, n, k, cu(q), w_scale=w1, residual=res, out_dtype=torch.bfloat16, algo=3, **kw)
    lin = torch.from_numpy(ref_dequant(packed, q, rs, np.float32(1.1), s)).to(torch.bfloat16)
    assert torch.equal(a.cpu(), res.cpu() + lin)
    wg = rng.integers(-1, 2, size=(ffn, k)).astype(np.float32)
    wu = rng.integers(-1, 2, size=(ffn, k)).astype(np.float32)
    pg, pu = O.pack_ternary(wg)[0], O.pack_ternary(wu)[0]
    tile_gu = ops.i2s_to_tile(ops.interleave_rows16(cu(pg), cu(pu)))
    sg, su = np.float32(0.9), np.float32(1.6)
    g = torch.from_numpy(ref_dequant(pg, q, rs, sg, s)).to(torch.bfloat16)
    u = torch.from_numpy(ref_dequant(pu, q, rs, su, s)).to(torch.bfloat16)
    got = ops.ternary_linear(tile_gu, 2 * ffn, k, cu(q), w_scale=torch.tensor([float(sg), float(su)]).cuda(), relu2_mul=True,
                             out_dtype=torch.bfloat16, algo=3, **kw).cpu()
    assert got.shape == (m, ffn) and torch.equal(got, torch.relu(g).square() * u)


@pytest.mark.parametrize("nk", [(3840, 2560), (2560, 6912)])
def test_tc_prefill_size_properties(nk):
    """BASELINE config 3 size: M = 8 x 2048 tokens."""
    from wf_bitnet import ops
    n, k = nk
    m = 16384
    rng = np.random.default_rng(n)
    packed = O.pack_ternary(rng.integers(-1, 2, size=(n, k)).astype(np.float32))[0]
    tile = ops.i2s_to_tile(cu(packed))
    g = torch.Generator(device="cuda").manual_seed(1)
    q1 = torch.randint(-64, 64, (m, k), generator=g, device="cuda", dtype=torch.int8)
    q2 = torch.randint(-63, 64, (m, k), generator=g, device="cuda", dtype=torch.int8)
    y1 = ops.ternary_linear(tile, n, k, q1, mode="i32_plus1", algo=3)
    y2 = ops.ternary_linear(tile, n, k, q2, mode="i32_plus1", algo=3)
    y12 = ops.ternary_linear(tile, n, k, q1 + q2, mode="i32_plus1", algo=3)
    assert torch.equal(y1 + y2, y12)                                                   # linearity (exact integers)
    assert torch.equal(y1, ops.ternary_linear(tile, n, k, q1, mode="i32_plus1", algo=2))   # the pinned mma.sync kernel
    rows = rng.choice(m, size=64, replace=False)
    want = O.gemm_i32_fast(packed, q1[torch.from_numpy(rows).cuda()].cpu().numpy(), 0)
    assert np.array_equal(y1[torch.from_numpy(rows).cuda()].cpu().numpy(), want)       # oracle on a random sample of tokens
