"""GPU parity: ternary GEMV/GEMM.  int32 accumulators bit-exact vs the oracle (and vs the reference's own
SIMD kernel in oracle/_ref), compat outputs bit-exact in fp32, dequantised outputs within 1e-3 relative."""
import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu

SHAPES_2B = [(2560, 2560), (640, 2560), (6912, 2560), (2560, 6912), (3840, 2560), (13824, 2560)]


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def ref_dequant(packed, q, rs, ws, s):
    """The reference's dequantised linear output with ITS fp32 rounding sequence (scripts/direct_inference_full.py:202-206):
    fl32( fl32( fl32(S * s_w) - fl32(s_w * sum a) ) * s_a ),  S = sum (w+1) a (the kernel's accumulator).  ws: scalar or [N]."""
    S = O.gemm_i32_fast(packed, q, 0).astype(np.float32)
    ws = np.asarray(ws, np.float32)
    wsr = ws[None, :] if ws.ndim else ws
    y = S * wsr
    c = (wsr * rs.astype(np.float32)[:, None]).astype(np.float32)
    return ((y - c).astype(np.float32) * s.astype(np.float32)[:, None]).astype(np.float32)


def make(n, k, m, seed):
    rng = np.random.default_rng(seed)
    w = rng.integers(-1, 2, size=(n, k)).astype(np.float32)
    packed = O.pack_ternary(w)[0]
    x = (rng.standard_normal((m, k)) * 3).astype(np.float32)
    q, s, rs = O.quantize_act_i8_rows(x)
    return w, packed, x, q, s, rs


@pytest.mark.parametrize("nk", SHAPES_2B + [(128, 128), (48, 384), (17, 256), (1024, 4096), (4096, 14336)])
@pytest.mark.parametrize("m", [1, 3, 8])
def test_accumulators_bit_exact_small_batch(nk, m):
    from wf_bitnet import ops
    n, k = nk
    _, packed, _, q, _, rs = make(n, k, m, n * 31 + k + m)
    want = O.gemm_i32_fast(packed, q, 0)
    pi = cu(packed)
    tile = ops.i2s_to_tile(pi)
    qd = cu(q)
    for layout, wt, algos in (("tile", tile, (1, 2)), ("i2s", pi, (2,))):
        for algo in algos:
            got = ops.ternary_linear(wt, n, k, qd, layout=layout, mode="i32_plus1", algo=algo).cpu().numpy()
            assert np.array_equal(got, want), (layout, algo)
            got2 = ops.ternary_linear(wt, n, k, qd, layout=layout, mode="i32_plus1", algo=algo).cpu().numpy()
            assert np.array_equal(got2, want), "second call: split-K workspace was not left clean"
    got = ops.ternary_linear(tile, n, k, qd, mode="i32", rowsum=cu(rs)).cpu().numpy()
    assert np.array_equal(got, O.gemm_i32_fast(packed, q, 1))


@pytest.mark.parametrize("nk,m", [((28672, 4096), 1), ((28672, 4096), 5), ((8192, 8192), 1), ((8192, 8192), 2), ((10240, 8192), 1),
                                  ((4096, 14336), 2), ((8208, 4096), 1), ((2576, 8192), 4)])
def test_gemv_large_shapes_every_slab_mode(nk, m):
    """8B / 70B-sized matrices: two CTAs per SM, whole-row-block slabs (28672x4096), one-group-per-CTA (4096x14336) and the
    even split with the split-K round (8192x8192, odd row-block counts); dp4a (m = 1) and mma (m > 1) consumers."""
    from wf_bitnet import ops
    n, k = nk
    rng = np.random.default_rng(n + 3 * k + m)
    w = rng.integers(-1, 2, size=(n, k)).astype(np.float32)
    packed = O.pack_ternary(w)[0]
    q = rng.integers(-128, 128, size=(m, k)).astype(np.int8)
    want = O.gemm_i32_fast(packed, q, 0)
    tile = ops.i2s_to_tile(cu(packed))
    for _ in range(2):   # twice: the self-cleaning workspace must be left zero
        got = ops.ternary_linear(tile, n, k, cu(q), mode="i32_plus1", algo=1).cpu().numpy()
        assert np.array_equal(got, want)


@pytest.mark.parametrize("nk", [(57344, 8192), (8192, 28672)])
def test_gemv_70b_shapes(nk):
    """The 70B matrices wf_bench_gemv times and reports bandwidth on: gate|up (57344 x 8192, 117 MB packed) and down
    (8192 x 28672, 224 k-blocks: the longest row dot of any BASELINE config, |acc| up to 7.3 M), M = 1 (dp4a), 2 and 8 (mma).
    The expected accumulators come straight from the ternary matrix in exact float64 row chunks (no unpack in between);
    the packed bytes are the C oracle's (quantize_i2_s restatement)."""
    from wf_bitnet import ops
    n, k = nk
    rng = np.random.default_rng(n + 5 * k)
    w = rng.integers(-1, 2, size=(n, k), dtype=np.int8)
    packed = np.concatenate([O.pack_ternary(w[r0:r0 + 8192].astype(np.float32))[0] for r0 in range(0, n, 8192)])
    tile = ops.i2s_to_tile(cu(packed))
    for m in (1, 2, 8):
        q = rng.integers(-128, 128, size=(m, k)).astype(np.int8)
        q[0, : k // 2] = 127          # drive one row towards the accumulator range limit
        want = np.empty((m, n), np.int32)
        qf = q.astype(np.float64).T
        for r0 in range(0, n, 4096):
            want[:, r0:r0 + 4096] = np.rint((w[r0:r0 + 4096].astype(np.float64) + 1.0) @ qf).astype(np.int64).astype(np.int32).T
        # algo 0 = the library's own choice: the streamed GEMV wherever the int8 rows fit in shared memory next to the weight ring,
        # the tiled GEMM for 8 rows of 28672 (they do not)
        for algo in ((0, 1) if m * k <= 4 * 28672 else (0,)):
            for _ in range(2):   # twice: the self-cleaning split-K workspace must be left zero
                got = ops.ternary_linear(tile, n, k, cu(q), mode="i32_plus1", algo=algo).cpu().numpy()
                assert np.array_equal(got, want), (nk, m, algo)


def test_gemv_large_relu2_epilogue():
    """fused gate|up with relu^2*up at the 8B ffn width through the GEMV (two-row-block finalisation groups, two CTAs per SM)."""
    from wf_bitnet import ops
    rng = np.random.default_rng(11)
    k, ffn, m = 4096, 14336, 3
    q = rng.integers(-100, 101, size=(m, k)).astype(np.int8)
    rs = q.astype(np.int64).sum(axis=1).astype(np.int32)
    s = np.array([0.011, 0.02, 0.007], np.float32)
    wg = rng.integers(-1, 2, size=(ffn, k)).astype(np.float32)
    wu = rng.integers(-1, 2, size=(ffn, k)).astype(np.float32)
    pg, pu = O.pack_ternary(wg)[0], O.pack_ternary(wu)[0]
    tile_gu = ops.i2s_to_tile(ops.interleave_rows16(cu(pg), cu(pu)))
    sg, su = np.float32(0.9), np.float32(1.6)
    g = torch.from_numpy(ref_dequant(pg, q, rs, sg, s)).to(torch.bfloat16)
    u = torch.from_numpy(ref_dequant(pu, q, rs, su, s)).to(torch.bfloat16)
    want_h = torch.relu(g).square() * u
    got = ops.ternary_linear(tile_gu, 2 * ffn, k, cu(q), mode="linear", rowsum=cu(rs), act_scale=cu(s),
                             w_scale=torch.tensor([float(sg), float(su)]).cuda(), relu2_mul=True, out_dtype=torch.bfloat16, algo=1).cpu()
    assert torch.equal(got, want_h)


@pytest.mark.parametrize("nk", [(2560, 2560), (13824, 2560), (2560, 6912), (200, 640)])
@pytest.mark.parametrize("m", [9, 16, 17, 32, 40, 64, 100])
def test_accumulators_bit_exact_batched(nk, m):
    from wf_bitnet import ops
    n, k = nk
    _, packed, _, q, _, _ = make(n, k, m, n + k * 7 + m)
    want = O.gemm_i32_fast(packed, q, 0)
    tile = ops.i2s_to_tile(cu(packed))
    got = ops.ternary_linear(tile, n, k, cu(q), mode="i32_plus1", algo=2).cpu().numpy()
    assert np.array_equal(got, want)
    got = ops.ternary_linear(cu(packed), n, k, cu(q), layout="i2s", mode="i32_plus1", algo=2).cpu().numpy()
    assert np.array_equal(got, want)


def test_raw_int8_activations_extremes():
    """extern/BitNet/gpu/test.py:32-61 style: random int8 in [-128,127], exact equality; plus the adversarial
    all-(+1) x 127 case where the reference's AVX2 int16 lanes wrap — exact oracle only."""
    from wf_bitnet import ops
    rng = np.random.default_rng(0)
    n, k = 2560, 6912
    w = rng.integers(-1, 2, size=(n, k)).astype(np.float32)
    packed = O.pack_ternary(w)[0]
    q = rng.integers(-128, 128, size=(4, k)).astype(np.int8)
    tile = ops.i2s_to_tile(cu(packed))
    for algo in (1, 2):
        assert np.array_equal(ops.ternary_linear(tile, n, k, cu(q), mode="i32_plus1", algo=algo).cpu().numpy(), O.gemm_i32_fast(packed, q, 0))
    ones = O.pack_ternary(np.ones((32, 4096), np.float32))[0]
    a = np.full((1, 4096), 127, np.int8)
    a2 = np.full((1, 4096), -128, np.int8)
    t1 = ops.i2s_to_tile(cu(ones))
    for algo in (1, 2):
        assert np.all(ops.ternary_linear(t1, 32, 4096, cu(a), mode="i32_plus1", algo=algo).cpu().numpy() == 2 * 127 * 4096)
        assert np.all(ops.ternary_linear(t1, 32, 4096, cu(a2), mode="i32_plus1", algo=algo).cpu().numpy() == -2 * 128 * 4096)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("nk", [(2560, 2560), (6912, 2560), (2560, 6912)])
def test_matches_reference_simd_kernel(nk):
    """The unmodified ggml_vec_dot_i2_i8_s (oracle/_ref) on the same packed bytes and int8 activations."""
    from sgl_kernel.quantization import bitnet_gemv, bitnet_gemm
    n, k = nk
    _, packed, _, q, _, _ = make(n, k, 3, n ^ k)
    y_ref = O.ref_gemv(packed, q[0], 0.0123)
    y = bitnet_gemv(cu(packed), cu(q[0]), 0.0123).cpu().numpy()
    assert np.array_equal(y, y_ref)                     # fp32 bit-exact: (float)sumi * scale
    Y = bitnet_gemm(cu(packed), cu(q), 1.7).cpu().numpy()
    assert np.array_equal(Y, O.ref_gemm(packed, q, 1.7))


def test_compat_golden(golden):
    from sgl_kernel.quantization import bitnet_gemv, bitnet_gemm, bitnet_quantize_activations
    g = golden["gemv"]
    for tag in ("a", "b", "c"):
        y = bitnet_gemv(cu(g[f"packed_{tag}"]), cu(g[f"xq_{tag}"]), 1.7).cpu().numpy()
        assert np.array_equal(y, g[f"y_{tag}"])
    assert np.all(bitnet_gemv(cu(g["ones_packed"]), torch.ones(256, dtype=torch.int8, device="cuda"), 2.0).cpu().numpy() == 1024.0)
    q, s = bitnet_quantize_activations(cu(g["gemm_X"]))
    assert np.array_equal(q.cpu().numpy(), g["gemm_Xq"]) and s == float(np.float32(float(g["gemm_sx"])))
    assert np.array_equal(bitnet_gemm(cu(g["gemm_packed"]), q, 0.75).cpu().numpy(), g["gemm_Y"])
    # GEMV == GEMM row 0 (tests/test_bitnet_weight_loading.py:270-302 asks cos > 0.999; integers give equality)
    assert np.array_equal(bitnet_gemv(cu(g["gemm_packed"]), q[0].contiguous(), 0.75).cpu().numpy(), g["gemm_Y"][0])


@pytest.mark.parametrize("nk", [(2560, 2560), (6912, 2560), (2560, 6912)])
@pytest.mark.parametrize("m", [1, 4, 64])
def test_linear_mode_within_1e3(nk, m):
    """dequantised outputs vs the reference formula (out - s_w*sum a)*s_a (direct_inference_full.py:186-208)."""
    from wf_bitnet import ops
    n, k = nk
    w, packed, x, q, s, rs = make(n, k, m, 99 + n + m)
    ws = 1.7969
    want = np.stack([O.bitnet_linear(x[i], packed, ws) for i in range(m)])
    tile = ops.i2s_to_tile(cu(packed))
    got = ops.ternary_linear(tile, n, k, cu(q), mode="linear", rowsum=cu(rs), act_scale=cu(s),
                             w_scale=torch.tensor([ws], device="cuda"), out_dtype=torch.float32).cpu().numpy()
    rel = np.linalg.norm(got - want) / np.linalg.norm(want)
    assert rel <= 1e-3, rel            # the north-star bar
    # the epilogue follows the reference's own rounding sequence fl32(fl32(fl32(S*s_w) - fl32(s_w*sum a)) * s_a), so the fp32
    # values are in fact IDENTICAL to the CPU path's, not merely close
    assert np.array_equal(got, want), np.abs(got - want).max()
    gb = ops.ternary_linear(tile, n, k, cu(q), mode="linear", rowsum=cu(rs), act_scale=cu(s),
                            w_scale=torch.tensor([ws], device="cuda"), out_dtype=torch.bfloat16).float().cpu().numpy()
    wb = torch.from_numpy(want).to(torch.bfloat16).float().numpy()
    assert np.array_equal(gb, wb)
    # and against the float matmul on the unquantised input (reference tests ask cos > 0.99)
    full = ws * (x @ w.T)
    cos = (got * full).sum() / (np.linalg.norm(got) * np.linalg.norm(full))
    assert cos > 0.99


@pytest.mark.parametrize("m", [1, 5, 32])
def test_fused_epilogues(m):
    from wf_bitnet import ops
    rng = np.random.default_rng(7)
    k, ffn, hid = 512, 256, 384
    # --- segments (fused q|k|v with three scales)
    n = 640
    _, packed, x, q, s, rs = make(n, k, m, 5)
    ws = torch.tensor([0.7, 1.3, 2.1])
    seg = [384, 512, 640]
    scale_n = np.concatenate([np.full(384, 0.7), np.full(128, 1.3), np.full(128, 2.1)]).astype(np.float32)
    want = ref_dequant(packed, q, rs, scale_n, s)
    tile = ops.i2s_to_tile(cu(packed))
    for algo in ((1, 2) if m <= 8 else (2,)):
        got = ops.ternary_linear(tile, n, k, cu(q), mode="linear", rowsum=cu(rs), act_scale=cu(s), w_scale=ws.cuda(), seg_end=seg,
                                 out_dtype=torch.float32, algo=algo).cpu().numpy()
        assert np.array_equal(got, want), algo
    # --- residual: hidden + proj in bf16
    res = torch.from_numpy(rng.standard_normal((m, n)).astype(np.float32) * 50).to(torch.bfloat16)
    w1 = torch.tensor([1.1])
    lin = torch.from_numpy(ref_dequant(packed, q, rs, np.float32(1.1), s)).to(torch.bfloat16)
    want_r = res + lin
    for algo in ((1, 2) if m <= 8 else (2,)):
        got = ops.ternary_linear(tile, n, k, cu(q), mode="linear", rowsum=cu(rs), act_scale=cu(s), w_scale=w1.cuda(),
                                 residual=res.cuda(), out_dtype=torch.bfloat16, algo=algo).cpu()
        assert torch.equal(got, want_r), algo
    # --- relu^2 * up on 16-row interleaved gate|up
    wg = rng.integers(-1, 2, size=(ffn, k)).astype(np.float32)
    wu = rng.integers(-1, 2, size=(ffn, k)).astype(np.float32)
    pg, pu = O.pack_ternary(wg)[0], O.pack_ternary(wu)[0]
    il = ops.interleave_rows16(cu(pg), cu(pu))
    tile_gu = ops.i2s_to_tile(il)
    sg, su = np.float32(0.9), np.float32(1.6)
    g = torch.from_numpy(ref_dequant(pg, q, rs, sg, s)).to(torch.bfloat16)
    u = torch.from_numpy(ref_dequant(pu, q, rs, su, s)).to(torch.bfloat16)
    want_h = torch.relu(g).square() * u
    for algo in ((1, 2) if m <= 8 else (2,)):
        got = ops.ternary_linear(tile_gu, 2 * ffn, k, cu(q), mode="linear", rowsum=cu(rs), act_scale=cu(s),
                                 w_scale=torch.tensor([float(sg), float(su)]).cuda(), relu2_mul=True, out_dtype=torch.bfloat16,
                                 algo=algo).cpu()
        assert got.shape == (m, ffn) and torch.equal(got, want_h), algo


def test_errors_match_reference_behaviour():
    from sgl_kernel.quantization import bitnet_gemv
    from wf_bitnet import ops
    with pytest.raises(ValueError, match="multiple of 128"):
        bitnet_gemv(torch.zeros(4, 16, dtype=torch.uint8, device="cuda"), torch.zeros(64, dtype=torch.int8, device="cuda"), 1.0)
    with pytest.raises(ValueError):
        ops.ternary_linear(torch.zeros(512, dtype=torch.uint8, device="cuda"), 16, 128, torch.zeros(1, 128, dtype=torch.int8, device="cuda"), mode="linear")
    with pytest.raises(ValueError):
        ops.i2s_to_tile(torch.zeros(4, 16, dtype=torch.uint8, device="cuda"))
    # empty batch is a no-op
    out = ops.ternary_linear(torch.zeros(512, dtype=torch.uint8, device="cuda"), 16, 128, torch.zeros(0, 128, dtype=torch.int8, device="cuda"))
    assert out.shape == (0, 16)


def test_wrapper_module_and_native_shim():
    from wrinklefree_inference.kernels import BitNetLinearNative, patch_model_with_native_kernels, NATIVE_KERNELS_AVAILABLE, benchmark_kernels
    from wrinklefree_inference.kernels.bitnet_patch import pack_ternary_weights
    from wrinklefree_inference.kernels import native
    assert NATIVE_KERNELS_AVAILABLE
    gen = torch.Generator().manual_seed(2)
    w = torch.randint(-1, 2, (384, 256), generator=gen).float()
    packed, scale = pack_ternary_weights(w)
    assert np.array_equal(packed.cpu().numpy(), O.pack_ternary(w.numpy())[0]) and scale.item() == 1.0
    lin = BitNetLinearNative(256, 384, packed, torch.tensor([1.5]), bias=torch.ones(384))
    x = torch.randn(2, 3, 256, generator=gen)
    y = lin(x.cuda()).cpu()
    ref = 1.5 * (x @ w.t()) + 1.0
    assert y.shape == (2, 3, 384)
    cos = torch.nn.functional.cosine_similarity(y.flatten(), ref.flatten(), dim=0).item()
    assert cos > 0.999
    # exact statement of what forward computes: per-tensor quant, exact integer dot, fp32 scales
    q, s = O.quantize_act_i8(x.reshape(-1, 256).numpy())
    exact = (O.gemm_i32_fast(O.pack_ternary(w.numpy())[0], q, 1).astype(np.float32) * np.float32(1.5)) * np.float32(s) + 1.0
    assert np.allclose(y.reshape(-1, 384).numpy(), exact, rtol=1e-6, atol=1e-5)

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a = torch.nn.Linear(256, 128, bias=False)
            self.b = torch.nn.Linear(100, 8)       # K % 128 != 0 -> left alone (bitnet_patch.py:186)
    net = Net()
    net.a.weight.data = torch.randint(-1, 2, (128, 256)).float()
    assert patch_model_with_native_kernels(net) == 1 and isinstance(net.a, BitNetLinearNative) and isinstance(net.b, torch.nn.Linear)
    # kernels/native.py shim: LSB layout in, scale * sum w*a out (reference tests/test_bitnet_model.py:7-91 cosine > 0.9999)
    kern = native.get_kernel()
    wn = np.random.default_rng(4).integers(-1, 2, size=(128, 256)).astype(np.float32)
    xa = np.random.default_rng(5).integers(-100, 100, size=(8, 256)).astype(np.int8)
    out = kern.gemm(native.pack_weights(wn), xa, 0.5)
    assert np.array_equal(out, (xa.astype(np.float32) @ wn.T) * 0.5)
    assert np.array_equal(kern.gemv(native.pack_weights(wn), xa[0], 2.0), (xa[0].astype(np.float32) @ wn.T) * 2.0)
    r = benchmark_kernels(in_features=256, out_features=256, batch_size=1, num_iterations=5)
    assert "native_ms" in r and r["native_ms"] > 0


@pytest.mark.parametrize("nk", [(4096, 4096), (4096, 14336), (6144, 4096), (28672, 4096), (2560, 6912), (3840, 2560), (13824, 2560), (1040, 1024), (8192, 8192)])
@pytest.mark.parametrize("m", [1, 2, 5, 8, 13, 16, 24, 32])
def test_streaming_gemm_accumulators_bit_exact(nk, m):
    """csrc/gemm_stream.cu on its own (wf_gemm_stream_acc): sum (w+1)*a as exact integers in fp32 == the oracle's int32 row dots, on one
    window (K <= 4096), two windows of 27 k-blocks (6912), four of 28 (14336), a short last row-block (1040 rows) — and again on a second
    and third launch into a re-zeroed accumulator (the shared tile and the ring leave no state behind)."""
    import ctypes as C
    from wf_bitnet import ops
    from wf_bitnet._lib import lib
    n, k = nk
    _, packed, _, q, _, _ = make(n, k, m, n * 7 + k + m)
    want = torch.from_numpy(O.gemm_i32_fast(packed, q, 0).astype(np.float32))
    tile = ops.i2s_to_tile(cu(packed))
    qd = cu(q)
    for rep in range(3):
        acc = torch.zeros(m, n, dtype=torch.float32, device="cuda")
        rc = lib().wf_gemm_stream_acc(tile.data_ptr(), n, k, qd.data_ptr(), m, k, acc.data_ptr(), 1, C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, lib().wf_last_error().decode()
        got = acc.cpu()
        bad = (got != want).nonzero()
        assert bad.numel() == 0, (rep, bad[:8].tolist(), got[tuple(bad[0])].item(), want[tuple(bad[0])].item())
