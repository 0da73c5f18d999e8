"""Pins the oracle (oracle/wf_oracle.c + oracle/model.py) against
 (a) golden vectors produced by importing the reference's own Python (tests/golden/make_golden.py),
 (b) the unmodified reference SIMD kernel built into oracle/_ref (when present).
CPU only."""
import math

import numpy as np
import pytest
import torch

import oracle as O
from oracle import model as OM


# ------------------------------------------------------------------ packing
@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_pack_matches_reference(golden, tag):
    g = golden["pack"]
    w = g[f"w_{tag}"].astype(np.float32)
    packed, scale = O.pack_ternary(w)
    assert np.array_equal(packed, g[f"packed_{tag}"])
    assert scale == g[f"scale_{tag}"][0]
    assert np.array_equal(O.unpack_i2s(packed), w)
    if w.shape[1] % 128 == 0:
        p2, s2 = O.quantize_i2_s(w)          # ggml restatement agrees on ternary input
        assert np.array_equal(p2, packed) and s2 == 1.0


def test_pack_float_rounding(golden):
    g = golden["pack"]
    packed, scale = O.pack_ternary(g["w_float"])
    assert np.array_equal(packed, g["packed_float"])
    assert scale == g["scale_float"][0]


def test_pack_rejects_bad_k():
    with pytest.raises(ValueError):
        O.pack_ternary(np.zeros((2, 6), np.float32))


@pytest.mark.parametrize("byte", [0x00, 0x55, 0xAA, 0x1B, 0xE4])
def test_layout_kats(golden, byte):
    g = golden["pack"]
    pk = np.full((2, 32), byte, np.uint8)
    assert np.array_equal(O.unpack_i2s(pk).astype(np.int8), g[f"kat_{byte:02x}"])


def test_hf_layout(golden):
    g = golden["hf"]
    un = O.unpack_hf(g["hf_packed"])
    assert np.array_equal(un.astype(np.int8), g["unpacked"])
    assert np.array_equal(O.hf_to_i2s(g["hf_packed"]), g["kernel_packed"])
    assert np.array_equal(O.pack_ternary(un)[0], g["kernel_packed"])


def test_native_lsb_layout(golden):
    g = golden["native"]
    assert np.array_equal(O.native_pack(g["w"].astype(np.float32)), g["packed"])


# ------------------------------------------------------------------ act quant
@pytest.mark.parametrize("tag", ["randn3", "vec", "zeros", "tiny", "ties", "bf16ish"])
def test_act_quant_matches_reference(golden, tag):
    g = golden["actquant"]
    q, s = O.quantize_act_i8(g[f"x_{tag}"])
    assert np.array_equal(q, g[f"q_{tag}"])
    assert s == float(g[f"s_{tag}"])


def test_act_quant_matches_torch_cpu_bulk():
    """Same formula straight from torch CPU (reference: tests/test_bitnet_core.py:139-146) on many rows."""
    gen = torch.Generator().manual_seed(42)
    for _ in range(20):
        x = torch.randn(4, 2560, generator=gen) * 3
        mv = x.abs().max().item()
        sc = mv / 127.0
        qt = (x / sc).round().clamp(-128, 127).to(torch.int8).numpy()
        q, s = O.quantize_act_i8(x.numpy())
        assert s == sc
        assert np.array_equal(q, qt)


def test_act_quant_rows():
    gen = torch.Generator().manual_seed(1)
    x = (torch.randn(6, 640, generator=gen) * 5).numpy()
    q, s, rs = O.quantize_act_i8_rows(x)
    for m in range(6):
        qm, sm = O.quantize_act_i8(x[m])
        assert np.array_equal(q[m], qm) and s[m] == np.float32(sm)
        assert rs[m] == int(qm.astype(np.int64).sum())


# ------------------------------------------------------------------ gemv / gemm
@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_gemv_matches_reference(golden, tag):
    g = golden["gemv"]
    y = O.gemv(g[f"packed_{tag}"], g[f"xq_{tag}"], 1.7)
    # reference computes matmul in fp32 then * scale; integer sums < 2^24 so both are exact products
    assert np.array_equal(y, g[f"y_{tag}"])
    acc = O.gemm_i32(g[f"packed_{tag}"], g[f"xq_{tag}"], mode=0)[0]
    assert np.array_equal((acc.astype(np.float32) * np.float32(1.7)), g[f"y_{tag}"])
    # sum w*a mode == python-fallback (LSB layout) semantics on the same ternary matrix
    acc1 = O.gemm_i32(g[f"packed_{tag}"], g[f"xq_{tag}"], mode=1)[0]
    assert np.allclose(acc1.astype(np.float32) * np.float32(1.7), g[f"lsb_y_{tag}"], rtol=1e-6, atol=1e-3)
    # native.py dot on the LSB-packed matrix gives the same integers
    lsb = g[f"lsb_packed_{tag}"]
    nat = np.array([O.native_dot(lsb[n], g[f"xq_{tag}"]) for n in range(lsb.shape[0])])
    assert np.array_equal(nat, acc1)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_linear_correction(golden, tag):
    g = golden["gemv"]
    out = O.bitnet_linear(g[f"x_{tag}"], g[f"packed_{tag}"], 1.7)
    ref = g[f"corrected_{tag}"]
    # fp32 evaluation order differs by at most an ulp or two of the uncorrected magnitude
    assert np.allclose(out, ref, rtol=1e-5, atol=1e-4 * np.abs(ref).max())


def test_all_ones_kat(golden):
    g = golden["gemv"]
    x = np.ones(256, np.int8)
    y = O.gemv(g["ones_packed"], x, 2.0)
    assert np.array_equal(y, g["ones_y"])
    assert np.all(y == 2.0 * 2 * 256)       # scale * sum (w+1)*a = 2 * 2K


def test_gemm_matches_reference(golden):
    g = golden["gemv"]
    q, s = O.quantize_act_i8(g["gemm_X"])
    assert np.array_equal(q, g["gemm_Xq"]) and s == float(g["gemm_sx"])
    Y = O.gemm(g["gemm_packed"], q, 0.75)
    assert np.array_equal(Y, g["gemm_Y"])


def test_gemv_errors():
    with pytest.raises(ValueError):
        O.gemv(np.zeros((4, 16), np.uint8), np.zeros(64, np.int8), 1.0)       # K=64 not multiple of 128
    with pytest.raises(ValueError):
        O.gemv(np.zeros((4, 32), np.uint8), np.zeros(64, np.int8), 1.0)       # length mismatch


# ------------------------------------------------------------------ oracle vs the real reference kernel (_ref)
needs_ref = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (no /root/reference here)")


@needs_ref
@pytest.mark.parametrize("shape", [(64, 128), (256, 512), (640, 2560), (2560, 2560), (320, 6912)])
def test_ref_kernel_agrees_with_oracle(shape):
    n, k = shape
    rng = np.random.default_rng(n * 7 + k)
    w = rng.integers(-1, 2, size=(n, k)).astype(np.float32)
    packed, _ = O.pack_ternary(w)
    rp, rs = O.ref_quantize_i2_s(w)
    assert np.array_equal(rp, packed) and rs == 1.0          # quantize_i2_s == _pack_ternary_weights bytes
    x = (rng.standard_normal(k) * 3).astype(np.float32)
    q, _ = O.quantize_act_i8(x)
    y_ref = O.ref_gemv(packed, q, 1.0)
    acc = O.gemm_i32(packed, q, 0)[0]
    assert np.array_equal(y_ref.astype(np.int64), acc.astype(np.int64))
    Y_ref = O.ref_gemm(packed, np.stack([q, -q]), 0.5)
    assert np.array_equal(Y_ref, O.gemm(packed, np.stack([q, -q]), 0.5))


@needs_ref
def test_ref_int16_wrap_is_modelled_not_reproduced():
    """the AVX2 kernel wraps on adversarial data; the oracle of record is exact."""
    k = 4096
    w = np.ones((1, k), np.float32)
    packed, _ = O.pack_ternary(w)
    a = np.full(k, 64, np.int8)
    exact = O.vec_dot(packed[0], a)
    assert exact == 2 * 64 * k
    wrapped = int(O.ref_gemv(packed, a, 1.0)[0])
    assert wrapped != exact
    assert O.vec_dot_avx2model(packed[0], a) == wrapped       # our scalar model of the SIMD data path
    a2 = np.full(k, 32, np.int8)
    assert int(O.ref_gemv(packed, a2, 1.0)[0]) == O.vec_dot(packed[0], a2)


# ------------------------------------------------------------------ attention
@pytest.mark.parametrize("tag,tol", [("fp32", 2e-6), ("bf16", 0.0)])
def test_attention_mha_matches_reference(golden, tag, tol):
    g = golden["attention"]
    dt = torch.float32 if tag == "fp32" else torch.bfloat16
    ks = torch.from_numpy(g[f"k_{tag}"]).to(dt)
    vs = torch.from_numpy(g[f"v_{tag}"]).to(dt)
    qs = torch.from_numpy(g[f"q_{tag}"]).to(dt)
    S = ks.shape[0]
    for pos in range(S):
        keys = ks[: pos + 1, 0, 0].unsqueeze(0)       # [1,S,H,D]
        vals = vs[: pos + 1, 0, 0].unsqueeze(0)
        out = OM.attention_mha_reference(qs[pos], keys, vals).float().numpy()
        assert np.allclose(out, g[f"out_{tag}"][pos], atol=tol, rtol=0)


def test_attention_gqa_forms_agree():
    gen = torch.Generator().manual_seed(3)
    q = torch.randn(8, 64, generator=gen).to(torch.bfloat16)
    k = torch.randn(37, 2, 64, generator=gen).to(torch.bfloat16)
    v = torch.randn(37, 2, 64, generator=gen).to(torch.bfloat16)
    a = OM.attention_gqa_reference(q, k, v).float()
    b = OM.attention_gqa_fp32(q, k, v)
    assert torch.allclose(a, b, atol=2e-2, rtol=2e-2)


def test_kv_memory_formula(golden):
    g = golden["attention"]
    per_tok = 2 * 5 * 128 * 2 * 30 / 1024
    assert math.isclose(float(g["mem_per_token_kb"]), per_tok)
    assert math.isclose(float(g["mem_total_mb"]), 4096 * 30 * 2560 / 2**20)


# ------------------------------------------------------------------ paged KV reference
def test_paged_reference_semantics():
    c = OM.PagedKVReference(num_layers=3, num_heads=2, head_dim=8, page_size=4, max_pages=10)
    assert c.num_free_pages() == 10
    p = c.allocate_pages(4)
    assert p.tolist() == [0, 1, 2, 3] and c.num_free_pages() == 6
    assert c.allocate_pages(7).numel() == 0 and c.num_free_pages() == 6
    k = torch.randn(4, 16)
    v = torch.randn(4, 16)
    slots = torch.tensor([0, 1, 2, 3], dtype=torch.int32)
    c.scatter_kv(k, v, p, slots, 1)
    ko, vo = torch.zeros(4, 16), torch.zeros(4, 16)
    c.gather_kv(ko, vo, p, slots, 1)
    assert torch.equal(ko, k) and torch.equal(vo, v)
    c.free_pages(p)
    assert c.num_free_pages() == 10


def test_kv_int8_quantiser_matches_reference_kvcache(golden):
    """tests/golden/kv_int8.npz: the reference's KVCache(INT8).update / .get fed one (token, head) row per call
    (kv_cache.py:198-228).  The oracle restatement the GPU tests check against must give the same codes, the same stored
    fp32 scale and the same dequantised bf16 row."""
    import torch
    from oracle import model as OM
    g = golden["kv_int8"]
    x = torch.from_numpy(g["x"])
    codes, sc = OM.kv_int8_quant(x)
    assert np.array_equal(codes.numpy(), g["codes"])
    assert np.array_equal(sc[:, 0].numpy(), g["scales"].astype(np.float32))
    assert np.array_equal(OM.kv_int8_roundtrip(x).float().numpy(), g["deq"])
    assert g["scales"][0] == np.float32(1.0 / 127.0) and np.all(g["codes"][0] == 0)      # all-zero row: absmax -> 1
