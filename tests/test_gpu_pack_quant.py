"""GPU parity: weight formats and activation quantisation vs the oracle / golden vectors (bit-exact)."""
import numpy as np
import pytest
import torch

import oracle as O
from oracle import model as OM

pytestmark = pytest.mark.gpu


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_pack_matches_golden(golden, tag):
    from wf_bitnet import ops
    g = golden["pack"]
    w = g[f"w_{tag}"].astype(np.float32)
    packed, scale = ops.pack_ternary(cu(w))
    assert np.array_equal(packed.cpu().numpy(), g[f"packed_{tag}"])
    assert scale.item() == g[f"scale_{tag}"][0]
    assert np.array_equal(ops.unpack_i2s(packed).cpu().numpy(), w)


def test_pack_float_and_dtypes(golden):
    from wf_bitnet import ops
    g = golden["pack"]
    packed, scale = ops.pack_ternary(cu(g["w_float"]))
    assert np.array_equal(packed.cpu().numpy(), g["packed_float"]) and scale.item() == g["scale_float"][0]
    w = torch.randint(-1, 2, (48, 384), generator=torch.Generator().manual_seed(5)).float()
    ref = O.pack_ternary(w.numpy())[0]
    for dt in (torch.float32, torch.bfloat16, torch.float16):
        assert np.array_equal(ops.pack_ternary(w.cuda().to(dt))[0].cpu().numpy(), ref)
    z, s = ops.pack_ternary(torch.zeros(4, 128, device="cuda"))
    assert s.item() == 1.0 and torch.all(z == 0x55)
    with pytest.raises(ValueError):
        ops.pack_ternary(torch.zeros(2, 6, device="cuda"))


def test_pack_large_shapes_vs_oracle():
    from wf_bitnet import ops
    rng = np.random.default_rng(12)
    for n, k in [(2560, 2560), (640, 2560), (6912, 2560), (2560, 6912)]:
        w = rng.integers(-1, 2, size=(n, k)).astype(np.float32)
        assert np.array_equal(ops.pack_ternary(cu(w))[0].cpu().numpy(), O.pack_ternary(w)[0])


def test_hf_to_i2s(golden):
    from wf_bitnet import ops
    g = golden["hf"]
    assert np.array_equal(ops.hf_to_i2s(cu(g["hf_packed"])).cpu().numpy(), g["kernel_packed"])
    rng = np.random.default_rng(3)
    hf = rng.integers(0, 3, size=(640, 2560, 4)).astype(np.uint8)
    hfb = (hf[..., 0] | (hf[..., 1] << 2) | (hf[..., 2] << 4) | (hf[..., 3] << 6)).astype(np.uint8)
    assert np.array_equal(ops.hf_to_i2s(cu(hfb)).cpu().numpy(), O.hf_to_i2s(hfb))


@pytest.mark.parametrize("shape", [(16, 128), (17, 256), (640, 2560), (2560, 6912), (100, 384)])
def test_tile_layout_roundtrip_and_definition(shape):
    from wf_bitnet import ops
    n, k = shape
    rng = np.random.default_rng(n + k)
    packed = rng.integers(0, 256, size=(n, k // 4)).astype(np.uint8)
    tile = ops.i2s_to_tile(cu(packed))
    assert tile.numel() == ((n + 15) // 16) * (k // 128) * 512
    assert np.array_equal(ops.tile_to_i2s(tile, n, k).cpu().numpy(), packed)
    # definition check: lane l, word i <- row 16rb + (l>>2) + 8(i&1), bytes 4(l&3) + 16(i>>1) of block kb
    t = tile.cpu().numpy().reshape((n + 15) // 16, k // 128, 32, 4, 4)
    rb, kb, l, i = (n - 1) // 16 // 2, (k // 128) - 1, 13, 2
    row = rb * 16 + (l >> 2) + 8 * (i & 1)
    j = 4 * (l & 3) + 16 * (i >> 1)
    if row < n:
        assert np.array_equal(t[rb, kb, l, i], packed[row, kb * 32 + j: kb * 32 + j + 4])
    if n % 16:
        assert np.all(t[-1, :, 31, 1] == 0x55)     # padded rows carry code 1 (weight 0)


def test_interleave_rows16():
    from wf_bitnet import ops
    rng = np.random.default_rng(9)
    a = rng.integers(0, 256, size=(64, 160)).astype(np.uint8)
    b = rng.integers(0, 256, size=(64, 160)).astype(np.uint8)
    out = ops.interleave_rows16(cu(a), cu(b)).cpu().numpy()
    for blk in range(4):
        assert np.array_equal(out[32 * blk: 32 * blk + 16], a[16 * blk: 16 * blk + 16])
        assert np.array_equal(out[32 * blk + 16: 32 * blk + 32], b[16 * blk: 16 * blk + 16])


# ------------------------------------------------------------------ activation quant
@pytest.mark.parametrize("tag", ["randn3", "vec", "zeros", "tiny", "ties", "bf16ish"])
def test_act_quant_golden(golden, tag):
    from wf_bitnet import ops
    g = golden["actquant"]
    x = g[f"x_{tag}"]
    q, s, rs = ops.quantize_activations(cu(x), per_token=False)
    assert np.array_equal(q.cpu().numpy(), g[f"q_{tag}"])
    assert s.item() == np.float32(float(g[f"s_{tag}"]))
    assert np.array_equal(rs.cpu().numpy()[: max(1, x.reshape(-1, x.shape[-1]).shape[0])],
                          g[f"q_{tag}"].reshape(-1, x.shape[-1]).astype(np.int64).sum(1).astype(np.int32))


@pytest.mark.parametrize("M,K", [(1, 2560), (1, 6912), (4, 2560), (64, 6912), (7, 4096), (3, 14336), (2, 28672), (33, 128)])
def test_act_quant_bit_exact_vs_oracle(M, K):
    from wf_bitnet import ops
    gen = torch.Generator().manual_seed(42 + M + K)
    x = torch.randn(M, K, generator=gen) * 3
    qo, so = O.quantize_act_i8(x.numpy())
    q, s, _ = ops.quantize_activations(x.cuda(), per_token=False)
    assert s.item() == np.float32(so) and np.array_equal(q.cpu().numpy(), qo)
    qr, sr, rr = O.quantize_act_i8_rows(x.numpy())
    q2, s2, r2 = ops.quantize_activations(x.cuda(), per_token=True)
    assert np.array_equal(q2.cpu().numpy(), qr) and np.array_equal(s2.cpu().numpy(), sr) and np.array_equal(r2.cpu().numpy(), rr)
    xb = x.to(torch.bfloat16)
    q3, s3, _ = ops.quantize_activations(xb.cuda(), per_token=True)
    qb, sb, _ = O.quantize_act_i8_rows(xb.float().numpy())
    assert np.array_equal(q3.cpu().numpy(), qb) and np.array_equal(s3.cpu().numpy(), sb)


def test_sgl_kernel_quantize_names():
    from sgl_kernel.quantization import bitnet_quantize_activations, quantize_activations_i8
    x = torch.randn(8, 512, generator=torch.Generator().manual_seed(0))
    q, s = bitnet_quantize_activations(x.cuda())
    qo, so = O.quantize_act_i8(x.numpy())
    assert isinstance(s, float) and s == float(np.float32(so)) and q.dtype == torch.int8
    assert np.array_equal(q.cpu().numpy(), qo)
    q1, s1 = quantize_activations_i8(x[0].cuda())
    assert q1.shape == (512,)


@pytest.mark.parametrize("M,K", [(1, 2560), (5, 2560), (64, 6912), (2, 256)])
def test_rmsnorm_quant_vs_oracle(M, K):
    from wf_bitnet import ops
    gen = torch.Generator().manual_seed(K + M)
    x = (torch.randn(M, K, generator=gen) * 4).to(torch.bfloat16)
    gamma = (1 + 0.1 * torch.randn(K, generator=gen)).to(torch.bfloat16)
    y_ref = OM.rms_norm(x, gamma, 1e-5)
    q, s, rs, y = ops.rmsnorm_quant(x.cuda(), gamma.cuda(), 1e-5, want_y=True)
    y = y.cpu()
    # fp32 sum order differs from torch's: allow a rare 1-ulp bf16 flip of the rsqrt factor, nothing more
    mism = (y != y_ref).float().mean().item()
    assert mism < 0.02, mism
    assert torch.allclose(y.float(), y_ref.float(), rtol=2e-2, atol=1e-6)
    # the quantiser applied to OUR normalised row must be bit-exact
    qo, so, ro = O.quantize_act_i8_rows(y.float().numpy())
    assert np.array_equal(q.cpu().numpy(), qo) and np.array_equal(s.cpu().numpy(), so) and np.array_equal(rs.cpu().numpy(), ro)
