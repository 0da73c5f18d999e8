"""GPU parity: paged KV cache, RoPE append and split-KV flash-decode attention vs the oracle
(oracle/model.py restates kv_cache.py:316-369 and direct_inference_full.py:210-244)."""
import math

import numpy as np
import pytest
import torch

from oracle import model as OM

pytestmark = pytest.mark.gpu

# reference test configuration (tests/test_kv_cache_cpp.py:136-142)
BITNET_CONFIG = {"num_layers": 32, "num_heads": 20, "head_dim": 128, "page_size": 256, "max_pages": 100}


class TestKVCacheManagerVsReference:
    """Same assertions as TestCppVsPython (tests/test_kv_cache_cpp.py:279-373) against our GPU manager."""

    def test_allocation(self):
        from sgl_kernel.kvcache import KVCacheManager, check_kernel_available
        assert check_kernel_available()
        py, cpp = OM.PagedKVReference(**BITNET_CONFIG), KVCacheManager(**BITNET_CONFIG)
        assert py.num_free_pages() == cpp.num_free_pages()
        a, b = py.allocate_pages(10), cpp.allocate_pages(10)
        assert a.shape == b.shape and torch.equal(a, b) and py.num_free_pages() == cpp.num_free_pages()
        py.free_pages(a)
        cpp.free_pages(b)
        assert py.num_free_pages() == cpp.num_free_pages()
        assert cpp.allocate_pages(101).shape == (0,)

    def test_roundtrip_and_layers(self):
        from sgl_kernel.kvcache import KVCacheManager
        cfg = dict(BITNET_CONFIG, num_layers=4, max_pages=20)
        T = 16
        py, cpp = OM.PagedKVReference(**cfg), KVCacheManager(**cfg)
        pages = cpp.allocate_pages(T)
        py.allocate_pages(T)
        slots = torch.zeros(T, dtype=torch.int32)
        ks, vs = [], []
        for layer in range(cfg["num_layers"]):
            k = torch.randn(T, cfg["num_heads"] * cfg["head_dim"])
            v = torch.randn(T, cfg["num_heads"] * cfg["head_dim"])
            ks.append(k); vs.append(v)
            cpp.scatter_kv(k, v, pages, slots, layer)
            py.scatter_kv(k, v, pages, slots, layer)
        for layer in range(cfg["num_layers"]):
            ko, vo = torch.zeros_like(ks[0]), torch.zeros_like(vs[0])
            cpp.gather_kv(ko, vo, pages, slots, layer)
            pk, pv = torch.zeros_like(ks[0]), torch.zeros_like(vs[0])
            py.gather_kv(pk, pv, pages, slots, layer)
            assert torch.equal(ko, ks[layer]) and torch.equal(vo, vs[layer])
            assert torch.equal(ko, pk) and torch.equal(vo, pv)

    def test_slots_are_honoured(self):
        from sgl_kernel.kvcache import KVCacheManager
        cfg = dict(num_layers=2, num_heads=3, head_dim=128, page_size=16, max_pages=4)
        py, cpp = OM.PagedKVReference(**cfg), KVCacheManager(**cfg)
        T = 40
        pos = torch.arange(T)
        pages, slots = (pos // 16).to(torch.int32), (pos % 16).to(torch.int32)
        k, v = torch.randn(T, 3 * 128), torch.randn(T, 3 * 128)
        cpp.scatter_kv(k, v, pages, slots, 1)
        py.scatter_kv(k, v, pages, slots, 1)
        idx = torch.randperm(T)
        ko, vo = torch.zeros(T, 384, device="cuda"), torch.zeros(T, 384, device="cuda")
        cpp.gather_kv(ko, vo, pages[idx], slots[idx], 1)
        assert torch.equal(ko.cpu(), k[idx]) and torch.equal(vo.cpu(), v[idx])
        assert torch.equal(cpp.kv.pool.cpu().permute(0, 1, 2, 4, 3, 5), py.store)   # [page,L,2,slot,H,D] view agrees


def _ref_attention(q, kpool, block_table, seq_lens, layer, H, Hkv, page_size):
    """oracle attention on the tokens the block table points at. q [B,H,D] bf16 (cpu)."""
    outs = []
    for b in range(q.shape[0]):
        L = int(seq_lens[b])
        toks = torch.arange(L)
        pages = block_table[b, toks // page_size].long()
        slots = toks % page_size
        k = kpool[pages, layer, 0, :, slots]        # [L, Hkv, D]
        v = kpool[pages, layer, 1, :, slots]
        outs.append(OM.attention_gqa_fp32(q[b], k, v))
    return torch.stack(outs)


@pytest.mark.parametrize("B,H,Hkv,ctx,page", [
    (1, 20, 5, 1, 16), (1, 20, 5, 15, 16), (1, 20, 5, 17, 16), (2, 20, 5, 513, 16), (3, 32, 8, 300, 16),
    (2, 64, 8, 1000, 16), (4, 20, 5, 4097, 16), (2, 8, 8, 70, 32), (2, 20, 5, 700, 256), (64, 20, 5, 200, 16),
])
def test_paged_decode_attention(B, H, Hkv, ctx, page):
    from wf_bitnet import ops
    gen = torch.Generator().manual_seed(B * 1000 + ctx)
    D, layers, layer = 128, 2, 1
    blocks = (ctx + page - 1) // page
    max_pages = B * blocks + 3
    kv = ops.PagedKV(max_pages, layers, Hkv, page, D)
    kv.pool.copy_(torch.randn(kv.pool.shape, generator=gen).to(torch.bfloat16))
    perm = torch.randperm(max_pages, generator=gen)[: B * blocks].reshape(B, blocks).to(torch.int32)
    seq_lens = torch.randint(max(1, ctx // 2), ctx + 1, (B,), generator=gen).to(torch.int32)
    seq_lens[0] = ctx
    q = torch.randn(B, H, D, generator=gen).to(torch.bfloat16)
    out = ops.paged_decode_attn(q.cuda(), kv, layer, perm.cuda(), seq_lens.cuda(), out_dtype=torch.float32).cpu()
    want = _ref_attention(q, kv.pool.cpu(), perm, seq_lens, layer, H, Hkv, page)
    # P is rounded to bf16 before P@V (as in every flash kernel): bf16-level tolerance
    assert torch.allclose(out, want, atol=2e-2, rtol=2e-2), (out - want).abs().max()
    assert (out - want).abs().mean() < 2e-3
    outb = ops.paged_decode_attn(q.cuda(), kv, layer, perm.cuda(), seq_lens.cuda(), out_dtype=torch.bfloat16).cpu()
    assert torch.equal(outb, out.to(torch.bfloat16))
    # against torch SDPA on bf16, the op the reference actually calls (direct_inference_full.py:241)
    b = 0
    toks = torch.arange(int(seq_lens[b]))
    k = kv.pool.cpu()[perm[b, toks // page].long(), layer, 0, :, toks % page]
    v = kv.pool.cpu()[perm[b, toks // page].long(), layer, 1, :, toks % page]
    sd = OM.attention_gqa_reference(q[b], k, v).float()
    assert torch.allclose(outb[b].float(), sd, atol=3e-2, rtol=3e-2)


def test_attention_ignores_garbage_beyond_seq_len():
    from wf_bitnet import ops
    kv = ops.PagedKV(8, 1, 5, 16, 128)
    kv.pool.fill_(float("nan"))
    gen = torch.Generator().manual_seed(1)
    good = torch.randn(2, 1, 2, 5, 16, 128, generator=gen).to(torch.bfloat16)
    kv.pool[:2] = good.cuda()
    kv.pool[1, :, :, :, 5:] = float("nan")          # tokens 21.. of page 1 are garbage
    bt = torch.tensor([[0, 1, 7, 7]], dtype=torch.int32).cuda()
    q = torch.randn(1, 20, 128, generator=gen).to(torch.bfloat16)
    out = ops.paged_decode_attn(q.cuda(), kv, 0, bt, torch.tensor([21], dtype=torch.int32).cuda(), out_dtype=torch.float32).cpu()
    assert torch.isfinite(out).all()
    k = torch.cat([good[0, 0, 0].permute(1, 0, 2), good[1, 0, 0].permute(1, 0, 2)[:5]])
    v = torch.cat([good[0, 0, 1].permute(1, 0, 2), good[1, 0, 1].permute(1, 0, 2)[:5]])
    assert torch.allclose(out[0], OM.attention_gqa_fp32(q[0], k, v), atol=2e-2, rtol=2e-2)


def test_rope_kv_append_bit_exact():
    from wf_bitnet import ops
    from wf_bitnet.synth import rope_tables
    gen = torch.Generator().manual_seed(4)
    B, H, Hkv, D, page = 3, 20, 5, 128, 16
    rope = OM.RotaryEmbedding(D, 4096, 500000.0)
    cos, sin = rope_tables(D, 4096, 500000.0)
    assert torch.equal(cos, rope.cos_cache) and torch.equal(sin, rope.sin_cache)
    qkv = (torch.randn(B, (H + 2 * Hkv) * D, generator=gen) * 3).to(torch.bfloat16)
    positions = torch.tensor([0, 37, 4095], dtype=torch.int32)
    kv = ops.PagedKV(B * 256, 2, Hkv, page, D)
    bt = torch.randperm(B * 256, generator=gen).reshape(B, 256).to(torch.int32)
    qrot = ops.rope_kv_append(qkv.cuda(), H, kv, 1, positions.cuda(), bt.cuda(), cos.cuda(), sin.cuda()).cpu()
    pool = kv.pool.cpu()
    for b in range(B):
        pos = int(positions[b])
        q = qkv[b, : H * D].view(H, D)
        k = qkv[b, H * D: (H + Hkv) * D].view(Hkv, D)
        v = qkv[b, (H + Hkv) * D:].view(Hkv, D)
        assert torch.equal(qrot[b], rope.rotate(q, pos))
        pg, sl = int(bt[b, pos // page]), pos % page
        assert torch.equal(pool[pg, 1, 0, :, sl], rope.rotate(k, pos))
        assert torch.equal(pool[pg, 1, 1, :, sl], v)
    assert pool[:, 0].abs().sum() == 0      # other layer untouched


def test_attention_with_kv_cache_mirror_vs_reference_golden(golden):
    """tests/golden/attention.npz `*_d128` was produced by the reference's own attention_with_kv_cache (bf16 cache)."""
    from wrinklefree_inference.kv_cache import KVCache, KVCacheConfig, KVCacheDtype, attention_with_kv_cache
    g = golden["attention"]
    ks, vs, qs, want = (torch.from_numpy(g[f"{n}_d128"]) for n in ("k", "v", "q", "out"))
    S, H, D = ks.shape[0], ks.shape[3], ks.shape[4]
    cache = KVCache(KVCacheConfig(max_seq_len=64, num_layers=2, num_heads=H, head_dim=D, dtype=KVCacheDtype.BF16))
    for pos in range(S):
        out = attention_with_kv_cache(qs[pos].to(torch.bfloat16).cuda(), cache, 0, ks[pos].to(torch.bfloat16).cuda(),
                                      vs[pos].to(torch.bfloat16).cuda(), pos)
        assert out.shape == (1, H, 1, D) and out.dtype == torch.bfloat16
        assert torch.allclose(out.float().cpu(), want[pos], atol=2e-2, rtol=2e-2), pos
    assert cache.current_seq_len == S
    k, v = cache.get(0)
    assert torch.equal(k.cpu(), ks[:, 0, 0].to(torch.bfloat16)) and torch.equal(v.cpu(), vs[:, 0, 0].to(torch.bfloat16))
    with pytest.raises(ValueError):
        cache.update(0, ks[0].to(torch.bfloat16), vs[0].to(torch.bfloat16), 64)
    assert KVCache(KVCacheConfig(max_seq_len=32, num_layers=1, num_heads=2, dtype=KVCacheDtype.FP8_E4M3)).storage_dtype == torch.bfloat16   # every dtype exists: tests/test_gpu_kv_int8.py
    cache.clear()
    assert cache.current_seq_len == 0


def test_lm_head_argmax():
    from wf_bitnet import ops
    gen = torch.Generator().manual_seed(8)
    V, K = 128256, 2560
    E = (0.02 * torch.randn(V, K, generator=gen)).to(torch.bfloat16)
    for B in (1, 5, 16, 17, 33, 64):     # <= 16: mma.sync kernel, above: tcgen05 kind::f16 kernel (lm_head_tc.cu), partial token tiles included
        h = torch.randn(B, K, generator=gen).to(torch.bfloat16)
        nxt, logits = ops.lm_head_argmax(h.cuda(), E.cuda())
        ref = (h.float() @ E.float().t())
        lg = logits.cpu()
        assert torch.allclose(lg, ref, rtol=2e-2, atol=2e-2)
        assert torch.equal(lg, lg.to(torch.bfloat16).float())          # bf16-rounded like F.linear on bf16
        assert torch.equal(nxt.cpu().long(), lg.argmax(dim=1))         # lowest index on ties == torch.argmax on our logits
        top = ref.topk(2, dim=1)
        clear = (top.values[:, 0] - top.values[:, 1]) > 0.05
        assert torch.equal(nxt.cpu().long()[clear], top.indices[:, 0][clear])
    tok = torch.tensor([3, 0, 128255], dtype=torch.int32)
    assert torch.equal(ops.embed(E.cuda(), tok.cuda()).cpu(), E[tok.long()])
