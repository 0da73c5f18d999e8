"""int8 KV option (the reference's KVCacheDtype.INT8, kv_cache/kv_cache.py:198-228) on the GPU path:
codes / scales bit-exact against the oracle's restatement of the reference arithmetic (per (token, head) row), attention
against the oracle attention on the DEQUANTISED cache (what KVCache.get hands to SDPA), engine logits against the oracle
model with the same cache, fused phase path == per-op path."""
import numpy as np
import pytest
import torch

from oracle import model as OM

pytestmark = pytest.mark.gpu


def _pages_slots(T, page):
    pos = torch.arange(T)
    return (pos // page).to(torch.int32), (pos % page).to(torch.int32)


@pytest.mark.parametrize("Hkv,page", [(5, 16), (8, 16), (3, 256)])
def test_scatter_quantises_like_the_reference_and_gather_dequantises(Hkv, page):
    from wf_bitnet import ops
    gen = torch.Generator().manual_seed(Hkv)
    T, D, layers = 70, 128, 3
    kv = ops.PagedKV((T + page - 1) // page + 1, layers, Hkv, page, D, dtype=torch.int8)
    k = (torch.randn(T, Hkv * D, generator=gen) * 2).to(torch.bfloat16)
    v = (torch.randn(T, Hkv * D, generator=gen) * 0.3).to(torch.bfloat16)
    k[3] = 0                                   # all-zero row: absmax < 1e-6 -> scale 1/127 (kv_cache.py:201-203)
    v[5, :D] = 1e-8
    pages, slots = _pages_slots(T, page)
    ops.kv_scatter(kv, k.cuda(), v.cuda(), pages.cuda(), slots.cuda(), 1)
    kc, ks = OM.kv_int8_quant(k.view(T, Hkv, D))
    vc, vs = OM.kv_int8_quant(v.view(T, Hkv, D))
    pool, scales = kv.codes().cpu(), kv.scales.cpu()
    assert torch.equal(pool[pages.long(), 1, 0, :, slots.long()], kc)            # int8 codes: bit-exact
    assert torch.equal(pool[pages.long(), 1, 1, :, slots.long()], vc)
    assert torch.equal(scales[pages.long(), 1, 0, :, slots.long()], ks[..., 0])  # fp32 scales: bit-exact
    assert torch.equal(scales[pages.long(), 1, 1, :, slots.long()], vs[..., 0])
    assert kv.pool[:, 0].sum() == 0 and kv.pool[:, 2].sum() == 0                # other layers untouched
    idx = torch.randperm(T, generator=gen)
    ko = torch.zeros(T, Hkv * D, dtype=torch.bfloat16, device="cuda")
    vo = torch.zeros_like(ko)
    ops.kv_gather(kv, ko, vo, pages[idx].cuda(), slots[idx].cuda(), 1)
    assert torch.equal(ko.cpu().view(T, Hkv, D), OM.kv_int8_roundtrip(k.view(T, Hkv, D))[idx])
    assert torch.equal(vo.cpu().view(T, Hkv, D), OM.kv_int8_roundtrip(v.view(T, Hkv, D))[idx])


def _fill_int8(kv, gen, k_std=1.0, v_std=1.0):
    """random bf16 K/V -> quantised into the pool through the oracle's arithmetic; returns the dequantised bf16 pool."""
    shp = kv.pool.shape                                      # [page][layer][2][head][slot][D]
    x = torch.randn(shp, generator=gen)
    x[:, :, 0] *= k_std
    x[:, :, 1] *= v_std
    x = x * (0.25 + 2 * torch.rand(shp[:-1], generator=gen)).unsqueeze(-1)      # per-token magnitudes differ
    codes, sc = OM.kv_int8_quant(x.to(torch.bfloat16))
    kv.set_codes(codes)
    kv.scales.copy_(sc[..., 0])
    return (codes.float() * sc).to(torch.bfloat16)


@pytest.mark.parametrize("B,H,Hkv,ctx,page", [
    (1, 20, 5, 1, 16), (1, 20, 5, 15, 16), (1, 20, 5, 17, 16), (2, 20, 5, 513, 16), (3, 32, 8, 300, 16),
    (2, 64, 8, 1000, 16), (4, 20, 5, 4097, 16), (2, 8, 8, 70, 32), (2, 20, 5, 700, 256), (64, 20, 5, 200, 16),
])
def test_paged_decode_attention_int8_kv(B, H, Hkv, ctx, page):
    from test_gpu_attention import _ref_attention
    from wf_bitnet import ops
    gen = torch.Generator().manual_seed(B * 1000 + ctx)
    D, layers, layer = 128, 2, 1
    blocks = (ctx + page - 1) // page
    max_pages = B * blocks + 3
    kv = ops.PagedKV(max_pages, layers, Hkv, page, D, dtype=torch.int8)
    deq = _fill_int8(kv, gen)
    perm = torch.randperm(max_pages, generator=gen)[: B * blocks].reshape(B, blocks).to(torch.int32)
    seq_lens = torch.randint(max(1, ctx // 2), ctx + 1, (B,), generator=gen).to(torch.int32)
    seq_lens[0] = ctx
    q = torch.randn(B, H, D, generator=gen).to(torch.bfloat16)
    out = ops.paged_decode_attn(q.cuda(), kv, layer, perm.cuda(), seq_lens.cuda(), out_dtype=torch.float32).cpu()
    want = _ref_attention(q, deq, perm, seq_lens, layer, H, Hkv, page)
    # same bar as the bf16 cache: P is rounded to 16 bits before P@V, everything else fp32
    assert torch.allclose(out, want, atol=2e-2, rtol=2e-2), (out - want).abs().max()
    assert (out - want).abs().mean() < 2e-3
    cos = torch.nn.functional.cosine_similarity(out.flatten(), want.flatten(), dim=0).item()
    assert cos > 0.9995, cos


def test_int8_attention_ignores_garbage_beyond_seq_len():
    from test_gpu_attention import _ref_attention
    from wf_bitnet import ops
    kv = ops.PagedKV(8, 1, 5, 16, 128, dtype=torch.int8)
    gen = torch.Generator().manual_seed(1)
    deq = _fill_int8(kv, gen)
    kv.scales[1, :, :, :, 5:] = float("nan")          # tokens 21.. of page 1: uninitialised scales
    kv.scales[7] = float("inf")
    bt = torch.tensor([[0, 1, 7, 7]], dtype=torch.int32)
    q = torch.randn(1, 20, 128, generator=gen).to(torch.bfloat16)
    sl = torch.tensor([21], dtype=torch.int32)
    out = ops.paged_decode_attn(q.cuda(), kv, 0, bt.cuda(), sl.cuda(), out_dtype=torch.float32).cpu()
    assert torch.isfinite(out).all()
    assert torch.allclose(out, _ref_attention(q, deq, bt, sl, 0, 20, 5, 16), atol=2e-2, rtol=2e-2)


def test_rope_kv_append_int8_codes_bit_exact():
    from wf_bitnet import ops
    from wf_bitnet.synth import rope_tables
    gen = torch.Generator().manual_seed(4)
    B, H, Hkv, D, page = 3, 20, 5, 128, 16
    rope = OM.RotaryEmbedding(D, 4096, 500000.0)
    cos, sin = rope_tables(D, 4096, 500000.0)
    qkv = (torch.randn(B, (H + 2 * Hkv) * D, generator=gen) * 3).to(torch.bfloat16)
    positions = torch.tensor([0, 37, 4095], dtype=torch.int32)
    kv = ops.PagedKV(B * 256, 2, Hkv, page, D, dtype=torch.int8)
    bt = torch.randperm(B * 256, generator=gen).reshape(B, 256).to(torch.int32)
    qrot = ops.rope_kv_append(qkv.cuda(), H, kv, 1, positions.cuda(), bt.cuda(), cos.cuda(), sin.cuda()).cpu()
    pool, scales = kv.codes().cpu(), kv.scales.cpu()
    for b in range(B):
        pos = int(positions[b])
        q = qkv[b, : H * D].view(H, D)
        k = qkv[b, H * D: (H + Hkv) * D].view(Hkv, D)
        v = qkv[b, (H + Hkv) * D:].view(Hkv, D)
        assert torch.equal(qrot[b], rope.rotate(q, pos))
        pg, sl = int(bt[b, pos // page]), pos % page
        kc, ks = OM.kv_int8_quant(rope.rotate(k, pos))
        vc, vs = OM.kv_int8_quant(v)
        assert torch.equal(pool[pg, 1, 0, :, sl], kc) and torch.equal(scales[pg, 1, 0, :, sl], ks[:, 0])
        assert torch.equal(pool[pg, 1, 1, :, sl], vc) and torch.equal(scales[pg, 1, 1, :, sl], vs[:, 0])
    assert kv.pool[:, 0].sum() == 0


# ------------------------------------------------------------------------------------------------ engine
def _cfg(**kw):
    from wf_bitnet.engine import EngineConfig
    base = dict(hidden=256, layers=2, heads=4, kv_heads=2, head_dim=128, ffn=384, vocab=512, max_seq=96, page_size=16, max_batch=1,
                kv_dtype="int8")
    base.update(kw)
    return EngineConfig(**base)


def test_engine_int8_kv_cache_and_attention_match_oracle_model():
    """One-layer model, so every input of the attention phase is bit-identical to the oracle's: the codes / scales the engine
    appends must equal the oracle's quantisation of its own k / v exactly, and the attention output must match the oracle's
    fp32 attention over the cache contents (codes * scales).  The kernel uses the exact product code * scale; the reference
    rounds the dequantised copy to bf16 first (kv_cache.py:226-228) - on this synthetic model, whose scores are in the
    thousands, that 2^-9 rounding alone reorders the softmax, so the comparison is made on the unrounded cache.  (Logits of deeper models are not compared position by position: the int8 activation
    re-quantisation downstream amplifies ANY sub-ulp change of the attention output - the oracle against itself with the KV
    scales rounded to bf16 already moves the logits of this toy model by 2-50 % - so they say nothing about this path.)"""
    from test_gpu_engine import oracle_cfg
    from wf_bitnet.engine import Engine
    from wf_bitnet.synth import make_weights
    cfg = _cfg(layers=1)
    w = make_weights(cfg, seed=3)
    eng = Engine(cfg)
    eng.load_weights(w)
    om = OM.OracleModel(oracle_cfg(cfg), w, use_ref=True, kv_int8=True)
    toks = [7, 300, 12, 99, 5, 410, 33, 2, 77, 150, 9, 256, 31, 8, 64, 200, 1, 500, 45, 6]    # crosses a page boundary
    kq, vq = [], []
    pages = cfg.max_seq // cfg.page_size
    elems = pages * cfg.layers * 2 * cfg.kv_heads * cfg.page_size * cfg.head_dim
    for pos, t in enumerate(toks):
        eng.decode([t])
        om.forward(t, pos)
        raw = eng.buffer("pool", (elems + elems // cfg.head_dim * 4,), torch.uint8).cpu()
        codes = (raw[:elems].view(pages, 1, 2, cfg.kv_heads, cfg.page_size, cfg.head_dim).to(torch.int16) - 128).to(torch.int8)
        scales = raw[elems:].view(torch.float32).view(pages, 1, 2, cfg.kv_heads, cfg.page_size)
        bt = eng.buffer("block_table", (1, pages), torch.int32).cpu()
        pg, sl = int(bt[0, pos // cfg.page_size]), pos % cfg.page_size
        for kv, name in ((0, "k"), (1, "v")):
            c, sc = OM.kv_int8_quant(om.taps[(name, 0)])
            assert torch.equal(codes[pg, 0, kv, :, sl], c), (pos, name)
            assert torch.equal(scales[pg, 0, kv, :, sl], sc[:, 0]), (pos, name)
            (kq, vq)[kv].append(c.float() * sc)
        got = eng.buffer("attn", (1, cfg.heads * cfg.head_dim), torch.bfloat16)[0].float().cpu()
        want = OM.attention_gqa_fp32(om.taps[("q", 0)], torch.stack(kq), torch.stack(vq)).flatten().float()
        assert torch.allclose(got, want, atol=2e-2 * want.abs().max().item(), rtol=2e-2), (pos, (got - want).abs().max())
        cos = torch.nn.functional.cosine_similarity(got, want, dim=0).item()
        assert cos > 0.9995, (pos, cos)


def test_engine_int8_kv_greedy_decode_is_close_to_bf16_kv():
    """End to end the int8 cache is a small perturbation: on the synthetic model most greedy tokens of the first steps agree
    with the bf16-cache engine and the logits stay strongly correlated."""
    from wf_bitnet.engine import Engine
    from wf_bitnet.synth import make_weights
    cfgs = [_cfg(kv_dtype="bf16"), _cfg()]
    w = make_weights(cfgs[0], seed=3)
    toks = [7, 300, 12, 99, 5, 410, 33, 2, 77, 150, 9, 256]
    outs = []
    for cfg in cfgs:
        eng = Engine(cfg)
        eng.load_weights(w)
        o = []
        for t in toks:
            eng.decode([t])
            o.append(eng.buffer("logits", (1, cfg.vocab), torch.float32)[0].cpu())
        outs.append(o)
        eng.close()
    cos = [torch.nn.functional.cosine_similarity(a, b, dim=0).item() for a, b in zip(*outs)]
    assert np.median(cos) > 0.99, cos
    assert min(cos) > 0.8, cos


@pytest.mark.parametrize("B", [1, 3, 16, 40])
def test_engine_int8_kv_phase_kernels_equal_per_op_kernels(B, monkeypatch):
    from wf_bitnet.engine import Engine
    from wf_bitnet.synth import make_weights
    cfg = _cfg(max_batch=B)
    w = make_weights(cfg, seed=9)
    gen = np.random.default_rng(B)
    steps = [gen.integers(0, cfg.vocab, B).tolist() for _ in range(20)]
    logits = []
    for fused in ("1", "0"):
        monkeypatch.setenv("WF_FUSED", fused)
        eng = Engine(cfg)
        eng.load_weights(w)
        out = []
        for t in steps:
            nxt = eng.decode(t)
            out.append((nxt.copy(), eng.buffer("logits", (B, cfg.vocab), torch.float32).cpu()))
        logits.append(out)
        eng.close()
    for (n1, l1), (n0, l0) in zip(*logits):
        assert np.array_equal(n1, n0)
        assert torch.equal(l1, l0)


def test_int8_pool_is_half_the_bytes():
    from wf_bitnet._lib import WF_BF16, WF_I8, lib
    b16 = lib().wf_kv_pool_bytes(100, 30, 5, 16, 128, WF_BF16)
    i8 = lib().wf_kv_pool_bytes(100, 30, 5, 16, 128, WF_I8)
    assert i8 == b16 // 2 + b16 // 2 // 128 * 4


def test_kvcache_mirror_int8_update_get_attention(golden):
    """wrinklefree_inference.kv_cache with KVCacheDtype.INT8 (the reference's default dtype, kv_cache.py:56): update / get
    reproduce the reference's row arithmetic (golden rows from its own KVCache), attention_with_kv_cache equals the oracle
    attention over what get() returns."""
    from wrinklefree_inference.kv_cache import KVCache, KVCacheConfig, KVCacheDtype, attention_with_kv_cache, compute_kv_cache_memory
    g = golden["kv_int8"]
    x = torch.from_numpy(g["x"])                                     # [96, 128] rows
    H, D, S, R0 = 4, 128, 8, 32                                    # rows 32..63: magnitudes 0.1 .. 10
    cfg = KVCacheConfig(max_seq_len=32, num_layers=2, num_heads=H, head_dim=D, dtype=KVCacheDtype.INT8)
    assert cfg.bytes_per_token_per_layer == 2 * H * D
    cache = KVCache(cfg)
    rows = x[R0: R0 + S * H].view(S, H, D)
    for pos in range(S):
        cache.update(1, rows[pos].view(1, 1, H, D), rows[pos].flip(0).contiguous().view(1, 1, H, D), pos)
    k, v = cache.get(1)
    want = torch.from_numpy(g["deq"])[R0: R0 + S * H].view(S, H, D)
    assert torch.equal(k.float().cpu(), want)
    assert torch.equal(v.float().cpu(), want.flip(1))
    assert cache.k_scales[1].item() == np.float32(g["scales"][R0 + S * H - 1])
    # attention at the next position over the int8 cache
    gen = torch.Generator().manual_seed(5)
    q = torch.randn(1, H, 1, D, generator=gen)
    nk, nv = torch.randn(1, 1, H, D, generator=gen), torch.randn(1, 1, H, D, generator=gen)
    out = attention_with_kv_cache(q, cache, 1, nk, nv, S)
    k2, v2 = cache.get(1)
    ref = OM.attention_gqa_fp32(q.view(H, D).to(torch.bfloat16), k2.cpu(), v2.cpu())
    assert out.shape == (1, H, 1, D)
    assert torch.allclose(out.view(H, D).cpu(), ref, atol=2e-2 * ref.abs().max().item(), rtol=2e-2)
    m8 = compute_kv_cache_memory({"num_layers": 30, "num_heads": 5, "head_dim": 128}, 4096, KVCacheDtype.INT8)
    m16 = compute_kv_cache_memory({"num_layers": 30, "num_heads": 5, "head_dim": 128}, 4096, KVCacheDtype.BF16)
    assert m8["total_mb"] * 2 == m16["total_mb"]
    cache.clear()
    assert cache.current_seq_len == 0


@pytest.mark.parametrize("mode", ["FP8_E4M3", "FP8_E5M2"])
def test_kvcache_mirror_fp8_modes_match_the_reference_class(golden, mode):
    """The reference's two "FP8" cache modes are int8 storage with an absmax / 127 scale (kv_cache.py:150-196).
    tests/golden/kv_modes.npz holds what the reference's own KVCache stores and returns in those modes, row by row:
    the mirror's update / get must reproduce the dequantised rows and the stored scale bit for bit, and the pool the
    codes."""
    from wrinklefree_inference.kv_cache import KVCache, KVCacheConfig, KVCacheDtype, attention_with_kv_cache
    g = golden["kv_modes"]
    x = torch.from_numpy(g["x"])
    H, D, S, R0 = 4, 128, 8, 32
    cfg = KVCacheConfig(max_seq_len=32, num_layers=2, num_heads=H, head_dim=D, dtype=getattr(KVCacheDtype, mode))
    assert cfg.bytes_per_token_per_layer == 2 * H * D
    cache = KVCache(cfg)
    rows = x[R0: R0 + S * H].view(S, H, D)
    for pos in range(S):
        cache.update(1, rows[pos].view(1, 1, H, D), rows[pos].flip(0).contiguous().view(1, 1, H, D), pos)
    k, v = cache.get(1)
    want = torch.from_numpy(g[f"{mode}_deq"])[R0: R0 + S * H].view(S, H, D)
    assert torch.equal(k.float().cpu(), want)
    assert torch.equal(v.float().cpu(), want.flip(1))
    assert cache.k_scales[1].item() == np.float32(g[f"{mode}_scales"][R0 + S * H - 1])
    codes = cache.kv.codes().cpu()[0, 1, 0].permute(1, 0, 2)[:S]          # page 0, layer 1, K: [slot][head][D]
    assert torch.equal(codes, torch.from_numpy(g[f"{mode}_codes"])[R0: R0 + S * H].view(S, H, D))
    gen = torch.Generator().manual_seed(6)
    q = torch.randn(1, H, 1, D, generator=gen)
    nk, nv = torch.randn(1, 1, H, D, generator=gen), torch.randn(1, 1, H, D, generator=gen)
    out = attention_with_kv_cache(q, cache, 1, nk, nv, S)
    k2, v2 = cache.get(1)
    ref = OM.attention_gqa_fp32(q.view(H, D).to(torch.bfloat16), k2.cpu(), v2.cpu())
    assert torch.allclose(out.view(H, D).cpu(), ref, atol=2e-2 * ref.abs().max().item(), rtol=2e-2)


def test_kvcache_mirror_fp16_and_fp32_caches(golden):
    """FP16 / FP32 caches: update / get are exact (the golden rows are what the reference's FP16 cache returns), and
    attention_with_kv_cache over an FP16 cache matches the output of the reference's own attention_with_kv_cache
    (tests/golden/kv_modes.npz attn_*) within bf16 tolerance."""
    from wrinklefree_inference.kv_cache import KVCache, KVCacheConfig, KVCacheDtype, attention_with_kv_cache
    g = golden["kv_modes"]
    x = torch.from_numpy(g["x"])
    R, D = x.shape
    cache = KVCache(KVCacheConfig(max_seq_len=R, num_layers=1, num_heads=1, head_dim=D, dtype=KVCacheDtype.FP16))
    for r in range(R):
        cache.update(0, x[r].view(1, 1, 1, D), (x[r] * 0.5).view(1, 1, 1, D), r)
    k, v = cache.get(0)
    assert k.dtype == torch.float16
    assert np.array_equal(k.float().cpu().reshape(R, D).numpy(), g["FP16_k"])
    assert np.array_equal(v.float().cpu().reshape(R, D).numpy(), g["FP16_v"])
    ks, vs, q = torch.from_numpy(g["attn_k"]), torch.from_numpy(g["attn_v"]), torch.from_numpy(g["attn_q"]).to(torch.bfloat16)
    S, H = ks.shape[0], ks.shape[1]
    for dt in (KVCacheDtype.FP16, KVCacheDtype.FP32):
        cache = KVCache(KVCacheConfig(max_seq_len=32, num_layers=1, num_heads=H, head_dim=D, dtype=dt))
        for pos in range(S - 1):
            cache.update(0, ks[pos].view(1, 1, H, D), vs[pos].view(1, 1, H, D), pos)
        out = attention_with_kv_cache(q, cache, 0, ks[S - 1].view(1, 1, H, D), vs[S - 1].view(1, 1, H, D), S - 1)
        want = torch.from_numpy(g["attn_out"])
        assert out.shape == want.shape
        assert torch.allclose(out.float().cpu(), want, atol=2e-2, rtol=2e-2), (dt, (out.float().cpu() - want).abs().max())


def test_kvcache_config_default_dtype_is_the_references():
    from wrinklefree_inference.kv_cache import KVCacheConfig, KVCacheDtype
    assert KVCacheConfig().dtype == KVCacheDtype.INT8          # kv_cache.py:56-57
