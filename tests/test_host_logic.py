"""Host-side logic (no GPU): synthetic-weight packer, LSB<->I2_S re-blocking, page allocator,
KV config arithmetic, sgl_kernel argument errors."""
import numpy as np
import pytest
import torch

import oracle as O
from oracle import model as OM


def test_synth_packer_equals_oracle_pack():
    from wf_bitnet.synth import pack_i2s_numpy
    rng = np.random.default_rng(0)
    for n, k in [(16, 128), (40, 640), (64, 2560)]:
        w = rng.integers(-1, 2, size=(n, k)).astype(np.float32)
        assert np.array_equal(pack_i2s_numpy(w), O.pack_ternary(w)[0])


def test_native_numpy_helpers_match_reference_golden(golden):
    from wrinklefree_inference.kernels import native
    g = golden["native"]
    w = g["w"].astype(np.float32)
    assert np.array_equal(native.pack_weights(w), g["packed"])
    assert np.array_equal(native.unpack_weights(g["packed"]).astype(np.int8), g["unpacked"])
    h = golden["hf"]
    assert np.array_equal(native.repack_hf_weights(h["hf_packed"]), h["native_repack"])
    a = golden["actquant"]
    q, s = native.quantize_activations(a["x_vec"])
    assert np.array_equal(q, a["native_q_vec"]) and s == float(a["native_s_vec"])


def test_lsb_to_i2s_is_the_same_ternary_matrix():
    from wrinklefree_inference.kernels import native
    rng = np.random.default_rng(1)
    w = rng.integers(-1, 2, size=(8, 256)).astype(np.float32)
    assert np.array_equal(native._lsb_to_i2s(native.pack_weights(w)), O.pack_ternary(w)[0])
    with pytest.raises(ValueError):
        native._lsb_to_i2s(np.zeros((2, 16), np.uint8))   # K=64


def test_page_allocator_matches_reference_semantics():
    from sgl_kernel.kvcache import PageAllocator
    ref = OM.PagedKVReference(num_layers=1, num_heads=1, head_dim=8, page_size=16, max_pages=100)
    mine = PageAllocator(100)
    assert mine.num_free_pages() == ref.num_free_pages() == 100
    p1, p2 = mine.allocate_pages(10), ref.allocate_pages(10)
    assert p1.dtype == torch.int32 and torch.equal(p1, p2)
    assert len(set(p1.tolist())) == 10
    assert mine.allocate_pages(91).shape == (0,) and mine.num_free_pages() == 90    # exhaustion: empty tensor, nothing taken
    allp = mine.allocate_pages(90)
    assert allp.shape == (90,) and mine.allocate_page() == -1
    mine.free_pages(p1)
    ref.free_pages(p2)
    assert mine.num_free_pages() == 10
    mine.free_page(1000)     # out of range ids are ignored
    assert mine.num_free_pages() == 10
    assert torch.equal(mine.allocate_pages(3), torch.tensor(p1.tolist()[::-1][:3], dtype=torch.int32))   # LIFO like the list


def test_kv_config_arithmetic(golden):
    from wrinklefree_inference.kv_cache import KVCacheConfig, KVCacheDtype, compute_kv_cache_memory
    g = golden["attention"]
    m = compute_kv_cache_memory({"num_layers": 30, "num_heads": 5, "head_dim": 128}, 4096, KVCacheDtype.BF16)
    assert m["total_mb"] == float(g["mem_total_mb"]) and m["per_token_kb"] == float(g["mem_per_token_kb"])
    assert KVCacheConfig(dtype=KVCacheDtype.INT8).bytes_per_token_per_layer == 2 * 32 * 128
    assert KVCacheConfig(dtype=KVCacheDtype.FP32).bytes_per_token_per_layer == 2 * 32 * 128 * 4


def test_sgl_kernel_argument_errors_are_value_errors():
    from sgl_kernel.quantization import bitnet_gemm, bitnet_gemv, BITNET_BLOCK_SIZE, QK_I2_S
    assert BITNET_BLOCK_SIZE == QK_I2_S == 128
    with pytest.raises(ValueError, match="multiple of 128"):
        bitnet_gemv(torch.zeros(4, 16, dtype=torch.uint8), torch.zeros(64, dtype=torch.int8), 1.0)
    with pytest.raises(ValueError, match="in_features"):
        bitnet_gemv(torch.zeros(4, 32, dtype=torch.uint8), torch.zeros(64, dtype=torch.int8), 1.0)
    with pytest.raises(ValueError):
        bitnet_gemm(torch.zeros(4, 32, dtype=torch.uint8), torch.zeros(128, dtype=torch.int8), 1.0)   # 1-D to gemm
    with pytest.raises(ValueError):
        bitnet_gemv(torch.zeros(4, 32, dtype=torch.uint8), torch.zeros(128, dtype=torch.float32), 1.0)


def test_torch_ops_are_registered_under_reference_names():
    import sgl_kernel  # noqa: F401
    for name in ("bitnet_gemv_cpu", "bitnet_gemm_cpu", "bitnet_quantize_activations_cpu"):
        assert hasattr(torch.ops.sgl_kernel, name)
    # fake (meta) kernels give the reference's shapes (docs/research/sglang_optimization_plan.md:59-75)
    pw = torch.empty(24, 32, dtype=torch.uint8, device="meta")
    x = torch.empty(5, 128, dtype=torch.int8, device="meta")
    assert torch.ops.sgl_kernel.bitnet_gemm_cpu(pw, x, 1.0).shape == (5, 24)
    assert torch.ops.sgl_kernel.bitnet_gemv_cpu(pw, x[0], 1.0).shape == (24,)
    q, s = torch.ops.sgl_kernel.bitnet_quantize_activations_cpu(torch.empty(3, 128, device="meta"))
    assert q.dtype == torch.int8 and q.shape == (3, 128) and s.shape == (1,)
    # the optional fused ops (sglang_optimization_plan.md:326-331)
    for name in ("bitnet_qkv_forward_cpu", "bitnet_mlp_forward_cpu"):
        assert hasattr(torch.ops.sgl_kernel, name)
    xm = torch.empty(5, 128, dtype=torch.bfloat16, device="meta")
    kw = torch.empty(8, 32, dtype=torch.uint8, device="meta")
    qo, ko, vo = torch.ops.sgl_kernel.bitnet_qkv_forward_cpu(xm, pw, kw, kw, 1.0, 1.0, 1.0)
    assert qo.shape == (5, 24) and ko.shape == (5, 8) and vo.shape == (5, 8) and qo.dtype == torch.bfloat16
    dw = torch.empty(128, 6, dtype=torch.uint8, device="meta")          # down: [hidden 128, ffn 24 / 4]
    y = torch.ops.sgl_kernel.bitnet_mlp_forward_cpu(xm, pw, pw, dw, 1.0, 1.0, 1.0, torch.empty(24, device="meta"), 1e-5)
    assert y.shape == (5, 128)


def test_oracle_model_runs_tiny_config():
    from wf_bitnet.synth import make_weights
    cfg = OM.ModelConfig(hidden=256, layers=2, heads=2, kv_heads=1, head_dim=128, ffn=384, vocab=512, max_seq=64)
    w = make_weights(cfg, seed=3)
    m = OM.OracleModel(cfg, w, use_ref=False)
    toks, gaps = m.generate_greedy([1, 2, 3], 4)
    assert len(toks) == 4 and all(0 <= t < 512 for t in toks) and all(g >= 0 for g in gaps)
    m2 = OM.OracleModel(cfg, w, use_ref=True)          # reference SIMD kernel inside the same forward
    toks2, _ = m2.generate_greedy([1, 2, 3], 4)
    assert toks2 == toks


def test_act_scale_needs_no_fp64():
    """csrc/common.cuh act_scale_from_absmax evaluates the reference's scale = fl32(fl64(max) / 127.0)
    (tests/test_bitnet_core.py:139-146) in fp32 only.  127 = 2^7 - 1 makes the quotient's expansion 7-periodic, so the
    double rounding can never differ from one fp32 rounding; and (double)max < 1e-6 <=> max <= fl32(1e-6)."""
    import numpy as np
    rng = np.random.default_rng(0)
    bits = rng.integers(0x00800000, 0x7F000000, size=4_000_000, dtype=np.int64).astype(np.uint32)
    a = np.concatenate([bits.view(np.float32), np.arange(1, 300000, dtype=np.float32) * np.float32(127.0),
                        np.float32(2.0) ** np.arange(-100, 100, dtype=np.float32), np.linspace(1e-7, 1e-5, 200001).astype(np.float32)])
    ref = (a.astype(np.float64) / 127.0).astype(np.float32)
    one = a / np.float32(127.0)
    assert np.array_equal(ref, one)
    lim = np.float32(1e-6)
    assert float(lim) < 1e-6 < float(np.nextafter(lim, np.float32(1)))
    assert np.array_equal(a.astype(np.float64) < 1e-6, a <= lim)


def test_cache_key_matches_reference_and_separates_pack_formats():
    """cache/cache_key.py:9-45 — golden keys produced by the reference's own function (imported from /root/reference when this
    test was written); the B200 tile format gets a key of its own."""
    from wrinklefree_inference.cache import compute_cache_key
    assert compute_cache_key("microsoft/bitnet-b1.58-2B-4T") == "microsoft_bitnet_b1_58_2B_4T_7573223ca919"
    assert compute_cache_key("microsoft/bitnet-b1.58-2B-4T", "abc123") == "microsoft_bitnet_b1_58_2B_4T_e3cc8c56ba2c"
    assert compute_cache_key("HF1BitLLM/Llama3-8B-1.58-100B-tokens") == "HF1BitLLM_Llama3_8B_1_58_100B_tokens_61cd867e670d"
    k = compute_cache_key("microsoft/bitnet-b1.58-2B-4T", pack_format="b200_tile_v1")
    assert k.startswith("microsoft_bitnet_b1_58_2B_4T_") and k != "microsoft_bitnet_b1_58_2B_4T_7573223ca919"


def test_sglang_model_mirror_imports_without_gpu():
    """tests/test_vllm_cpu_stub.py:346-354: the model module must import on a CUDA-less host; detection helpers are host logic."""
    import torch
    from sglang.srt.models.bitnet import BitNetForCausalLM, BitNetLinear, EntryClass, _is_packed_weight, _is_ternary_float
    assert EntryClass is BitNetForCausalLM and BitNetLinear is not None
    assert _is_packed_weight(torch.zeros(8, 4, dtype=torch.uint8)) and not _is_packed_weight(torch.zeros(8, 16))
    assert _is_ternary_float(torch.tensor([[-1.0, 0.0, 1.0]])) and not _is_ternary_float(torch.tensor([[0.5]]))


def test_loader_shards_like_tp_shard_weights():
    """The checkpoint loader's on-the-fly sharding (device tensors) must cut exactly what wf_bitnet.tp.shard_rows / shard_cols
    cut from the numpy form: q/k/v/gate/up by output rows, o/down by input columns in whole 128-blocks (32 packed bytes)."""
    import numpy as np
    import torch
    from wf_bitnet import loader, tp
    rng = np.random.default_rng(0)
    full = rng.integers(0, 256, size=(64, 96), dtype=np.uint8)          # N = 64, K = 384 (three 128-blocks)
    t = torch.from_numpy(full)
    for size in (1, 2):
        for rank in range(size):
            for name in ("q", "k", "v", "gate", "up"):
                want = full if size == 1 else tp.shard_rows(full, rank, size)
                assert np.array_equal(loader._shard(t, name, rank, size).numpy(), want)
    for rank in range(3):
        for name in ("o", "down"):
            assert np.array_equal(loader._shard(t, name, rank, 3).numpy(), tp.shard_cols(full, rank, 3))
    import pytest
    with pytest.raises(ValueError):
        loader._shard(t, "o", 0, 2)          # 96 bytes do not split into two runs of whole 32-byte blocks
    with pytest.raises(ValueError):
        loader._shard(t, "q", 0, 5)
