"""`sglang.srt.models.bitnet` mirror, written the way the reference's own tests for the fork's module read:
tests/test_bitnet_weight_loading.py:15-135, tests/test_bitnet_comparison.py:10-74, tests/test_vllm_cpu_stub.py:225-295.
Tensors are handed in on the CPU like there; the work happens on the GPU through the C ABI."""
import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu


def test_hf_unpack_known_bytes():
    from sglang.srt.models.bitnet import _unpack_ternary_weights
    for byte, val in ((0x00, -1.0), (0x55, 0.0), (0xAA, 1.0)):
        un = _unpack_ternary_weights(torch.full((64, 256), byte, dtype=torch.uint8))
        assert un.shape == (256, 256) and un.dtype == torch.float32
        assert torch.all(un == val)


def test_hf_to_kernel_format_conversion_roundtrip():
    from sglang.srt.models.bitnet import _pack_ternary_weights, _unpack_i2_to_ternary, _unpack_ternary_weights
    out_features, in_features = 256, 512
    g = torch.Generator().manual_seed(3)
    codes = torch.randint(0, 3, (4, out_features // 4, in_features), generator=g, dtype=torch.uint8)   # the 81 valid bytes
    hf_packed = codes[0] | (codes[1] << 2) | (codes[2] << 4) | (codes[3] << 6)
    unpacked = _unpack_ternary_weights(hf_packed)
    assert unpacked.shape == (out_features, in_features)
    assert torch.equal(unpacked, codes.reshape(out_features, in_features).float() - 1.0)
    kernel_packed, scale = _pack_ternary_weights(unpacked)
    assert kernel_packed.shape == (out_features, in_features // 4) and kernel_packed.dtype == torch.uint8
    assert scale.numel() == 1
    assert np.array_equal(kernel_packed.numpy(), O.pack_ternary(unpacked.numpy())[0])          # bit-exact vs oracle
    assert torch.equal(_unpack_i2_to_ternary(kernel_packed, out_features, in_features), unpacked)


def test_weight_kind_detection():
    from sglang.srt.models.bitnet import _is_packed_weight, _is_ternary_float
    assert _is_packed_weight(torch.zeros(256, 128, dtype=torch.uint8)) is True
    assert _is_packed_weight(torch.zeros(256, 512)) is False
    assert _is_ternary_float(torch.randint(-1, 2, (100, 100)).float()) is True
    assert _is_ternary_float(torch.randn(100, 100)) is False


def test_weight_scale_not_overwritten():
    from sglang.srt.models.bitnet import BitNetLinear
    linear = BitNetLinear(256, 512)
    linear.weight_scale.data.copy_(torch.tensor([1.7969]))
    assert linear.weight_scale.item() == pytest.approx(1.7969, rel=1e-3)
    assert linear.qweight.shape == (512, 64) and linear.qweight.dtype == torch.uint8
    with pytest.raises(ValueError):
        BitNetLinear(100, 64)


@pytest.mark.parametrize("batch", [1, 4, 33])
@pytest.mark.parametrize("finalized", [False, True])
def test_bitnet_linear_vs_oracle_and_float_matmul(batch, finalized):
    from sglang.srt.models.bitnet import BitNetLinear, _pack_ternary_weights
    out_features, in_features = 512, 256
    g = torch.Generator().manual_seed(11)
    weight = torch.randint(-1, 2, (out_features, in_features), generator=g).float()
    linear = BitNetLinear(in_features, out_features)
    packed, scale = _pack_ternary_weights(weight)
    linear.qweight.data.copy_(packed)
    linear.weight_scale.data.fill_(1.7)
    if finalized:
        linear.finalize()
    x = torch.randn(batch, in_features, generator=g) * 3
    y = linear(x)
    assert y.shape == (batch, out_features) and y.device == x.device
    # oracle: per-tensor absmax int8, exact integer dot, (sum (w+1)a * s_w - s_w * sum a) * s_a (direct_inference_full.py:186-208)
    q, s_a = O.quantize_act_i8(x.numpy())
    acc = q.astype(np.int64) @ weight.numpy().astype(np.int64).T
    want = acc.astype(np.float64) * 1.7 * s_a
    err = np.linalg.norm(y.numpy().astype(np.float64) - want) / np.linalg.norm(want)
    assert err <= 1e-3, err                                                               # north-star tolerance for dequantised outputs
    ref = torch.matmul(x, weight.t()) * 1.7
    cos = torch.nn.functional.cosine_similarity(ref.flatten().unsqueeze(0), y.flatten().unsqueeze(0)).item()
    assert cos > 0.99, cos


def test_forward_uses_weight_scale():
    from sglang.srt.models.bitnet import BitNetLinear, _pack_ternary_weights
    weight = torch.randint(-1, 2, (128, 256)).float()
    l1, l2 = BitNetLinear(256, 128), BitNetLinear(256, 128)
    packed, _ = _pack_ternary_weights(weight)
    l1.qweight.data.copy_(packed)
    l2.qweight.data.copy_(packed)
    l1.weight_scale.data.fill_(1.0)
    l2.weight_scale.data.fill_(2.0)
    x = torch.randn(1, 256)
    assert torch.allclose(l2(x), 2.0 * l1(x), rtol=1e-6, atol=0)


def test_causal_lm_matches_engine_and_oracle():
    """BitNetForCausalLM(config).load_weights(HF-named tensors) decodes the same greedy ids / logits as the engine fed the
    in-memory weights, which the engine tests pin to the oracle model."""
    from sglang.srt.models.bitnet import BitNetForCausalLM
    from test_gpu_loader import _state_dict
    from wf_bitnet.engine import Engine, EngineConfig
    from wf_bitnet.synth import make_weights
    hf_cfg = dict(hidden_size=256, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2, head_dim=128,
                  intermediate_size=384, vocab_size=512, rope_theta=500000.0, rms_norm_eps=1e-5, max_position_embeddings=64)
    model = BitNetForCausalLM(hf_cfg, max_batch=2)
    cfg = model.engine_config
    weights = make_weights(cfg, seed=5, keep_ternary=True)
    model.load_weights(_state_dict(cfg, weights, "hf").items())
    ref = Engine(EngineConfig(**{**cfg.__dict__}))
    ref.load_weights(weights)
    toks = [[3, 400], [17, 9], [250, 250]]
    for t in toks:
        got = model(torch.tensor(t))
        want = ref.decode(t)
        assert got.tolist() == want.tolist()
        assert torch.equal(model.logits(), ref.buffer("logits", (2, cfg.vocab), torch.float32))
    model.reset()
    ref.reset()
    assert model.generate([3, 400, 17], 6) == ref.generate_greedy([3, 400, 17], 6)


@pytest.mark.parametrize("batch", [1, 5, 40])
def test_fused_qkv_and_mlp_ops_match_the_oracle_step(batch):
    """torch.ops.sgl_kernel.bitnet_qkv_forward_cpu / bitnet_mlp_forward_cpu (sglang_optimization_plan.md:326-331) against the
    oracle's restatement of scripts/direct_inference_full.py:186-208 (linear) and :253-264 (MLP), row by row (each token is an
    independent copy of the reference's bs=1 path): bf16 outputs must be IDENTICAL — same integer accumulators, same rounding points."""
    import sgl_kernel  # noqa: F401
    from oracle import model as OM
    hidden, ffn, nq, nkv = 256, 384, 256, 128
    rng = np.random.default_rng(batch)
    mats = {n: rng.integers(-1, 2, size=shp).astype(np.float32) for n, shp in
            (("q", (nq, hidden)), ("k", (nkv, hidden)), ("v", (nkv, hidden)), ("gate", (ffn, hidden)), ("up", (ffn, hidden)), ("down", (hidden, ffn)))}
    packed = {n: O.pack_ternary(w)[0] for n, w in mats.items()}
    scales = {n: float(np.float32(rng.uniform(0.5, 2.5))) for n in mats}
    g = torch.Generator().manual_seed(batch)
    x = (torch.randn(batch, hidden, generator=g) * 2).to(torch.bfloat16)
    gamma = (1.0 + 0.1 * torch.randn(ffn, generator=g)).to(torch.bfloat16)
    dev = {n: torch.from_numpy(p).cuda() for n, p in packed.items()}
    q, k, v = torch.ops.sgl_kernel.bitnet_qkv_forward_cpu(x.cuda(), dev["q"], dev["k"], dev["v"], scales["q"], scales["k"], scales["v"])
    y = torch.ops.sgl_kernel.bitnet_mlp_forward_cpu(x.cuda(), dev["gate"], dev["up"], dev["down"], scales["gate"], scales["up"], scales["down"], gamma.cuda(), 1e-5)
    assert q.shape == (batch, nq) and k.shape == (batch, nkv) and y.shape == (batch, hidden) and y.dtype == torch.bfloat16
    for m in range(batch):
        for name, got in (("q", q), ("k", k), ("v", v)):
            want = OM.bitnet_linear(x[m], packed[name], scales[name], use_ref=False)
            assert torch.equal(got[m].cpu(), want), (name, m)
        gate = OM.bitnet_linear(x[m], packed["gate"], scales["gate"])
        up = OM.bitnet_linear(x[m], packed["up"], scales["up"])
        act = OM.rms_norm(torch.relu(gate).square() * up, gamma, 1e-5)
        want = OM.bitnet_linear(act, packed["down"], scales["down"])
        assert torch.equal(y[m].cpu(), want), ("mlp", m, (y[m].cpu().float() - want.float()).abs().max())
    # 1-D input (the reference's bs=1 call shape) and a second call through the cached fused tiles
    q1, _, _ = torch.ops.sgl_kernel.bitnet_qkv_forward_cpu(x[0].cuda(), dev["q"], dev["k"], dev["v"], scales["q"], scales["k"], scales["v"])
    assert q1.shape == (nq,) and torch.equal(q1.cpu(), q[0].cpu())
