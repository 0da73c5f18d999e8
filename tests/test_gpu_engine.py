"""GPU parity: the whole decode step (wf_engine) vs the oracle forward (scripts/direct_inference_full.py
restated in oracle/model.py) on identical synthetic random-init ternary weights."""
import numpy as np
import pytest
import torch

from oracle import model as OM

pytestmark = pytest.mark.gpu


def small_cfg(max_batch=1, **kw):
    from wf_bitnet.engine import EngineConfig
    base = dict(hidden=256, layers=2, heads=2, kv_heads=1, head_dim=128, ffn=384, vocab=512, max_seq=128, page_size=16, max_batch=max_batch)
    base.update(kw)
    return EngineConfig(**base)


def oracle_cfg(c):
    return OM.ModelConfig(hidden=c.hidden, layers=c.layers, heads=c.heads, kv_heads=c.kv_heads, head_dim=c.head_dim, ffn=c.ffn,
                          vocab=c.vocab, rope_theta=c.rope_theta, eps=c.eps, max_seq=c.max_seq)


@pytest.fixture(scope="module")
def small_model():
    from wf_bitnet.engine import Engine
    from wf_bitnet.synth import make_weights
    cfg = small_cfg(max_batch=4)
    w = make_weights(cfg, seed=1234)
    eng = Engine(cfg)
    eng.load_weights(w)
    yield cfg, w, eng
    eng.close()


def test_logits_track_the_oracle(small_model):
    cfg, w, eng = small_model
    om = OM.OracleModel(oracle_cfg(cfg), w, use_ref=True)
    eng.reset()
    toks = [5, 17, 400, 3, 3, 250, 9, 77]
    for pos, t in enumerate(toks):
        want = om.forward(t, pos).float()
        eng.decode([t])
        got = eng.buffer("logits", (cfg.max_batch, cfg.vocab), torch.float32)[0].cpu()
        rel = (got - want).norm() / want.norm()
        assert rel < 3e-2, (pos, rel.item())          # bf16 chain: a few bf16 ulps accumulate over the layers
        cos = torch.nn.functional.cosine_similarity(got, want, dim=0).item()
        assert cos > 0.999, (pos, cos)


def test_greedy_tokens_match_oracle(small_model):
    cfg, w, eng = small_model
    om = OM.OracleModel(oracle_cfg(cfg), w, use_ref=True)
    prompt = [11, 2, 300, 41, 7, 7, 128, 64]
    want, gaps = om.generate_greedy(prompt, 24)
    got = eng.generate_greedy(prompt, 24)
    # exact token equality wherever the oracle's own top-2 logit gap exceeds bf16 resolution; after a
    # near-tie divergence the sequences legitimately differ, so stop comparing there.
    compared = 0
    for i, (a, b, gap) in enumerate(zip(got, want, gaps)):
        if a != b:
            assert gap < 0.05, f"token {i}: got {a}, oracle {b}, gap {gap}"
            break
        compared += 1
    assert compared >= 12, f"only {compared} of 24 greedy tokens were compared before a near-tie divergence (gaps {gaps[:compared + 1]})"


def test_batched_equals_single_sequences(small_model):
    cfg, w, eng = small_model
    seqs = [[1, 2, 3, 4, 5, 6], [500, 400, 300, 200, 100, 50], [9, 9, 9, 9, 9, 9]]
    singles = []
    for s in seqs:
        eng.reset()
        outs = []
        for t in s:
            outs.append(int(eng.decode([t])[0]))
        singles.append(outs)
    eng.reset()
    batched = [[] for _ in seqs]
    for step in range(6):
        nxt = eng.decode([s[step] for s in seqs])
        for i in range(3):
            batched[i].append(int(nxt[i]))
    assert batched == singles          # per-token quantisation + exact integer GEMM => batch-invariant


def test_device_loop_equals_host_loop(small_model):
    cfg, w, eng = small_model
    eng.reset()
    first = int(eng.decode([42])[0])
    host = [first]
    for _ in range(10):
        host.append(int(eng.decode([host[-1]])[0]))
    eng.reset()
    eng.decode([42])                                  # leaves next token in the device buffer; feed it back on device
    tok = eng.buffer("next", (cfg.max_batch,), torch.int32)
    assert int(tok[0]) == first
    # device loop: copy next -> tokens once, then run 10 graph steps with on-device feedback
    from wf_bitnet._lib import lib, check
    check(lib().wf_copy_d2d(lib().wf_engine_buffer(eng._h, b"tokens"), lib().wf_engine_buffer(eng._h, b"next"), 4, None))
    torch.cuda.synchronize()
    dev = []
    for _ in range(10):
        eng.decode_device(1, use_graph=True)
        dev.append(int(eng.buffer("next", (cfg.max_batch,), torch.int32)[0]))
    assert dev == host[1:]
    assert eng.launches_per_step() >= 2 * cfg.layers + 3      # embed-phase + {attention, layer tail} per layer + final glue, LM head, advance


def test_full_size_2b_step_runs_and_is_finite():
    from wf_bitnet.engine import Engine, EngineConfig
    cfg = EngineConfig.bitnet_2b(max_seq=256, max_batch=2)
    eng = Engine(cfg)
    eng.init_synthetic(7)
    for slot in range(2):
        eng.set_context(slot, 40 + slot, fill_random=True)
    out = eng.decode([1, 2])
    assert out.shape == (2,) and all(0 <= int(t) < cfg.vocab for t in out)
    lg = eng.buffer("logits", (2, cfg.vocab), torch.float32)
    assert torch.isfinite(lg).all() and lg.abs().max() > 0
    out2 = eng.decode([int(out[0]), int(out[1])])
    assert out2.shape == (2,)
    eng.close()


def _run_tokens(cfg, w, fused, steps, monkeypatch, ctx=None, synthetic=None, even="0", taps=()):
    """logits of every step from a fresh engine built with WF_FUSED=fused"""
    from wf_bitnet.engine import Engine
    monkeypatch.setenv("WF_FUSED", "1" if fused else "0")
    # bit-identity across the two step forms needs the same split of the softmax sums in both: the even work split of the fused
    # attention (one resident wave, range boundaries anywhere) changes the fp32 summation order, so it is compared on its own below
    monkeypatch.setenv("WF_ATTN_EVEN", even)
    eng = Engine(cfg)
    if synthetic is not None:
        eng.init_synthetic(synthetic)
    else:
        eng.load_weights(w)
    B = len(steps[0])
    if ctx:
        for slot in range(B):
            eng.set_context(slot, ctx + 3 * slot, fill_random=True)
    outs = []
    for toks in steps:
        nxt = eng.decode(toks)
        outs.append((eng.buffer("logits", (cfg.max_batch, cfg.vocab), torch.float32)[:B].cpu(), nxt.copy(), eng.launches_per_step()) +
                    tuple(eng.buffer(n, sh, dt).cpu() for (n, sh, dt) in taps))
    eng.close()
    return outs


@pytest.mark.parametrize("B,ctx,kv", [(64, 300, "bf16"), (16, 1200, "bf16"), (64, 1800, "bf16"), (33, 1500, "int8")])
def test_attention_even_split_matches_chunked_split(B, ctx, kv, monkeypatch):
    """The fused attention's even work split (one resident wave of CTAs, each an equal contiguous range of the (sequence, kv head,
    block) space, partials merged by the last CTA of a pair) against the chunked grid on the same engine state: the two differ only
    in where the online-softmax partial sums are cut, i.e. by fp32 summation order — the bf16 attention output of a 1-layer model
    must agree to a bf16 ulp or two, and the greedy tokens must be the same."""
    from wf_bitnet.engine import EngineConfig
    cfg = EngineConfig.bitnet_2b(max_seq=2048, max_batch=B, layers=1, vocab=4096)
    if kv == "int8":
        cfg.kv_dtype = "int8"
    rng = np.random.default_rng(70 + B)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(2)]
    taps = (("attn", (cfg.max_batch, cfg.heads * cfg.head_dim), torch.bfloat16),)
    a = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5, even="1", taps=taps)
    b = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5, even="0", taps=taps)
    (la, na, _, xa), (lb, nb, _, xb) = a[0], b[0]
    xa, xb = xa[:B].float(), xb[:B].float()
    assert torch.isfinite(xa).all() and xa.abs().max() > 0
    err = (xa - xb).abs().max().item()
    assert err <= 2.0 ** -7 * xb.abs().max().item(), err          # bf16: 8 significant bits
    assert (xa != xb).float().mean().item() < 0.2                # most elements round to the same bf16
    assert (na == nb).mean() >= 0.9                              # random-init logits have near ties: a flipped last bit may flip an argmax


@pytest.mark.parametrize("B", [1, 3, 8, 12, 33, 40, 64])
def test_phase_kernels_equal_per_op_kernels(B, monkeypatch):
    """The fused phase path (glue prologues, raw fp32 accumulators, fused attention) and the one-kernel-per-op path share
    every rounding point of the reference step, so their logits must be IDENTICAL, not merely close."""
    from wf_bitnet.synth import make_weights
    cfg = small_cfg(max_batch=B, layers=3)
    w = make_weights(cfg, seed=99)
    rng = np.random.default_rng(B)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(20)]   # crosses a page boundary (16)
    a = _run_tokens(cfg, w, True, steps, monkeypatch)
    b = _run_tokens(cfg, w, False, steps, monkeypatch)
    assert a[0][2] < b[0][2]          # fewer launches
    for i, ((la, na, _), (lb, nb, _)) in enumerate(zip(a, b)):
        assert torch.equal(la, lb), (i, (la - lb).abs().max().item())
        assert (na == nb).all()


@pytest.mark.parametrize("B", [1, 3, 8, 12, 33, 64])
def test_layer_tail_kernel_equals_phase_kernels(B, monkeypatch):
    """Two fused forms exist: the persistent layer-tail kernels (default: o -> gate|up -> down -> next q|k|v in one launch, grid
    barriers between the phases, whole-row-block ownership; layer_tail.cu for B <= 8, layer_tail_batch.cu above) and the per-phase
    kernels with split-K REDs (WF_TAIL=0, the form the tensor-parallel engines use).  Same rounding points, exact integer
    contraction: identical logits, and far fewer launches."""
    from wf_bitnet.synth import make_weights
    cfg = small_cfg(max_batch=B, layers=3)
    w = make_weights(cfg, seed=98)
    rng = np.random.default_rng(40 + B)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(20)]
    monkeypatch.setenv("WF_TAIL", "1")
    monkeypatch.setenv("WF_TAIL_BATCH", "1")     # the batched form is opt-in (the phase kernels measure faster at batch 64)
    monkeypatch.setenv("WF_SMALL_MAX", "8")      # and so are the GEMV-style kernels above batch 1: keep both covered
    a = _run_tokens(cfg, w, True, steps, monkeypatch)
    monkeypatch.setenv("WF_TAIL", "0")
    b = _run_tokens(cfg, w, True, steps, monkeypatch)
    assert a[0][2] <= 2 * cfg.layers + 6 and b[0][2] > a[0][2]      # embed-phase + {attention, tail} per layer + final glue, LM head (+ argmax), advance
    for i, ((la, na, _), (lb, nb, _)) in enumerate(zip(a, b)):
        assert torch.equal(la, lb), (i, (la - lb).abs().max().item())
        assert (na == nb).all()


@pytest.mark.parametrize("kv", ["bf16", "int8"])
def test_attention_two_class_grid_matches_uniform_grid(kv, monkeypatch):
    """320 (sequence, kv head) pairs on 296 resident CTA slots: the two-class grid (296 whole pairs as one CTA each, the other 24 pairs
    cut into short chunks that fill one more wave) against the uniform grid.  Whole pairs are not split at all, so their softmax sums
    are cut differently from the uniform grid's chunks: same tolerance as the even-split test (a bf16 ulp or two, same greedy tokens
    up to near ties)."""
    from wf_bitnet.engine import EngineConfig
    B, ctx = 64, 1800
    cfg = EngineConfig.bitnet_2b(max_seq=2048, max_batch=B, layers=1, vocab=4096)
    if kv == "int8":
        cfg.kv_dtype = "int8"
    rng = np.random.default_rng(71)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(2)]
    taps = (("attn", (cfg.max_batch, cfg.heads * cfg.head_dim), torch.bfloat16),)
    monkeypatch.setenv("WF_ATTN_TWO", "1")
    a = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5, taps=taps)
    monkeypatch.setenv("WF_ATTN_TWO", "0")
    b = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5, taps=taps)
    (la, na, _, xa), (lb, nb, _, xb) = a[0], b[0]
    xa, xb = xa[:B].float(), xb[:B].float()
    assert torch.isfinite(xa).all() and xa.abs().max() > 0
    assert (xa - xb).abs().max().item() <= 2.0 ** -7 * xb.abs().max().item()
    assert (xa != xb).float().mean().item() < 0.2
    assert (na == nb).mean() >= 0.9


@pytest.mark.parametrize("kv", ["bf16", "int8"])
def test_attention_per_sequence_chunks_match_fixed_chunks(kv, monkeypatch):
    """Sequences of 1100 tokens under a 2048-token length hint (max_seq 4096): fixed chunks of the hint leave the later chunks of a
    pair nearly empty, per-sequence chunking (default) cuts the actual length evenly.  Same attention output up to the summation order
    of the softmax partials (tolerance of the two-class test)."""
    from wf_bitnet.engine import EngineConfig
    B, ctx = 64, 1100
    cfg = EngineConfig.bitnet_2b(max_seq=4096, max_batch=B, layers=1, vocab=4096)
    if kv == "int8":
        cfg.kv_dtype = "int8"
    rng = np.random.default_rng(72)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(2)]
    taps = (("attn", (cfg.max_batch, cfg.heads * cfg.head_dim), torch.bfloat16),)
    monkeypatch.setenv("WF_ATTN_ADAPT", "1")
    a = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5, taps=taps)
    monkeypatch.setenv("WF_ATTN_ADAPT", "0")
    b = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5, taps=taps)
    (la, na, _, xa), (lb, nb, _, xb) = a[0], b[0]
    xa, xb = xa[:B].float(), xb[:B].float()
    assert torch.isfinite(xa).all() and xa.abs().max() > 0
    assert (xa - xb).abs().max().item() <= 2.0 ** -7 * xb.abs().max().item()
    assert (xa != xb).float().mean().item() < 0.2
    assert (na == nb).mean() >= 0.9


@pytest.mark.parametrize("shape", ["2b", "8b"])
def test_layer_tail_handoff_forms_equal_phase_kernels(shape, monkeypatch):
    """Batch 1 at the real layer widths: the persistent tail kernel with flag-in-data hand-offs (default at the 2B shape; the 8B ffn row
    is too long for the register glue, so there the default IS the grid-barrier form), the same kernel with grid barriers (WF_TAIL_LAM=0) and the
    per-phase kernels (WF_TAIL=0) must produce identical logits over a run of steps — a lost or early hand-off would flip bits."""
    import dataclasses
    from wf_bitnet.engine import EngineConfig
    mk = EngineConfig.bitnet_2b if shape == "2b" else EngineConfig.llama3_8b
    cfg = dataclasses.replace(mk(max_seq=1024, max_batch=1), layers=3, vocab=4096)
    rng = np.random.default_rng(3)
    steps = [[int(rng.integers(0, cfg.vocab))] for _ in range(12)]
    runs = {}
    monkeypatch.setenv("WF_SMALL_MAX", "1")   # (at the 8B widths batch 1 defaults to the row glue + streaming GEMM; this test is about the GEMV-style forms)
    for name, env in (("handoff", {"WF_TAIL": "1", "WF_TAIL_LAM": "1"}), ("barrier", {"WF_TAIL": "1", "WF_TAIL_LAM": "0"}), ("phase", {"WF_TAIL": "0", "WF_TAIL_LAM": "1"})):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        runs[name] = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=600, synthetic=9)
    assert runs["handoff"][0][2] == runs["barrier"][0][2] < runs["phase"][0][2]          # launches per step
    for name in ("handoff", "barrier"):
        for i, ((la, na, _), (lb, nb, _)) in enumerate(zip(runs[name], runs["phase"])):
            assert torch.equal(la, lb), (name, i, (la - lb).abs().max().item())
            assert (na == nb).all()


@pytest.mark.parametrize("B,ctx", [(1, 500), (4, 700), (16, 1200), (40, 100), (64, 300)])
def test_phase_kernels_equal_per_op_kernels_2b_shape(B, ctx, monkeypatch):
    """same identity at the BitNet-2B layer shapes (split attention chunks, 54 k-blocks) on a 4-layer slice"""
    from wf_bitnet.engine import EngineConfig
    cfg = EngineConfig.bitnet_2b(max_seq=2048, max_batch=B, layers=4, vocab=4096)
    rng = np.random.default_rng(7 + B)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(3)]
    a = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5)
    b = _run_tokens(cfg, None, False, steps, monkeypatch, ctx=ctx, synthetic=5)
    for i, ((la, na, _), (lb, nb, _)) in enumerate(zip(a, b)):
        assert torch.isfinite(la).all()
        assert torch.equal(la, lb), (i, (la - lb).abs().max().item())
        assert (na == nb).all()


@pytest.mark.parametrize("B,ctx", [(40, 100), (64, 300)])
def test_tcgen05_decode_linears_equal_per_op_kernels(B, ctx, monkeypatch):
    """WF_DECODE_TC=1: the batched decode linears on tcgen05 kind::i8 with split-K fp32 REDs (gemm_tc.cu, decode form) give the
    same logits, bit for bit, as the one-kernel-per-op path at the 2B layer shapes."""
    from wf_bitnet.engine import EngineConfig
    cfg = EngineConfig.bitnet_2b(max_seq=2048, max_batch=B, layers=3, vocab=4096)
    rng = np.random.default_rng(17 + B)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(3)]
    monkeypatch.setenv("WF_DECODE_TC", "1")
    a = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5)
    monkeypatch.setenv("WF_DECODE_TC", "0")      # the mma.sync one-shot GEMM (the default below WF_DECODE_TC_MIN tokens) at the same batch
    c = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5)
    b = _run_tokens(cfg, None, False, steps, monkeypatch, ctx=ctx, synthetic=5)
    for i, ((la, na, _), (lb, nb, _), (lc, nc, _)) in enumerate(zip(a, b, c)):
        assert torch.equal(la, lb), (i, (la - lb).abs().max().item())
        assert torch.equal(lc, lb), (i, (lc - lb).abs().max().item())
        assert (na == nb).all() and (nc == nb).all()


@pytest.mark.parametrize("shape,B,ctx", [("2b", 2, 80), ("2b", 13, 200), ("2b", 32, 120), ("8b", 1, 90), ("8b", 5, 90), ("8b", 16, 150), ("8b", 32, 60)])
def test_streaming_decode_gemm_equals_default_kernels(shape, B, ctx, monkeypatch):
    """WF_DECODE_STREAM=2: every decode linear of 2..32 rows on the streaming batched GEMM (gemm_stream.cu: K windows, TMA ring, masked
    2-bit fields on mma.sync with M/8 n-tiles, fp32 REDs) — same logits, bit for bit, as the step on the default linears (one-shot /
    tcgen05 forms; those are held against the one-kernel-per-op path by the tests above).  Both runs share the attention kernel and its
    chunking, so any difference is the GEMM's.  2B widths: one window (K = 2560) and two windows of 27 k-blocks (K = 6912); 8B widths:
    one window of 32 and four of 28."""
    import dataclasses
    from wf_bitnet.engine import EngineConfig
    mk = EngineConfig.bitnet_2b if shape == "2b" else EngineConfig.llama3_8b
    cfg = dataclasses.replace(mk(max_seq=1024, max_batch=B), layers=2, vocab=4096)
    rng = np.random.default_rng(23 + B)
    steps = [list(map(int, rng.integers(0, cfg.vocab, B))) for _ in range(3)]
    if B == 1:
        monkeypatch.setenv("WF_SMALL_MAX", "0")   # batch 1 through the row glue + batched GEMM in both runs (the default at the 8B widths)
    monkeypatch.setenv("WF_DECODE_STREAM", "2")
    a = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5)
    monkeypatch.setenv("WF_DECODE_STREAM", "0")
    c = _run_tokens(cfg, None, True, steps, monkeypatch, ctx=ctx, synthetic=5)
    for i, ((la, na, _), (lc, nc, _)) in enumerate(zip(a, c)):
        assert torch.isfinite(la).all()
        assert torch.equal(la, lc), (i, (la - lc).abs().max().item())
        assert (na == nc).all()


# ---------------------------------------------------------------------------------------------------------------------
# BASELINE.json configs[0]: the full BitNet-2B-4T shape (30 layers, hidden 2560, 20/5 heads, ffn 6912, vocab 128256),
# batch 1, 512-token prompt, 128 greedy tokens -- the protocol of extern/BitNet/utils/e2e_benchmark.py:36-53 run through
# the generate loop of scripts/direct_inference_full.py:328-371 -- GPU engine vs the reference CPU path (oracle forward on
# the reference's own unmodified AVX2 kernel) on identical synthetic random-init ternary weights.
#
# What "the same tokens within bf16 tolerance" can mean for THIS model has to be measured, not assumed: a deep random-init
# network with int8 re-quantisation at every linear is chaotic.  The oracle run against ITSELF with one bf16-level change
# (attention evaluated in fp32 and rounded once, instead of torch's bf16 SDPA kernel: OracleModel(attn_fp32=True)) ends, after
# 30 layers and a few dozen tokens of context, 40-60 % away in normwise logit distance with most greedy tokens changed; with
# 4 layers and several hundred tokens of context it is already percent-level.  No two implementations that differ anywhere in
# attention rounding can agree token for token there, the reference with itself included.  So the tests pin what CAN be
# pinned and bound the rest by that measured sensitivity:
#   (1) full depth, position 0 (attention over one token is the identity, every other op has the reference's rounding points):
#       the final normalised hidden state must be BIT-IDENTICAL to the oracle's after 30 layers, the greedy token equal;
#   (2) ONE layer of the same model (same width, GQA 4:1, RoPE, 128256-row LM head) through the whole 512 + 128 protocol:
#       every greedy token outside bf16 near-ties equal, logits within 1e-2 normwise;
#   (3) 4 layers and the full 30 layers through the whole protocol, teacher-forced: the engine must stay as close to the oracle
#       as the oracle stays to itself under the fp32-attention change, measured over the same 128 generated steps.
CFG0_PROMPT, CFG0_NEW = 512, 128
# A greedy token is only defined up to the resolution of the logits.  The reference's logits are bf16 (F.linear on bf16
# tensors, direct_inference_full.py:297-300) — with 128256 random logits the top two are often EQUAL in bf16.  Steps whose
# oracle top-2 gap is below CFG0_TIE_ULPS bf16 ulps (ulp = 2^-7 * 2^floor(log2 |top|)) are near-ties and are not compared.
CFG0_TIE_ULPS = 2.0


def _bf16_ulp(x: float) -> float:
    import math
    return 2.0 ** (math.floor(math.log2(max(abs(x), 1e-30))) - 7)


@pytest.fixture(scope="module")
def cfg0_weights():
    from wf_bitnet.engine import EngineConfig
    from wf_bitnet.synth import make_weights
    cfg = EngineConfig.bitnet_2b(max_seq=CFG0_PROMPT + CFG0_NEW + 16, max_batch=1)
    return cfg, make_weights(cfg, seed=1234)


def _sliced(cfg, w, layers):
    import dataclasses
    c = dataclasses.replace(cfg, layers=layers)
    ww = dict(w)
    ww["layers"] = w["layers"][:layers]
    return c, ww


def test_config0_full_depth_position0_is_bit_identical(cfg0_weights):
    from wf_bitnet.engine import Engine
    cfg, w = cfg0_weights
    om = OM.OracleModel(oracle_cfg(cfg), w, use_ref=True)
    eng = Engine(cfg)
    eng.load_weights(w)
    for t in (0, 7, 4242, 65000, 128255):
        eng.reset()
        logits, hid = om.forward(t, 0, return_hidden=True)
        got = int(eng.decode([t])[0])
        hg = eng.buffer("normed", (1, cfg.hidden), torch.bfloat16)[0].cpu()
        lg = eng.buffer("logits", (1, cfg.vocab), torch.float32)[0].cpu()
        assert torch.equal(hg, hid), f"token {t}: {int((hg != hid).sum())} of {cfg.hidden} final hidden values differ after 30 layers"
        rel = float((lg - logits.float()).norm() / logits.float().norm())
        assert rel < 2e-3, (t, rel)                                  # fp32-accumulated LM head vs the reference's bf16 logits
        top2 = torch.topk(logits.float(), 2)
        if float(top2.values[0] - top2.values[1]) >= CFG0_TIE_ULPS * _bf16_ulp(float(top2.values[0])):
            assert got == int(top2.indices[0]), t
    eng.close()


def _protocol(cfg, w, with_sensitivity):
    """512-token prompt + 128 greedy steps, engine (and the fp32-attention oracle) teacher-forced with the oracle's tokens.
    Returns per generated step: (step, engine token, oracle token, top-2 gap, near-tie, rel(engine, oracle), rel(oracle32, oracle),
    oracle32 token)."""
    from wf_bitnet.engine import Engine
    om = OM.OracleModel(oracle_cfg(cfg), w, use_ref=True)
    om32 = OM.OracleModel(oracle_cfg(cfg), w, use_ref=True, attn_fp32=True) if with_sensitivity else None
    eng = Engine(cfg)
    eng.load_weights(w)
    prompt = np.random.default_rng(0).integers(0, cfg.vocab, size=CFG0_PROMPT)   # randint(0, 128256), seed 0
    for pos, t in enumerate(prompt):       # the reference's generate() prefill is exactly this token-by-token loop (:343-347)
        om.forward(int(t), pos)
        eng.decode([int(t)])
        if om32 is not None:
            om32.forward(int(t), pos)
    nxt, pos = int(prompt[-1]), CFG0_PROMPT   # the reference feeds the last prompt token again at pos = len(prompt) (:349-353)
    steps = []
    for step in range(CFG0_NEW):
        lo = om.forward(nxt, pos).float()
        got = int(eng.decode([nxt])[0])
        top2 = torch.topk(lo, 2)
        want, gap = int(top2.indices[0]), float(top2.values[0] - top2.values[1])
        lg = eng.buffer("logits", (1, cfg.vocab), torch.float32)[0].cpu()
        rel32, tok32 = float("nan"), -1
        if om32 is not None:
            l32 = om32.forward(nxt, pos).float()
            rel32, tok32 = float((l32 - lo).norm() / lo.norm()), int(l32.argmax())
        steps.append((step, got, want, gap, gap < CFG0_TIE_ULPS * _bf16_ulp(float(top2.values[0])), float((lg - lo).norm() / lo.norm()), rel32, tok32))
        nxt, pos = want, pos + 1           # teacher forcing: every side keeps seeing the same context
    eng.close()
    return steps


def test_config0_2b_width_1_layer_greedy_128_after_512(cfg0_weights):
    """(2): full width / GQA / RoPE / vocab, one layer, the whole 512 + 128 protocol: exact greedy tokens outside near-ties."""
    cfg, w = _sliced(*cfg0_weights, 1)
    steps = _protocol(cfg, w, False)
    compared = [s for s in steps if not s[4]]
    bad = [s for s in compared if s[1] != s[2]]
    worst = max(s[5] for s in steps)
    print(f"config0 / 1 layer: compared {len(compared)}/{CFG0_NEW} steps ({CFG0_NEW - len(compared)} near-ties skipped), mismatches {len(bad)}, "
          f"worst normwise logit error {worst:.5f}")
    assert not bad, f"first divergence at step {bad[0][0]}: engine {bad[0][1]} != oracle {bad[0][2]} at top-2 gap {bad[0][3]:.4f}"
    assert len(compared) >= 64, f"only {len(compared)} of {CFG0_NEW} steps had a top-2 gap above the bf16 tie bound"
    # bf16 attention in the engine vs the oracle's bf16-rounded fp32 attention: an int8 activation code flipping by one step
    # moves the normwise logit error to ~2e-2 (measured worst 0.0204 over the 128 steps); a wrong kernel is O(1)
    assert worst < 5e-2


@pytest.mark.parametrize("layers", [4, 30])
def test_config0_2b_greedy_128_after_512(cfg0_weights, layers):
    """(3): 4 layers and the full 30-layer model through the whole protocol, judged against the oracle's own measured sensitivity."""
    cfg, w = _sliced(*cfg0_weights, layers)
    steps = _protocol(cfg, w, True)
    eng_rel = float(np.median([s[5] for s in steps]))
    self_rel = float(np.median([s[6] for s in steps]))
    eng_agree = sum(1 for s in steps if s[1] == s[2]) / len(steps)
    self_agree = sum(1 for s in steps if s[7] == s[2]) / len(steps)
    first_div = next((s for s in steps if s[1] != s[2] and not s[4]), None)
    print(f"config0 / {layers} layers, 128 generated steps after the 512-token prompt: engine-vs-oracle median logit error {eng_rel:.4f}, greedy-token agreement "
          f"{eng_agree:.2f}; oracle(fp32 attention)-vs-oracle {self_rel:.4f}, {self_agree:.2f}; first engine divergence outside near-ties: "
          f"{first_div[:4] if first_div else None}")
    assert eng_rel <= 1.5 * self_rel + 0.02, (eng_rel, self_rel)     # no further from the oracle than the oracle is from itself
    assert eng_agree >= self_agree - 0.2, (eng_agree, self_agree)
