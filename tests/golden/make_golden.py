#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by IMPORTING THE REFERENCE'S OWN PYTHON.

Run in the build container only (needs /root/reference; the GPU box does not have it):

    python tests/golden/make_golden.py

Sources imported (never copied):
  /root/reference/tests/test_bitnet_core.py           (_pack_ternary_weights, _unpack_ternary_weights,
                                                       _unpack_i2_to_ternary, quantize_activations_i8,
                                                       bitnet_gemv_reference — verbatim copies of the
                                                       missing sgl-kernel / sglang-bitnet Python)
  /root/reference/src/wrinklefree_inference/cache/bitnet_converter.py   (_pack_ternary_weights, _is_ternary_float)
  /root/reference/src/wrinklefree_inference/kv_cache/kv_cache.py        (KVCache, attention_with_kv_cache)
  /root/reference/src/wrinklefree_inference/kernels/native.py           (numpy pack/unpack/repack/quantize)
  /root/reference/legacy/benchmark/test_sglang_bitnet_kernel.py         (the sgl-kernel Python fallback)

Everything is seeded; the outputs are small .npz files committed next to this script.
"""
import importlib.util
import sys
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, str(path))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def main():
    core = _load("ref_test_bitnet_core", REF / "tests/test_bitnet_core.py")
    sys.path.insert(0, str(REF / "src"))
    conv = _load("ref_bitnet_converter", REF / "src/wrinklefree_inference/cache/bitnet_converter.py")
    kvmod = _load("ref_kv_cache", REF / "src/wrinklefree_inference/kv_cache/kv_cache.py")
    native = _load("ref_native", REF / "src/wrinklefree_inference/kernels/native.py")
    fallback = _load("ref_fallback", REF / "legacy/benchmark/test_sglang_bitnet_kernel.py")

    g = torch.Generator().manual_seed(20261104)

    # ---- 1. packing (blocked I2_S), incl. the K<128 special case and the converter twin
    pack = {}
    for tag, (n, k) in {"a": (64, 256), "b": (16, 128), "c": (8, 64), "d": (40, 640)}.items():
        w = torch.randint(-1, 2, (n, k), generator=g).float()
        p1, s1 = core._pack_ternary_weights(w)
        p2, s2 = conv._pack_ternary_weights(w)
        assert torch.equal(p1, p2) and torch.equal(s1, s2)
        un = core._unpack_i2_to_ternary(p1, n, k)
        assert torch.equal(un, w)
        pack[f"w_{tag}"] = w.numpy().astype(np.int8)
        pack[f"packed_{tag}"] = p1.numpy()
        pack[f"scale_{tag}"] = s1.numpy()
    # non-ternary float input: rounding / clamping / scale behaviour
    wf = (torch.randn(8, 128, generator=g) * 0.8)
    pf, sf = core._pack_ternary_weights(wf)
    pack["w_float"] = wf.numpy()
    pack["packed_float"] = pf.numpy()
    pack["scale_float"] = sf.numpy()
    pack["is_ternary_float_true"] = np.array(conv._is_ternary_float(torch.randint(-1, 2, (4, 8)).float()))
    pack["is_ternary_float_false"] = np.array(conv._is_ternary_float(torch.randn(4, 8) * 3))
    # layout KATs (tests/test_bitnet_core.py:186-203 uses the HF function; same bytes through the I2_S unpack)
    for byte in (0x00, 0x55, 0xAA, 0x1B, 0xE4):
        pk = torch.full((2, 32), byte, dtype=torch.uint8)
        pack[f"kat_{byte:02x}"] = core._unpack_i2_to_ternary(pk, 2, 128).numpy().astype(np.int8)
    np.savez_compressed(OUT / "pack.npz", **pack)

    # ---- 2. HF checkpoint layout
    hf = {}
    valid = np.array([b0 | (b1 << 2) | (b2 << 4) | (b3 << 6)
                      for b0 in range(3) for b1 in range(3) for b2 in range(3) for b3 in range(3)], np.uint8)
    idx = torch.randint(0, len(valid), (16, 256), generator=g).numpy()
    hfp = valid[idx]
    un = core._unpack_ternary_weights(torch.from_numpy(hfp))
    kp, _ = core._pack_ternary_weights(un)
    hf["hf_packed"] = hfp
    hf["unpacked"] = un.numpy().astype(np.int8)
    hf["kernel_packed"] = kp.numpy()
    # native.py's (interleaved) view of the same bytes — pinned only by its own round trip
    hf["native_repack"] = native.repack_hf_weights(hfp)
    np.savez_compressed(OUT / "hf.npz", **hf)

    # ---- 3. activation quantisation
    aq = {}
    cases = {
        "randn3": torch.randn(4, 256, generator=g) * 3,
        "vec": torch.randn(2560, generator=g),
        "zeros": torch.zeros(3, 128),
        "tiny": torch.full((1, 128), 5e-7),
        "ties": torch.tensor([[0.5, 1.5, 2.5, -0.5, -1.5, -2.5, 127.0, -127.0] * 16]),
        "bf16ish": (torch.randn(2, 384, generator=g) * 7).to(torch.bfloat16).float(),
    }
    for tag, x in cases.items():
        q, s = core.quantize_activations_i8(x)
        q2, s2 = fallback.quantize_activations_i8(x)
        assert torch.equal(q, q2) and s == s2
        aq[f"x_{tag}"] = x.numpy()
        aq[f"q_{tag}"] = q.numpy()
        aq[f"s_{tag}"] = np.array(s, np.float64)
    # native.py variant truncates instead of rounding (documented divergence)
    xn = cases["vec"].numpy()
    qn, sn = native.quantize_activations(xn)
    aq["native_q_vec"] = qn
    aq["native_s_vec"] = np.array(sn, np.float64)
    np.savez_compressed(OUT / "actquant.npz", **aq)

    # ---- 4. GEMV / GEMM semantic reference + Sigma-a correction
    gv = {}
    for tag, (n, k) in {"a": (128, 128), "b": (64, 256), "c": (96, 640)}.items():
        w = torch.randint(-1, 2, (n, k), generator=g).float()
        p, _ = core._pack_ternary_weights(w)
        x = torch.randn(k, generator=g) * 3
        xq, sa = core.quantize_activations_i8(x)
        ws = 1.7
        y = core.bitnet_gemv_reference(p, xq, ws)
        act_sum = xq.float().sum()
        corrected = (y - ws * act_sum) * sa
        gv[f"packed_{tag}"] = p.numpy()
        gv[f"x_{tag}"] = x.numpy()
        gv[f"xq_{tag}"] = xq.numpy()
        gv[f"sa_{tag}"] = np.array(sa, np.float64)
        gv[f"y_{tag}"] = y.numpy()
        gv[f"corrected_{tag}"] = corrected.numpy()
        # python-fallback semantics (LSB-sequential layout, sum w*a): same ternary matrix
        pl = fallback.pack_ternary_weights(w)
        gv[f"lsb_packed_{tag}"] = pl.numpy()
        gv[f"lsb_y_{tag}"] = fallback.bitnet_gemv(pl, xq, ws).numpy()
    # all-ones KAT (tests/test_bitnet_weight_loading.py:141-169): W=+1, x=1, scale 2.0
    k = 256
    w1 = torch.ones(32, k)
    p1, _ = core._pack_ternary_weights(w1)
    x1 = torch.ones(k, dtype=torch.int8)
    gv["ones_packed"] = p1.numpy()
    gv["ones_y"] = core.bitnet_gemv_reference(p1, x1, 2.0).numpy()
    # batched: per-tensor quant then row-wise reference
    w = torch.randint(-1, 2, (48, 256), generator=g).float()
    p, _ = core._pack_ternary_weights(w)
    X = torch.randn(5, 256, generator=g) * 2
    Xq, sx = core.quantize_activations_i8(X)
    Y = torch.stack([core.bitnet_gemv_reference(p, Xq[i], 0.75) for i in range(5)])
    gv["gemm_packed"] = p.numpy()
    gv["gemm_X"] = X.numpy()
    gv["gemm_Xq"] = Xq.numpy()
    gv["gemm_sx"] = np.array(sx, np.float64)
    gv["gemm_Y"] = Y.numpy()
    np.savez_compressed(OUT / "gemv.npz", **gv)

    # ---- 5. attention_with_kv_cache (MHA, fp32 and bf16 caches; int8 cache kept as a documented variant)
    at = {}
    L, H, D, S = 2, 4, 32, 24
    for tag, dt, qdt in (("fp32", kvmod.KVCacheDtype.FP32, torch.float32),
                         ("bf16", kvmod.KVCacheDtype.BF16, torch.bfloat16),
                         ("int8", kvmod.KVCacheDtype.INT8, torch.float32)):
        cfg = kvmod.KVCacheConfig(max_seq_len=S, num_layers=L, num_heads=H, head_dim=D, dtype=dt)
        cache = kvmod.KVCache(cfg)
        gg = torch.Generator().manual_seed(7)
        ks = torch.randn(S, 1, 1, H, D, generator=gg)
        vs = torch.randn(S, 1, 1, H, D, generator=gg)
        qs = torch.randn(S, 1, H, 1, D, generator=gg)
        outs = []
        for pos in range(S):
            o = kvmod.attention_with_kv_cache(qs[pos].to(qdt), cache, 1, ks[pos].to(qdt), vs[pos].to(qdt), pos)
            outs.append(o.float())
        at[f"k_{tag}"] = ks.numpy()
        at[f"v_{tag}"] = vs.numpy()
        at[f"q_{tag}"] = qs.numpy()
        at[f"out_{tag}"] = torch.stack(outs).numpy()
    # head_dim 128 (the only head size of the BASELINE shapes; what the CUDA kernel implements), bf16 cache
    L, H, D, S = 2, 4, 128, 40
    cfg = kvmod.KVCacheConfig(max_seq_len=S, num_layers=L, num_heads=H, head_dim=D, dtype=kvmod.KVCacheDtype.BF16)
    cache = kvmod.KVCache(cfg)
    gg = torch.Generator().manual_seed(11)
    ks = torch.randn(S, 1, 1, H, D, generator=gg).to(torch.bfloat16)
    vs = torch.randn(S, 1, 1, H, D, generator=gg).to(torch.bfloat16)
    qs = torch.randn(S, 1, H, 1, D, generator=gg).to(torch.bfloat16)
    outs = [kvmod.attention_with_kv_cache(qs[pos], cache, 0, ks[pos], vs[pos], pos).float() for pos in range(S)]
    at["k_d128"] = ks.float().numpy()
    at["v_d128"] = vs.float().numpy()
    at["q_d128"] = qs.float().numpy()
    at["out_d128"] = torch.stack(outs).numpy()
    mem = kvmod.compute_kv_cache_memory({"num_layers": 30, "num_heads": 5, "head_dim": 128}, 4096, kvmod.KVCacheDtype.BF16)
    at["mem_total_mb"] = np.array(mem["total_mb"])
    at["mem_per_token_kb"] = np.array(mem["per_token_kb"])
    np.savez_compressed(OUT / "attention.npz", **at)

    # ---- 6. kernels/native.py numpy helpers (LSB-sequential)
    nv = {}
    w = torch.randint(-1, 2, (16, 64), generator=g).float().numpy()
    nv["w"] = w.astype(np.int8)
    nv["packed"] = native.pack_weights(w)
    nv["unpacked"] = native.unpack_weights(nv["packed"]).astype(np.int8)
    np.savez_compressed(OUT / "native.npz", **nv)

    for f in sorted(OUT.glob("*.npz")):
        print(f"{f.name}: {f.stat().st_size} bytes")


if __name__ == "__main__":
    main()
