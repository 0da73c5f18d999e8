#!/usr/bin/env python
"""BASELINE config 3: BitNet-2B-4T-shaped prefill, 8 x 2048 tokens (M = 16384) through the ternary linears of one
layer on the tcgen05 int8 path.  Prints one JSON line: int8 TOP/s per linear and per layer, next to the int8
tensor-pipe peak measured on the same box with torch._int_mm (cuBLASLt) at 8192^3."""
import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters


def clocks_under_load(fn, seconds=2.0):
    """SM clock (MHz) and board power (W) sampled with NVML while `fn` runs back to back: median over the loaded samples."""
    import threading
    import time
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(torch.cuda.current_device())
    stop, sm, pw = [False], [], []

    def sampler():
        while not stop[0]:
            sm.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
            pw.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
            time.sleep(0.02)
    t = threading.Thread(target=sampler)
    t0 = time.time()
    fn()
    torch.cuda.synchronize()
    t.start()
    n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < seconds:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    e1.record()
    e1.synchronize()
    stop[0] = True
    t.join()
    k = len(sm) // 2
    return {"sm_mhz_median": sorted(sm[k:])[len(sm[k:]) // 2] if sm[k:] else None, "power_w_median": round(sorted(pw[k:])[len(pw[k:]) // 2], 1) if pw[k:] else None,
            "ms_sustained": round(e0.elapsed_time(e1) / max(n, 1), 4), "samples": len(sm)}


def measure(tokens=16384, iters=10, algo=3, shape="2b", only="", layout="tile", clocks=False):
    """-> dict: per-linear and per-layer int8 TOP/s of the ternary GEMM at M = tokens, next to torch._int_mm (cuBLASLt) 8192^3 on the same GPU."""
    from wf_bitnet import _lib, ops
    _lib.require_gpu()
    M = tokens
    hid, ffn, qkv = (2560, 6912, 3840) if shape == "2b" else (4096, 14336, 6144)
    shapes = {"qkv": (qkv, hid), "o": (hid, hid), "gate_up": (2 * ffn, hid), "down": (hid, ffn)}
    g = torch.Generator(device="cuda").manual_seed(0)
    out = {"tokens": M, "shape": shape, "algo": algo, "kernel": "ternary_gemm_tc_kernel (tcgen05 kind::i8)" if algo == 3 else f"algo {algo}"}
    total_ms, total_ops = 0.0, 0.0
    for name, (n, k) in shapes.items():
        if only and name != only:
            continue
        tile = torch.randint(0, 256, (ops.tile_bytes(n, k),), generator=g, device="cuda", dtype=torch.uint8)
        if layout == "i2s":
            tile = tile[: n * k // 4].reshape(n, k // 4).contiguous()
        q = torch.randint(-127, 128, (M, k), generator=g, device="cuda", dtype=torch.int8)
        rs = q.sum(1, dtype=torch.int32)
        sc = torch.rand(M, generator=g, device="cuda") * 0.05 + 0.01
        ws = torch.tensor([1.7], device="cuda")
        fn = lambda: ops.ternary_linear(tile, n, k, q, mode="linear", rowsum=rs, act_scale=sc, w_scale=ws, out_dtype=torch.bfloat16, algo=algo, layout=layout)  # noqa: E731
        ms = timed(fn, iters)
        opsn = 2.0 * M * n * k
        out[name] = {"N": n, "K": k, "ms": round(ms, 4), "TOPs": round(opsn / ms / 1e9, 1)}
        if clocks:
            out[name]["under_load"] = clocks_under_load(fn)
            x8 = torch.randint(-127, 128, (M, k), device="cuda", dtype=torch.int8)
            w8 = torch.randint(-1, 2, (n, k), device="cuda", dtype=torch.int8).t()
            out[name]["cublaslt_under_load"] = clocks_under_load(lambda: torch._int_mm(x8, w8))
            del x8, w8
        total_ms += ms
        total_ops += opsn
        del tile, q
    out["layer_ms"] = round(total_ms, 4)
    out["layer_TOPs"] = round(total_ops / total_ms / 1e9, 1)
    out["prefill_30_layers_ms"] = round(30 * total_ms, 2)
    out["prefill_tok_s_linears_only"] = round(M / (30 * total_ms * 1e-3), 0)
    if only:
        return out
    a = torch.randint(-127, 128, (8192, 8192), device="cuda", dtype=torch.int8)
    b = torch.randint(-127, 128, (8192, 8192), device="cuda", dtype=torch.int8).t()
    try:
        ms = timed(lambda: torch._int_mm(a, b), 10)
        out["int8_peak_TOPs_cublaslt_8192"] = round(2 * 8192.0 ** 3 / ms / 1e9, 1)
        out["frac_of_int8_peak"] = round(out["layer_TOPs"] / out["int8_peak_TOPs_cublaslt_8192"], 4)
        # the same contraction shapes through the library (int8 x int8 -> int32, no unpack, no epilogue): what "matching cuBLASLt" means here
        lib_ms = 0.0
        for name, (n, k) in shapes.items():
            x = torch.randint(-127, 128, (M, k), device="cuda", dtype=torch.int8)
            w = torch.randint(-1, 2, (n, k), device="cuda", dtype=torch.int8).t()
            lib_ms += timed(lambda: torch._int_mm(x, w), 5)
            del x, w
        out["cublaslt_same_shapes_layer_ms"] = round(lib_ms, 4)
        out["cublaslt_same_shapes_layer_TOPs"] = round(total_ops / lib_ms / 1e9, 1)
        out["frac_of_cublaslt_same_shapes"] = round(lib_ms / total_ms, 4)
    except Exception as ex:  # noqa: BLE001
        out["int8_peak_error"] = str(ex)[:200]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=16384)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--algo", type=int, default=3)
    ap.add_argument("--shape", default="2b", choices=["2b", "8b"])
    ap.add_argument("--only", default="", help="time just this linear (qkv|o|gate_up|down)")
    ap.add_argument("--layout", default="tile", choices=["tile", "i2s"])
    ap.add_argument("--clocks", action="store_true", help="also run each linear (and the same shape through cuBLASLt) back to back for 2 s and report SM clock / power under load")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    print(json.dumps(measure(args.tokens, args.iters, args.algo, args.shape, args.only, args.layout, args.clocks)))


if __name__ == "__main__":
    main()
