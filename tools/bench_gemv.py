#!/usr/bin/env python
"""Kernel-level ternary GEMV bandwidth (BASELINE.json metric "ternary GEMV HBM GB/s vs peak"), in the style of the
reference's scripts/benchmark_kernels.py:28-66: bytes = N*K/4 + M*K + out per call, but the calls rotate over > L2 worth
of weight copies (wf_bench_gemv).  One JSON line per (shape, M, variant)."""
import argparse, ctypes as C, json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    sys.path.insert(0, p)
from wf_bitnet import _lib
from wf_bitnet._lib import check, lib

SHAPES = {   # SURVEY Appendix B; fused matrices as the engine stores them
    "2b_qkv": (3840, 2560), "2b_o": (2560, 2560), "2b_gateup": (13824, 2560), "2b_down": (2560, 6912),
    "8b_qkv": (6144, 4096), "8b_gateup": (28672, 4096), "8b_down": (4096, 14336),
    "70b_qkv": (10240, 8192), "70b_gateup": (57344, 8192), "70b_down": (8192, 28672),
}

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default=",".join(SHAPES))
    ap.add_argument("--M", default="1")
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--variants", default="0,1,2,3,5", help="wf_bench_gemv flag words")
    a = ap.parse_args()
    _lib.require_gpu()
    peak = 6650.0
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        peak = float(json.loads(f.read_text())["hbm_gbs"])
    for name in a.shapes.split(","):
        N, K = SHAPES[name]
        for M in [int(x) for x in a.M.split(",")]:
            for fl in [int(x) for x in a.variants.split(",")]:
                ms = C.c_float()
                check(lib().wf_bench_gemv(N, K, M, a.iters, fl, C.byref(ms)))
                out_b = (2 if fl & 4 else 4) * M * N
                byt = N * K // 4 + M * K + out_b
                gbs = byt / (ms.value * 1e-3) / 1e9
                print(json.dumps({"shape": name, "N": N, "K": K, "M": M, "pdl": bool(fl & 1), "stream_only": bool(fl & 2), "fused_epilogue": bool(fl & 4), "mma_path": bool(fl & 8), "cw": 16 if fl & 16 else (8 if fl & 32 else 0), "l2_ahead": bool(fl & 64), "grid_mult": ((fl >> 16) & 15) or (2 if fl & 128 else 1), "stages": (fl >> 8) & 255,
                                  "us_per_launch": round(ms.value * 1e3, 3), "bytes": byt, "GBs": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 4)}), flush=True)

if __name__ == "__main__":
    main()
