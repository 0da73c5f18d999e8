#!/usr/bin/env python
"""bench.py — decode throughput of the BitNet-2B-4T-shaped hot path on B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload bs64|bs1]

A "step" is one decode step (one new token for every sequence of the batch) of the
BitNet-2B-4T-shaped model: 30 layers x (rmsnorm+int8 quant, fused q|k|v ternary linear, RoPE + paged KV
append, paged flash-decode attention, sub-norm+quant, o linear, rmsnorm+quant, fused gate|up linear with
relu^2*up, sub-norm+quant, down linear) + final norm + tied bf16 LM head + greedy argmax.
Default workload = BASELINE.json configs[1]: batch 64, every sequence holding 4096 tokens of paged KV.
Weights are random-init ternary (synthetic), KV contents random.

N > 1: the 2B model has 5 KV heads and does not shard over 2/4/8 GPUs -> replicas only:
every rank runs its own batch of 64 (weak scaling, no data-path collective).

--impl reference: the reference's own CPU implementation of the same step (the unmodified AVX2 I2_S
kernel from oracle/_ref under OpenMP + torch CPU attention / LM head), on the host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "wrinklefree-inference-engine_b200"
for p in (str(ROOT), str(PKG)):
    if p not in sys.path:
        sys.path.insert(0, p)

if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":   # keeps "NCCL version ..." off stdout: rank 0 prints ONE JSON line
    os.environ["NCCL_DEBUG"] = "WARN"

# stdout carries exactly ONE JSON line: anything a library writes to fd 1 (NCCL prints its version banner there when torch
# initialises the communicator) is sent to stderr, and the JSON line goes to a private duplicate of the original stdout.
sys.stdout.flush()
_JSON_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(obj) -> None:
    _JSON_OUT.write(json.dumps(obj) + "\n")
    _JSON_OUT.flush()


import numpy as np
import torch

# BitNet-2B-4T shape: hidden 2560, 30 layers, 20/5 heads x 128, ffn 6912, vocab 128256
SHAPE = dict(hidden=2560, layers=30, heads=20, kv_heads=5, head_dim=128, ffn=6912, vocab=128256)
LAYER_WEIGHT_BYTES = 17_367_040          # packed ternary bytes per layer
KV_BYTES_PER_TOKEN_LAYER = 2 * 5 * 128 * 2
LM_HEAD_BYTES = 128256 * 2560 * 2
METRIC = "decode tok/s, BitNet-2B-4T-shaped, batch 64, 4K paged KV, 1 B200 (replicas for N>1)"

# tensor-parallel workloads (BASELINE.json configs[3], configs[4]): N GPUs = TP degree, strong scaling
TP_WORKLOADS = {
    "8b_bs1": dict(shape="llama3_8b", batch=1, ctx=512, name="BASELINE.json configs[3]: Llama-3-8B-shaped 1.58-bit model, random-init, batch=1 decode, 512-token context, TP = n_gpus"),
    "8b_bs32": dict(shape="llama3_8b", batch=32, ctx=2048, name="BASELINE.json configs[3]: Llama-3-8B-shaped 1.58-bit model, random-init, batch=32 decode, 2K-token paged KV, TP = n_gpus"),
    "70b_bs16": dict(shape="llama3_70b", batch=16, ctx=8192, name="BASELINE.json configs[4]: Llama-3-70B-shaped 1.58-bit model, random-init, batch=16 decode, 8K-token paged KV, TP = n_gpus"),
}


def run_gpu_tp(args):
    """Tensor-parallel decode: one rank per GPU, the model sharded over all ranks (wf_bitnet.tp), strong scaling."""
    import torch.distributed as dist
    from wf_bitnet import _lib, tp
    from wf_bitnet.engine import Engine, EngineConfig

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = TP_WORKLOADS[args.workload]
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.require_gpu()
    batch, ctx = wl["batch"], wl["ctx"]
    K, W = args.steps, max(3, args.warmup)
    page = 16
    extra = ((2 * (W + K) + 16 + page - 1) // page) * page + page
    cfg = getattr(EngineConfig, wl["shape"])(max_seq=ctx + extra, page_size=page, max_batch=batch)
    eng = Engine(cfg, tp_rank=rank, tp_size=world)
    eng.init_synthetic(4321)                      # same seed on every rank: replicated tensors agree, shards differ by rank
    if world > 1:
        tp.attach(eng)
    for slot in range(batch):
        eng.set_context(slot, ctx, fill_random=True)
    tokens = np.random.default_rng(0).integers(0, cfg.vocab, size=batch).astype(np.int32)
    ext = torch.cuda.ExternalStream(eng.stream)

    def barrier():
        eng.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    tokens = eng.decode(tokens)
    for _ in range(W):
        eng.decode_device(batch)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(K):
        eng.decode_device(batch)
    e1.record(ext)
    e1.synchronize()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = eng.launches_per_step()
    for slot in range(batch):
        eng.set_context(slot, ctx, fill_random=False)
    barrier()
    for _ in range(W):
        tokens = eng.decode(tokens)
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        tokens = eng.decode(tokens)
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()
    eng.close()
    t = torch.tensor([ms_total, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = float(t[0]), float(t[1])
    if rank == 0:
        peak, peak_src = measured_peaks()
        hd, ffn, hid, L = cfg.head_dim, cfg.ffn, cfg.hidden, cfg.layers
        w_bytes = L * ((cfg.heads + 2 * cfg.kv_heads) * hd * hid + hid * cfg.heads * hd + 3 * ffn * hid) // 4
        kv_bytes = 2 * cfg.kv_heads * hd * 2 * L * batch * (ctx + 1)
        lm_bytes = cfg.vocab * hid * 2
        per_gpu = (w_bytes + kv_bytes + lm_bytes) / world
        ms_step = ms_total / K
        achieved = per_gpu / (ms_step * 1e-3) / 1e9
        line = {
            "metric": f"decode tok/s, {wl['shape']}-shaped 1.58-bit model, batch {batch}, {ctx}-token KV, TP={world}", "value": batch * K / (ms_total * 1e-3),
            "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "i8 x i2 -> i32, bf16 glue", "data": "synthetic",
            "config": {"workload": wl["name"], "batch": batch, "ctx": ctx, "page_size": page, "tp": world,
                       "l2": "inputs larger than L2: every step streams the rank's weight shard, vocab slice and KV heads",
                       "multi_gpu": "tensor parallel: q|k|v, gate|up by output rows, o/down by input columns, KV by head; all-gather/all-reduce = NVLink peer stores/REDs fused into the attention and linear kernels (no NCCL on the data path)",
                       "graph": "CUDA graph replay, 1 graph launch per step per rank"},
            "clocks": clocks,
            "e2e": {"value": batch * K / (e2e_ms * 1e-3), "unit": "tok/s", "h2d_bytes_per_step": 4 * batch * world, "d2h_bytes_per_step": 4 * batch * world,
                    "call": "wf_engine_decode(host token ids) -> host greedy next ids on every rank, sync every step"},
            "gpu_launches": launches * K * world, "gpu_launches_per_step": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "kernel": "whole decode step, per GPU (weight shard + vocab slice + KV heads streamed once per step)",
                         "algorithmic_bytes_per_launch": per_gpu, "avg_launch_ms": ms_step},
            "cpu_baseline": None,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def measured_peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            d = json.loads(f.read_text())
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel):
    """per-launch dram bytes of the dominant kernel from the committed ncu summary, or None."""
    f = ROOT / "profiles" / "roofline_traffic.json"
    if f.exists():
        try:
            return json.loads(f.read_text()).get(kernel)
        except Exception:
            return None
    return None


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.proc:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    sm.append(float(f[0])); mx.append(float(f[1]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        finally:
            try:
                os.unlink(self.path)
            except OSError:
                pass
        if sm:
            out["sm_mhz"] = statistics.median(sm)
            out["sm_max_mhz"] = max(mx)
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference_step(batch, ctx, threads, reps):
    """One bounded sample of the reference CPU step: ONE of the 30 layers (7 ternary GEMMs through the
    reference's AVX2 kernel + GQA decode attention over `ctx` tokens for every sequence, torch CPU as the
    reference does) and the bf16 LM head.  Returns (seconds per full step extrapolated x30, description)."""
    import torch.nn.functional as F
    import oracle as O
    use_ref = O.ref_available()
    torch.set_num_threads(threads)
    if use_ref:
        O.ref().wfref_set_threads(threads)
    rng = np.random.default_rng(1234)
    H, HKV, D, HID, FFN, V = 20, 5, 128, 2560, 6912, 128256
    shapes = {"q": (H * D, HID), "k": (HKV * D, HID), "v": (HKV * D, HID), "o": (HID, H * D), "gate": (FFN, HID), "up": (FFN, HID), "down": (HID, FFN)}
    from wf_bitnet.synth import pack_i2s_numpy
    packed = {n: pack_i2s_numpy(rng.integers(-1, 2, size=s, dtype=np.int8)) for n, s in shapes.items()}
    acts = {K: rng.integers(-127, 128, size=(batch, K), dtype=np.int8) for K in (HID, FFN)}
    kc = torch.randn(batch, HKV, ctx, D).to(torch.bfloat16)
    vc = torch.randn(batch, HKV, ctx, D).to(torch.bfloat16)
    q = torch.randn(batch, H, 1, D).to(torch.bfloat16)
    E = (0.02 * torch.randn(V, HID)).to(torch.bfloat16)
    hcpu = torch.randn(batch, HID).to(torch.bfloat16)
    gemm = (lambda p, a: O.ref_gemm(p, a, 1.0)) if use_ref else (lambda p, a: O.gemm(p, a, 1.0))

    def layer():
        for n, (N, K) in shapes.items():
            gemm(packed[n], acts[K])
        F.scaled_dot_product_attention(q, kc, vc, enable_gqa=True)

    layer()  # warm
    t0 = time.perf_counter()
    for _ in range(reps):
        layer()
    t_layer = (time.perf_counter() - t0) / reps
    F.linear(hcpu, E)
    t0 = time.perf_counter()
    for _ in range(max(1, reps // 2)):
        F.linear(hcpu, E)
    t_lm = (time.perf_counter() - t0) / max(1, reps // 2)
    desc = (f"1 of 30 layers (7 ternary GEMMs, batch {batch}, {'unmodified reference AVX2 ggml_vec_dot_i2_i8_s + OpenMP' if use_ref else 'scalar C oracle port'}; "
            f"GQA SDPA over {ctx}-token KV for {batch} sequences, torch CPU) x30 + bf16 LM head, {reps} repetitions")
    return t_layer * 30 + t_lm, desc, ("reference" if use_ref else "port")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    batch, ctx = (64, 4096) if args.workload == "bs64" else (1, 512)
    cores = os.cpu_count() or 1
    times = []
    desc = kind = None
    for _ in range(args.warmup):
        cpu_reference_step(batch, ctx, cores, 1)
        break   # one warm-up sample is enough on the CPU; keeps the arm within minutes
    for _ in range(max(1, min(args.steps, 5))):
        t, desc, kind = cpu_reference_step(batch, ctx, cores, 1)
        times.append(t)
    t_step = statistics.median(times)
    val = batch / t_step
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tok/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i8 x i2 -> i32, bf16 glue",
        "data": "synthetic", "config": {"workload": workload_name(args.workload), "batch": batch, "ctx": ctx},
        "cpu_baseline": {"value": val, "unit": "tok/s", "cores": cores, "kind": kind, "sample": desc},
        "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)
    return 0


def workload_name(w):
    if w == "bs64":
        return "BASELINE.json configs[1]: BitNet-2B-4T-shaped batched decode, batch=64, 4K-token paged KV cache, single B200"
    return "BASELINE.json configs[0] shape on GPU: BitNet-2B-4T-shaped batch=1 decode at 512-token context"


# ------------------------------------------------------------------------------------------- GPU arm
def run_gpu(args):
    import torch.distributed as dist
    from wf_bitnet import _lib
    from wf_bitnet.engine import Engine, EngineConfig

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    _lib.require_gpu()

    batch, ctx = (64, 4096) if args.workload == "bs64" else (1, 512)
    K, W = args.steps, max(3, args.warmup)
    page = 16
    extra = ((W + K + 8 + page - 1) // page) * page + page
    cfg = EngineConfig(**SHAPE, max_seq=ctx + extra, page_size=page, max_batch=batch)
    eng = Engine(cfg)
    eng.init_synthetic(1234 + rank)
    for slot in range(batch):
        eng.set_context(slot, ctx, fill_random=True)
    rng = np.random.default_rng(rank)
    tokens = rng.integers(0, SHAPE["vocab"], size=batch).astype(np.int32)
    ext = torch.cuda.ExternalStream(eng.stream)

    def barrier():
        eng.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # ---------------- value: device-resident inputs, CUDA-graph replay, events on the engine stream
    eng.decode(tokens)                      # seeds the device token buffer (and captures the host-token graph)
    from wf_bitnet._lib import lib, check
    check(lib().wf_copy_d2d(lib().wf_engine_buffer(eng._h, b"tokens"), lib().wf_engine_buffer(eng._h, b"next"), 4 * batch, None))
    torch.cuda.synchronize()
    for _ in range(W):
        eng.decode_device(batch)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(K):
        eng.decode_device(batch)
    e1.record(ext)
    e1.synchronize()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = eng.launches_per_step()

    # ---------------- e2e: host tokens in, host tokens out, every step (the public call)
    for slot in range(batch):
        eng.set_context(slot, ctx, fill_random=False)
    for _ in range(W):
        tokens = eng.decode(tokens)
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        tokens = eng.decode(tokens)
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()

    # ---------------- roofline of the dominant kernel, measured live with CUDA events on the same stream
    for slot in range(batch):
        eng.set_context(slot, ctx, fill_random=False)
    peak, peak_src = measured_peaks()
    roof_iters = 60 if batch > 1 else 300
    if args.workload == "bs64":
        ms_attn = eng.time_kernel("attn", batch, roof_iters)
        alg_bytes = KV_BYTES_PER_TOKEN_LAYER * batch * (ctx + 1) + 2 * batch * 20 * 128 * 2   # K+V of every cached token + q + out
        dom = {"kernel": "paged_decode_attn_kernel (+attn_merge_kernel)", "ms": ms_attn, "bytes": alg_bytes,
               "launches_per_step": 30, "share_of_step": 30 * ms_attn / (ms_total / K)}
        traffic = ncu_traffic("paged_decode_attn_kernel")
    else:
        ms_lin = eng.time_kernel("linears", batch, roof_iters)
        alg_bytes = LAYER_WEIGHT_BYTES + 4 * (2560 + 6912) + 2 * (3840 + 2560 + 6912 + 2560)
        dom = {"kernel": "ternary_gemv_stream_kernel (q|k|v, o, gate|up, down of one layer)", "ms": ms_lin, "bytes": alg_bytes,
               "launches_per_step": 120, "share_of_step": 30 * ms_lin / (ms_total / K)}
        traffic = ncu_traffic("ternary_gemv_stream_kernel")
    achieved = dom["bytes"] / (dom["ms"] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": peak_src, "kernel": dom["kernel"], "algorithmic_bytes_per_launch": dom["bytes"], "avg_launch_ms": dom["ms"],
                "share_of_step": dom["share_of_step"]}

    # ---------------- per-component picture of the same step (explains `value`)
    breakdown = {}
    try:
        for comp in ("attn", "qkv", "o", "gateup", "down", "lm_head", "norm"):
            breakdown[comp + "_ms"] = round(eng.time_kernel(comp, batch, 30 if comp != "lm_head" else 10), 5)
        lin_ms = breakdown["qkv_ms"] + breakdown["o_ms"] + breakdown["gateup_ms"] + breakdown["down_ms"]
        breakdown["ternary_linears_GBs_per_layer"] = round(LAYER_WEIGHT_BYTES / (lin_ms * 1e-3) / 1e9, 1)
        breakdown["ternary_linears_frac_of_hbm_peak"] = round(LAYER_WEIGHT_BYTES / (lin_ms * 1e-3) / 1e9 / peak, 4)
        breakdown["lm_head_GBs"] = round(LM_HEAD_BYTES / (breakdown["lm_head_ms"] * 1e-3) / 1e9, 1)
    except Exception as ex:  # noqa: BLE001
        breakdown["error"] = str(ex)
    eng.close()

    # ---------------- batch-1 decode + GEMV roofline (the other half of BASELINE's metric), N = 1 only
    aux = {}
    if args.workload == "bs64" and world == 1 and not args.no_aux:
        try:
            c1 = EngineConfig(**SHAPE, max_seq=1024, page_size=page, max_batch=1)
            e1g = Engine(c1)
            e1g.init_synthetic(99)
            e1g.set_context(0, 512, fill_random=True)
            e1g.decode([1])
            for _ in range(5):
                e1g.decode_device(1)
            e1g.sync()
            ext1 = torch.cuda.ExternalStream(e1g.stream)
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n1 = 128
            a0.record(ext1)
            for _ in range(n1):
                e1g.decode_device(1)
            a1.record(ext1)
            a1.synchronize()
            ms1 = a0.elapsed_time(a1) / n1
            lin1 = e1g.time_kernel("linears", 1, 300)
            aux = {"bs1_ctx512_tok_s": 1e3 / ms1, "bs1_ms_per_token": ms1, "bs1_gpu_launches_per_token": e1g.launches_per_step(),
                   "bs1_gemv_GBs": LAYER_WEIGHT_BYTES / (lin1 * 1e-3) / 1e9, "bs1_gemv_frac_of_hbm_peak": LAYER_WEIGHT_BYTES / (lin1 * 1e-3) / 1e9 / peak,
                   "bs1_gemv_ms_per_layer": lin1,
                   "bs1_step_weight_stream_frac": (30 * LAYER_WEIGHT_BYTES + LM_HEAD_BYTES) / (ms1 * 1e-3) / 1e9 / peak}
            e1g.close()
        except Exception as ex:  # noqa: BLE001
            aux = {"error": str(ex)}

    # ---------------- the same step with the int8 KV option (the reference's KVCacheDtype.INT8): NOT the headline (other cache
    # precision), reported beside it because it halves the attention bytes of this configuration
    if args.workload == "bs64" and world == 1 and not args.no_aux:
        try:
            c8 = EngineConfig(**SHAPE, max_seq=ctx + 64, page_size=page, max_batch=batch, kv_dtype="int8")
            e8 = Engine(c8)
            e8.init_synthetic(1234)
            for slot in range(batch):
                e8.set_context(slot, ctx, fill_random=True)
            for _ in range(5):
                e8.decode_device(batch)
            e8.sync()
            ext8 = torch.cuda.ExternalStream(e8.stream)
            b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n8 = 30
            b0.record(ext8)
            for _ in range(n8):
                e8.decode_device(batch)
            b1.record(ext8)
            b1.synchronize()
            ms8 = b0.elapsed_time(b1) / n8
            attn8 = e8.time_kernel("attn", batch, 60)
            bytes8 = batch * (ctx + 36) * 2 * 5 * (128 + 4)
            aux["int8_kv"] = {"note": "same workload with kv_dtype=int8 (int8 codes + one fp32 scale per token and kv head); not the headline precision",
                              "ms_per_step": ms8, "tok_s": batch / ms8 * 1e3, "attn_ms_per_layer": attn8, "attn_algorithmic_bytes": bytes8,
                              "attn_GBs": bytes8 / (attn8 * 1e-3) / 1e9, "attn_frac_of_hbm_peak": bytes8 / (attn8 * 1e-3) / 1e9 / peak,
                              "kv_pool_bytes_ratio_vs_bf16": (128 + 4) / 256}
            e8.close()
        except Exception as ex:  # noqa: BLE001
            aux["int8_kv"] = {"error": str(ex)}

    # ---------------- kernel-level GEMV bandwidth (the third part of BASELINE's metric): M = 1, > L2 worth of weight copies
    if world == 1 and not args.no_aux:
        try:
            import ctypes as C
            rows = []
            for name, (N_, K_) in (("2b_gate|up", (13824, 2560)), ("2b_down", (2560, 6912)), ("8b_gate|up", (28672, 4096)), ("70b_gate|up", (57344, 8192)),
                                   ("70b_down", (8192, 28672))):
                ms = C.c_float()
                check(lib().wf_bench_gemv(N_, K_, 1, 200, 1, C.byref(ms)))
                byt = N_ * K_ // 4 + K_ + 4 * N_
                gbs = byt / (ms.value * 1e-3) / 1e9
                rows.append({"shape": name, "N": N_, "K": K_, "us_per_launch": round(ms.value * 1e3, 2), "algorithmic_bytes": byt, "GBs": round(gbs, 1),
                             "frac_of_hbm_peak": round(gbs / peak, 4)})
            aux["gemv_kernel_M1"] = {"method": "wf_bench_gemv: 200 graph-captured launches of ternary_gemv_stream_kernel rotating over > 320 MB of weight copies, "
                                               "PDL between launches, bytes = N*K/4 + K + 4*N (scripts/benchmark_kernels.py:61-66)", "rows": rows}
        except Exception as ex:  # noqa: BLE001
            aux["gemv_kernel_M1"] = {"error": str(ex)}

    # ---------------- CPU baseline beside it (rank 0, N = 1 only; bounded sample)
    cpu_baseline = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        t_step, desc, kind = cpu_reference_step(batch, ctx, cores, 2)
        cpu_baseline = {"value": batch / t_step, "unit": "tok/s", "cores": cores, "kind": kind, "sample": desc}

    # ---------------- max over ranks
    t = torch.tensor([ms_total, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = float(t[0]), float(t[1])
    if rank == 0:
        line = {
            "metric": METRIC, "value": world * batch * K / (ms_total * 1e-3), "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "i8 x i2 -> i32, bf16 glue", "data": "synthetic",
            "config": {"workload": workload_name(args.workload), "batch_per_gpu": batch, "ctx": ctx, "page_size": page,
                       "l2": "inputs larger than L2: every step streams 521 MB of ternary weights, 657 MB of LM head and the KV cache (20 GB at bs64)",
                       "multi_gpu": "replicas only (2B: 5 KV heads do not shard); no data-path collective", "graph": "CUDA graph replay, 1 graph launch per step"},
            "clocks": clocks,
            "e2e": {"value": world * batch * K / (e2e_ms * 1e-3), "unit": "tok/s", "h2d_bytes_per_step": 4 * batch, "d2h_bytes_per_step": 4 * batch,
                    "call": "wf_engine_decode(host token ids) -> host greedy next ids, sync every step"},
            "gpu_launches": launches * K,
            "gpu_launches_per_step": launches,
            "roofline": roofline,
            "cpu_baseline": cpu_baseline,
            "breakdown_ms": breakdown,
            "aux": aux,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="bs64", choices=["bs64", "bs1"] + sorted(TP_WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-aux", action="store_true", help="skip the batch-1 aux measurement")
    args = ap.parse_args()
    if args.workload in TP_WORKLOADS:
        if args.impl == "reference":
            emit({"impl": "reference", "unavailable": "the reference arm is defined for the 2B workloads (bs64 / bs1) only"})
            return 0
        return run_gpu_tp(args)
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
