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


# ------------------------------------------------------------------------------------------- tensor-parallel suite (aux.tp)
def _tp_cfg(wl, K, W, **over):
    from wf_bitnet.engine import EngineConfig
    page = 16
    extra = ((2 * (W + K) + 16 + page - 1) // page) * page + page
    cfg = getattr(EngineConfig, wl["shape"])(max_seq=wl["ctx"] + extra, page_size=page, max_batch=wl["batch"])
    for k, v in over.items():
        setattr(cfg, k, v)
    return cfg


def _tp_time(cfg, wl, K, W, rank, world, dist):
    """device-timed ms per decode step of one engine (TP = world when world > 1: max over ranks), CUDA-graph replay."""
    from wf_bitnet import tp
    from wf_bitnet.engine import Engine
    eng = Engine(cfg, tp_rank=rank if world > 1 else 0, tp_size=world)
    eng.init_synthetic(4321)
    if world > 1:
        tp.attach(eng)
    for slot in range(wl["batch"]):
        eng.set_context(slot, wl["ctx"], fill_random=True)
    tokens = np.random.default_rng(0).integers(0, cfg.vocab, size=wl["batch"]).astype(np.int32)
    ext = torch.cuda.ExternalStream(eng.stream)
    eng.decode(tokens)
    for _ in range(W):
        eng.decode_device(wl["batch"])
    eng.sync()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(K):
        eng.decode_device(wl["batch"])
    e1.record(ext)
    e1.synchronize()
    eng.sync()
    ms = e0.elapsed_time(e1) / K
    launches = eng.launches_per_step()
    eng.close()
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    return ms, launches


def _tp_parity(rank, world, dist):
    """TP = world vs TP = 1 at the 8B LAYER shape (hidden 4096, 32/8 heads, ffn 14336; 2 layers, vocab 8192): every rank builds the
    same seeded model and keeps its shard, rank 0 also runs the unsharded engine; the final normalised hidden state and the greedy
    tokens must be BIT-identical at batch 1 (GEMV phases, all-reduce by RED), 16 / 20 (streaming batched GEMM, reduce-scatter + pull) and 32 (one-shot / tcgen05 forms)."""
    from wf_bitnet import tp
    from wf_bitnet.engine import Engine, EngineConfig
    from wf_bitnet.synth import make_weights
    ok, detail = True, []
    cases = ((1, 5, 40), (16, 4, 24), (20, 3, 24), (32, 4, 24))   # 16 / 20: the streaming batched GEMM under the exchanges
    if os.environ.get("WF_TP_PARITY_BATCHES"):   # diagnostics: other batch sizes
        cases = tuple((int(b), 4, 24) for b in os.environ["WF_TP_PARITY_BATCHES"].split(","))
    for batch, steps, ctx in cases:
        cfg = EngineConfig(hidden=4096, layers=2, heads=32, kv_heads=8, head_dim=128, ffn=14336, vocab=8192, max_seq=128, page_size=16, max_batch=batch)
        weights = make_weights(cfg, seed=11 + batch)
        eng = Engine(cfg, tp_rank=rank, tp_size=world)
        eng.load_weights(tp.shard_weights(weights, rank, world))
        tp.attach(eng)
        ref = None
        if rank == 0:
            ref = Engine(cfg)
            ref.load_weights(weights)
        for e in ([eng, ref] if ref else [eng]):
            for slot in range(batch):
                e.set_context(slot, ctx + slot % 7, fill_random=False)
        toks = np.random.default_rng(batch).integers(0, cfg.vocab, size=batch).astype(np.int32)
        t_tp, t_ref = toks.copy(), toks.copy()
        same = True
        for _ in range(steps):
            t_tp = eng.decode(t_tp)
            if ref is not None:
                t_ref = ref.decode(t_ref)
                same = same and bool(torch.equal(eng.buffer("normed", (batch, cfg.hidden), torch.bfloat16), ref.buffer("normed", (batch, cfg.hidden), torch.bfloat16)))
                same = same and bool(np.array_equal(t_tp, t_ref))
        for _ in range(3):                       # the CUDA-graph replay path
            eng.decode_device(batch)
            if ref is not None:
                ref.decode_device(batch)
        eng.sync()
        if ref is not None:
            ref.sync()
            same = same and bool(torch.equal(eng.buffer("next", (batch,), torch.int32), ref.buffer("next", (batch,), torch.int32)))
            detail.append({"batch": batch, "steps": steps + 3, "bit_identical": same})
        flag = torch.tensor([1 if same else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = ok and bool(flag.item())
        eng.close()
        if ref is not None:
            ref.close()
        dist.barrier()
    return ok, detail


def run_tp_suite(rank, world, dist, K, W, peak):
    """BASELINE.json configs[3] / configs[4] in front of the driver: at N ranks, TP = N parity and timing of the
    8B bs1 / 8B bs32 / 70B bs16 workloads; rank 0 additionally times the TP = 1 engine and a single-GPU stand-in for one rank's
    shard (heads, KV heads, ffn and vocab divided by N: the rank's compute without the exchanges) on its own GPU, so every line
    carries speedup_vs_tp1 and the exchange cost measured on the same box in the same run."""
    out = {}
    parity, parity_detail = None, None
    if world > 1:
        try:
            parity, parity_detail = _tp_parity(rank, world, dist)
        except Exception as ex:  # noqa: BLE001
            parity, parity_detail = False, f"error: {ex}"
    for name, wl in TP_WORKLOADS.items():
        row = {"workload": wl["name"].replace("TP = n_gpus", f"TP = {world}")}
        try:
            cfg = _tp_cfg(wl, K, W)
            hd, ffn, hid, L = cfg.head_dim, cfg.ffn, cfg.hidden, cfg.layers
            w_bytes = L * ((cfg.heads + 2 * cfg.kv_heads) * hd * hid + hid * cfg.heads * hd + 3 * ffn * hid) // 4
            kv_bytes = 2 * cfg.kv_heads * hd * 2 * L * wl["batch"] * (wl["ctx"] + 1)
            per_gpu = (w_bytes + kv_bytes + cfg.vocab * hid * 2) / world
            ms_tp = None
            if world > 1:
                ms_tp, launches = _tp_time(cfg, wl, K, W, rank, world, dist)
            tp1 = shard = None
            if rank == 0:
                tp1, l1 = _tp_time(_tp_cfg(wl, K, W), wl, K, W, 0, 1, None)
                if world > 1:
                    v_slice = (((cfg.vocab + world - 1) // world) + 127) // 128 * 128
                    shard, _ = _tp_time(_tp_cfg(wl, K, W, heads=cfg.heads // world, kv_heads=cfg.kv_heads // world, ffn=cfg.ffn // world, vocab=v_slice),
                                        wl, K, W, 0, 1, None)
            if world > 1:
                dist.barrier()
            if rank == 0:
                ms = ms_tp if world > 1 else tp1
                row.update({"tp": world, "ms_per_step": round(ms, 4), "tok_s": round(wl["batch"] / ms * 1e3, 1), "tp1_ms_per_step": round(tp1, 4),
                            "speedup_vs_tp1": round(tp1 / ms, 3), "hbm_frac_per_gpu": round(per_gpu / (ms * 1e-3) / 1e9 / peak, 4),
                            "algorithmic_bytes_per_gpu": int(per_gpu), "parity_bit_identical": parity})
                if world > 1:
                    row.update({"gpu_launches_per_step": launches, "shard_compute_ms": round(shard, 4), "exchanges_per_step": 4 * L + 1,
                                "us_per_exchange": round((ms - shard) / (4 * L + 1) * 1e3, 2)})
        except Exception as ex:  # noqa: BLE001
            row["error"] = str(ex)[:300]
            if world > 1:
                try:
                    dist.barrier()
                except Exception:  # noqa: BLE001
                    pass
        out[name] = row
    out["parity"] = {"bit_identical": parity, "shape": "8B layer shape (hidden 4096, 32/8 heads x 128, ffn 14336), 2 layers, vocab 8192; TP=N vs TP=1 on rank 0",
                     "cases": parity_detail}
    out["method"] = ("strong scaling: one model sharded over all N ranks (q|k|v, gate|up by output rows, o/down by input columns, KV by head); exchanges are NVLink peer "
                     "stores / REDs fused into the producing kernels, no NCCL on the data path; device-timed CUDA-graph replay, max over ranks; "
                     "us_per_exchange = (TP step - single-GPU stand-in for one rank's shard) / exchanges")
    return out


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
class CpuReferenceStep:
    """The reference's CPU decode step at the BitNet-2B-4T shape, batched: a REAL 30-layer step (30 distinct sets of packed
    weights, 7 ternary GEMMs per layer through the reference's unmodified AVX2 ggml_vec_dot_i2_i8_s under OpenMP, and the glue the
    CPU path really executes with torch CPU ops as scripts/direct_inference_full.py does: RMSNorm x2, sub-norm x2, per-token int8
    quantisation x4, Sigma-a correction + dequant, RoPE, GQA SDPA over the KV cache, relu^2*up, residual adds) + final norm + bf16
    LM head + argmax.  The KV cache of a layer is 2 x batch x 5 x ctx x 128 bf16 (671 MB at batch 64 / 4K); two such buffers are
    alternated over the 30 layers (each far larger than the host caches) instead of allocating 20 GB of random numbers."""

    H, HKV, D, HID, FFN, V, L = 20, 5, 128, 2560, 6912, 128256, 30

    def __init__(self, batch, ctx):
        import oracle as O
        from wf_bitnet.synth import pack_i2s_numpy
        self.O = O
        self.batch, self.ctx = batch, ctx
        self.use_ref = O.ref_available()
        rng = np.random.default_rng(1234)
        shapes = {"q": (self.H * self.D, self.HID), "k": (self.HKV * self.D, self.HID), "v": (self.HKV * self.D, self.HID), "o": (self.HID, self.H * self.D),
                  "gate": (self.FFN, self.HID), "up": (self.FFN, self.HID), "down": (self.HID, self.FFN)}
        self.layers = []
        for _ in range(self.L):
            lw = {n: pack_i2s_numpy(rng.integers(-1, 2, size=s, dtype=np.int8)) for n, s in shapes.items()}
            lw["scales"] = {n: float(rng.uniform(0.5, 2.5)) for n in shapes}
            self.layers.append(lw)
        g = torch.Generator().manual_seed(7)
        self.gam = [(1.0 + 0.1 * torch.randn(n, generator=g)).to(torch.bfloat16) for n in (self.HID, self.H * self.D, self.HID, self.FFN)]
        self.kv = [(torch.randn(batch, self.HKV, ctx + 1, self.D, generator=g).to(torch.bfloat16),
                    torch.randn(batch, self.HKV, ctx + 1, self.D, generator=g).to(torch.bfloat16)) for _ in range(2)]
        self.E = (0.02 * torch.randn(self.V, self.HID, generator=g)).to(torch.bfloat16)
        self.tokens = torch.randint(0, self.V, (batch,), generator=g)
        self.cos = torch.cos(torch.rand(self.D // 2, generator=g)).to(torch.bfloat16)
        self.sin = torch.sin(torch.rand(self.D // 2, generator=g)).to(torch.bfloat16)

    def describe(self):
        k = "unmodified reference AVX2 ggml_vec_dot_i2_i8_s + OpenMP" if self.use_ref else "scalar C oracle port"
        return (f"REAL 30-layer step, batch {self.batch}: 210 ternary GEMMs ({k}), per-token int8 quant / norms / RoPE / relu^2*up / residuals in torch CPU, "
                f"GQA SDPA over {self.ctx}-token KV for {self.batch} sequences per layer, final norm + bf16 LM head + argmax")

    def _lin(self, x, lw, name):
        """direct_inference_full.py:186-208 for every row of x (bf16 [B, K]) -> bf16 [B, N]"""
        O = self.O
        q, s, rs = O.quantize_act_i8_rows(x.float().numpy())
        sc = lw["scales"][name]
        y = O.ref_gemm(lw[name], q, sc) if self.use_ref else O.gemm(lw[name], q, sc)
        out = (torch.from_numpy(y) - sc * torch.from_numpy(rs.astype(np.float32))[:, None]) * torch.from_numpy(s)[:, None]
        return out.to(torch.bfloat16)

    @staticmethod
    def _norm(x, w, eps=1e-5):
        var = x.float().pow(2).mean(-1, keepdim=True)
        return (x * torch.rsqrt(var + eps).to(x.dtype)) * w

    def _rope(self, x):
        half = self.D // 2
        x1, x2 = x[..., :half], x[..., half:]
        return torch.cat([x1 * self.cos - x2 * self.sin, x1 * self.sin + x2 * self.cos], dim=-1)

    def step(self, threads):
        import torch.nn.functional as F
        torch.set_num_threads(threads)
        if self.use_ref:
            self.O.ref().wfref_set_threads(threads)
        B = self.batch
        h = self.E[self.tokens]
        for li, lw in enumerate(self.layers):
            kc, vc = self.kv[li & 1]
            n = self._norm(h, self.gam[0])
            q = self._rope(self._lin(n, lw, "q").view(B, self.H, self.D))
            k = self._rope(self._lin(n, lw, "k").view(B, self.HKV, self.D))
            v = self._lin(n, lw, "v").view(B, self.HKV, self.D)
            kc[:, :, self.ctx] = k
            vc[:, :, self.ctx] = v
            a = F.scaled_dot_product_attention(q.unsqueeze(2), kc, vc, enable_gqa=True).reshape(B, self.H * self.D)
            h = h + self._lin(self._norm(a, self.gam[1]), lw, "o")
            n = self._norm(h, self.gam[2])
            act = torch.relu(self._lin(n, lw, "gate")).square() * self._lin(n, lw, "up")
            h = h + self._lin(self._norm(act, self.gam[3]), lw, "down")
        logits = F.linear(self._norm(h, self.gam[0]), self.E)
        self.tokens = logits.argmax(-1)

    def timed(self, threads, reps, pin8=False):
        old_aff = None
        if pin8 and hasattr(os, "sched_setaffinity"):
            try:
                old_aff = os.sched_getaffinity(0)
                os.sched_setaffinity(0, set(sorted(old_aff)[:8]))      # the reference's operating point: taskset -c 0-7 (launch_sglang_bitnet.sh:61)
            except OSError:
                old_aff = None
        try:
            self.step(threads)                                         # warm
            ts = []
            for _ in range(reps):
                t0 = time.perf_counter()
                self.step(threads)
                ts.append(time.perf_counter() - t0)
        finally:
            if old_aff is not None:
                os.sched_setaffinity(0, old_aff)
        return statistics.median(ts)


def cpu_baseline_block(batch, ctx, reps):
    cores = os.cpu_count() or 1
    ref = CpuReferenceStep(batch, ctx)
    t_all = ref.timed(cores, reps)
    out = {"value": batch / t_all, "unit": "tok/s", "cores": cores, "kind": "reference" if ref.use_ref else "port",
           "sample": f"{ref.describe()}; median of {reps} steps after 1 warm-up", "ms_per_step": t_all * 1e3}
    if cores > 8:
        t8 = ref.timed(8, max(1, reps - 1), pin8=True)
        out["threads8"] = {"value": batch / t8, "unit": "tok/s", "cores": 8, "ms_per_step": t8 * 1e3,
                           "note": "the reference's operating point: 8 threads pinned to 8 cores (scripts/launch_sglang_bitnet.sh:61 taskset -c 0-7)"}
    return out, t_all


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    batch, ctx = (64, 4096) if args.workload == "bs64" else (1, 512)
    reps = max(1, min(args.steps, 5))
    cb, t_step = cpu_baseline_block(batch, ctx, reps)
    val = batch / t_step
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tok/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i8 x i2 -> i32, bf16 glue",
        "data": "synthetic", "config": {"workload": workload_name(args.workload), "batch": batch, "batch_per_gpu": batch, "ctx": ctx,
                                        "note": "one host, all cores: under torchrun only rank 0 runs this arm, the figure does not scale with n_gpus"},
        "cpu_baseline": cb,
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
        cpu_baseline, _ = cpu_baseline_block(batch, ctx, 2)

    # ---------------- BASELINE configs[2]: prefill 8 x 2048 tokens through the tcgen05 int8 ternary GEMM (N = 1 only)
    if args.workload == "bs64" and world == 1 and not args.no_aux:
        try:
            sys.path.insert(0, str(ROOT / "tools"))
            import bench_prefill
            aux["prefill"] = bench_prefill.measure(16384, 8, 3, "2b")
            aux["prefill"]["workload"] = "BASELINE.json configs[2]: BitNet-2B-4T-shaped prefill, 8 x 2048-token prompts, the 4 fused ternary linears of a layer at M = 16384"
        except Exception as ex:  # noqa: BLE001
            aux["prefill"] = {"error": str(ex)[:300]}

    # ---------------- BASELINE configs[3] / [4]: tensor-parallel 8B / 70B-shaped decode at TP = N (every N, incl. the TP = 1 lines at N = 1)
    if args.workload == "bs64" and not args.no_aux and not args.no_tp:
        try:
            tp_aux = run_tp_suite(rank, world, dist, 20, 3, peak)
            if rank == 0:
                aux["tp"] = tp_aux
        except Exception as ex:  # noqa: BLE001
            aux["tp"] = {"error": str(ex)[:300]}

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
            "config": {"workload": workload_name(args.workload), "batch": batch, "batch_per_gpu": batch, "ctx": ctx, "page_size": page,
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
    ap.add_argument("--no-tp", action="store_true", help="skip the tensor-parallel suite (aux.tp)")
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
