#!/usr/bin/env python
"""One decode step of the BitNet-2B-4T-shaped engine bracketed by cudaProfilerStart/Stop, for ncu:

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches_bs64.csv python tools/profile_step.py --batch 64 --ctx 4096

Without ncu it prints the CUDA-event time of the same step (ms), so the plain run doubles as the
"exited 0 without ncu" gate of B200_PROFILING.md.
"""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--shape", default="2b", choices=["2b", "8b", "70b"])
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (0 = the shape's own)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--kv", default="bf16", choices=["bf16", "int8"], help="KV cache precision")
    ap.add_argument("--tp-shard", type=int, default=0, help="single-GPU stand-in for ONE rank of a TP=n engine: heads, kv heads and ffn divided by n "
                                                          "(the compute of a rank without the exchanges)")
    args = ap.parse_args()

    from wf_bitnet import _lib
    from wf_bitnet.engine import Engine, EngineConfig

    _lib.require_gpu()
    torch.cuda.set_device(0)
    mk = {"2b": EngineConfig.bitnet_2b, "8b": EngineConfig.llama3_8b, "70b": EngineConfig.llama3_70b}[args.shape]
    kw = dict(max_seq=args.ctx + 64, page_size=16, max_batch=args.batch)
    if args.kv == "int8":
        kw["kv_dtype"] = "int8"
    cfg = mk(**kw)
    if args.layers:
        cfg.layers = args.layers
    if args.tp_shard > 1:
        cfg.heads //= args.tp_shard
        cfg.kv_heads //= args.tp_shard
        cfg.ffn //= args.tp_shard
    eng = Engine(cfg)
    eng.init_synthetic(1234)
    for slot in range(args.batch):
        eng.set_context(slot, args.ctx, fill_random=False)   # zero KV: same bytes moved, no 16K fill launches
    eng.decode([1] * args.batch)
    for _ in range(args.warmup):
        eng.decode_device(args.batch, use_graph=not args.no_graph)
    eng.sync()
    ext = torch.cuda.ExternalStream(eng.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()
    e0.record(ext)
    for _ in range(args.steps):
        eng.decode_device(args.batch, use_graph=not args.no_graph)
    e1.record(ext)
    e1.synchronize()
    torch.cuda.profiler.stop()
    print(f"batch={args.batch} ctx={args.ctx} shape={args.shape} kv={args.kv} steps={args.steps} ms_per_step={e0.elapsed_time(e1) / args.steps:.4f} "
          f"launches_per_step={eng.launches_per_step()}")
    eng.close()


if __name__ == "__main__":
    main()
