#!/usr/bin/env python
"""Experiment: does a decode step of 64 sequences go faster as n concurrent sub-batches (n engines x 64/n sequences, one stream
each) than as one batch?  While one sub-batch streams its KV cache (HBM-bound) the others run their latency-bound linears.
Pessimistic stand-in for an in-engine split: weights and the LM head are read n times."""
import argparse
import sys
import time
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
    ap.add_argument("--parts", type=int, default=2)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--skew-us", type=float, default=0.0, help="host sleep between the launches of the parts (first step only)")
    args = ap.parse_args()
    from wf_bitnet import _lib
    from wf_bitnet.engine import Engine, EngineConfig
    _lib.require_gpu()
    torch.cuda.set_device(0)
    nb = args.batch // args.parts
    engs = []
    for i in range(args.parts):
        cfg = EngineConfig.bitnet_2b(max_seq=args.ctx + 64, page_size=16, max_batch=nb)
        e = Engine(cfg)
        e.init_synthetic(1234 + i)
        for slot in range(nb):
            e.set_context(slot, args.ctx, fill_random=False)
        e.decode([1] * nb)
        for _ in range(3):
            e.decode_device(nb)
        e.sync()
        engs.append(e)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(args.steps):
        for i, e in enumerate(engs):
            e.decode_device(nb)
            if s == 0 and args.skew_us and i + 1 < len(engs):
                time.sleep(args.skew_us * 1e-6)
    for e in engs:
        e.sync()
    dt = (time.perf_counter() - t0) / args.steps
    print(f"parts={args.parts} batch_each={nb} ctx={args.ctx} ms_per_step_all={dt * 1e3:.4f} tok_s={args.batch / dt:.0f}")
    for e in engs:
        e.close()


if __name__ == "__main__":
    main()
