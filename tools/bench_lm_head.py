#!/usr/bin/env python
"""LM head (bf16 tied embedding x hidden, fused per-CTA argmax) timed alone at several batch sizes."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "wrinklefree-inference-engine_b200"))
from wf_bitnet.engine import Engine, EngineConfig  # noqa: E402

for b in (64, 32, 16, 1):
    e = Engine(EngineConfig.bitnet_2b(max_batch=b, max_seq=256))
    e.init_synthetic(1)
    for s in range(b):
        e.set_context(s, 32, fill_random=True)
    e.decode_device(b)
    e.sync()
    us = e.time_kernel("lm_head", b, 20) * 1e3
    print(f"lm_head batch {b}: {us:.1f} us  {656.7e6 / us / 1e3:.0f} GB/s")
    e.close()
