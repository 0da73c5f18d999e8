#!/usr/bin/env python
"""WF_TRACE timeline of the persistent layer-tail kernel (B <= 8): globaltimer stamps of CTA 0 for every phase of one graph-replayed
decode step.  python tools/trace_tail.py [--batch 1] [--ctx 512]"""
import argparse, os, sys
from pathlib import Path
os.environ["WF_TRACE"] = "1"
ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
from wf_bitnet import _lib
from wf_bitnet.engine import Engine, EngineConfig
ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=1); ap.add_argument("--ctx", type=int, default=512)
a = ap.parse_args()
_lib.require_gpu(); torch.cuda.set_device(0)
cfg = EngineConfig.bitnet_2b(max_seq=a.ctx + 64, page_size=16, max_batch=a.batch)
eng = Engine(cfg); eng.init_synthetic(1)
for s in range(a.batch): eng.set_context(s, a.ctx)
eng.decode([1] * a.batch)
for _ in range(5): eng.decode_device(a.batch)
eng.sync()
L = cfg.layers
raw = eng.buffer("trace", (5 * L + 8, 32), torch.int64).cpu().numpy().astype(np.float64) / 1e3   # us
tail = raw[:L]
attn = raw[4 * L + 4: 4 * L + 4 + L]
names = ["o", "gate|up", "down", "qkv'"]
print("tail kernel, CTA 0 (us relative to kernel entry): wait = dependency (attention) done")
for l in (1, 2, 3):
    t = tail[l]; e0 = t[0]
    print(f"layer {l}: entry {e0 - tail[1][0]:8.2f} (since layer 1 entry)   dependency wait done +{t[1]-e0:6.2f}")
    prev = t[1]
    for pi in range(4):
        b, g, w, m, ep = t[2 + 5 * pi: 7 + 5 * pi]
        print(f"   {names[pi]:8s} barrier {b-prev:6.2f} | glue {g-b:6.2f} | first stage wait {w-g:6.2f} | main loop {m-w:6.2f} | epilogue {ep-m:6.2f} | phase total {ep-prev:6.2f}")
        prev = ep
    print(f"   tail total (entry -> last epilogue) {prev-e0:7.2f};  attention of layer {l+1}: entry {attn[l+1][0]-prev:6.2f} after, its end {attn[l+1][5]-prev:6.2f} after; next tail entry {tail[l+1][0]-prev:6.2f} after")
print("fused attention kernel, CTA 0 (us): entry -> dependency wait done -> q/RoPE/append done -> KV stream done -> CTA merge / store done -> end")
for l in (2, 3, 4):
    t = attn[l]
    print(f"layer {l}: wait {t[1]-t[0]:6.2f} | prologue {t[2]-t[1]:6.2f} | stream {t[3]-t[2]:6.2f} | merge+store {t[4]-t[3]:6.2f} | chunk merge {t[5]-t[4]:6.2f} | total {t[5]-t[0]:6.2f};"
          f"  entry {t[0]-tail[l-1][21]:6.2f} after the previous tail's last epilogue, next tail's dependency done {tail[l][1]-t[5]:6.2f} after its end")
tot = [tail[l][21] - tail[l][1] for l in range(1, L - 1)]
att = [tail[l + 1][1] - tail[l][21] for l in range(1, L - 2)]
print(f"median over layers: tail (dependency done -> last epilogue) {np.median(tot):.2f} us; gap tail end -> next tail's dependency done (attention) {np.median(att):.2f} us")
