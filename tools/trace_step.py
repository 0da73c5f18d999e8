#!/usr/bin/env python
"""WF_TRACE=1 python tools/trace_step.py [--shape 2b|8b] [--batch 1]: per-phase timeline of one graph-replayed decode step
(globaltimer stamps written by CTA 0 and the last CTA of every linear_small launch)."""
import argparse, os, sys
from pathlib import Path
os.environ["WF_TRACE"] = "1"
ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
from wf_bitnet import _lib
from wf_bitnet.engine import Engine, EngineConfig
ap = argparse.ArgumentParser(); ap.add_argument("--shape", default="2b"); ap.add_argument("--batch", type=int, default=1); ap.add_argument("--ctx", type=int, default=512)
a = ap.parse_args()
_lib.require_gpu(); torch.cuda.set_device(0)
mk = {"2b": EngineConfig.bitnet_2b, "8b": EngineConfig.llama3_8b}[a.shape]
cfg = mk(max_seq=a.ctx + 64, page_size=16, max_batch=a.batch)
eng = Engine(cfg); eng.init_synthetic(1)
for s in range(a.batch): eng.set_context(s, a.ctx)
eng.decode([1] * a.batch)
for _ in range(5): eng.decode_device(a.batch)
eng.sync()
n = 4 * cfg.layers
raw = eng.buffer("trace", (5 * cfg.layers + 8, 2, 16), torch.int64).cpu().numpy().astype(np.float64) / 1e3   # us
t = raw[:n]
at = raw[n + 4:n + 4 + cfg.layers, 0]
t0 = t[0, 0, 0]
names = ["qkv", "o", "gateup", "down"]
print("phase        entry   +wait   +flag   +prolog +gemv   +red   | lastCTA: entry +wait +prolog +gemv +red | gap to next entry")
for i in range(4, 12):
    c0, c1 = t[i, 0], t[i, 1]
    nxt = t[i + 1, 0, 0] if i + 1 < n else float("nan")
    print(f"L{i//4} {names[i%4]:7s} {c0[0]-t0:8.2f} " + " ".join(f"{c0[k]-c0[0]:7.2f}" for k in range(1, 6)) + "  | " + f"{c1[0]-c0[0]:6.2f} " + " ".join(f"{c1[k]-c1[0]:6.2f}" for k in (1, 3, 4, 5)) + f" | {nxt-c0[0]:6.2f}")
print("prologue detail (CTA0, row 0; us after the dependency wait): pass1 done, +bar, pass2 done, +bar, pass3 done")
for i in range(4, 8):
    c0 = t[i, 0]
    print(f"  {names[i%4]:7s} " + " ".join(f"{c0[8+k]-c0[2]:6.2f}" for k in range(5)) + f"   prologue end {c0[3]-c0[2]:6.2f}")
print("attention (CTA 0; us): entry(rel. qkv red end), wait done, q prologue done, kv loop done, warp merge+store done, end")
for l in (1, 2):
    q_end = t[4 * l, 0, 5]
    print(f"  L{l}: " + " ".join(f"{at[l, k]-q_end:7.2f}" for k in range(6)) + f"   o-proj wait done {t[4*l+1,0,1]-q_end:7.2f}")
d = np.diff(t[:, 0, 0])
for k in range(4): print(f"{names[k]:7s}: median entry->next entry {np.median(d[k::4]):.2f} us")
print("step (first qkv entry -> last down end): %.1f us" % (t[n-1, 0, 5] - t[0, 0, 0]))
