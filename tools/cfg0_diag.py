#!/usr/bin/env python
"""Diagnostic for BASELINE configs[0] parity (2B shape, bs1): where do the engine and the oracle part ways?
For L in --layers: teacher-forced steps, per-position normwise logit error, final normalised hidden-state equality, token agreement;
and the oracle against ITSELF with fp32 attention (sensitivity of a deep random-init model to a bf16-level attention change)."""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", default="1,4,30")
    ap.add_argument("--steps", type=int, default=40)
    args = ap.parse_args()
    from oracle import model as OM
    from wf_bitnet.engine import Engine, EngineConfig
    from wf_bitnet.synth import make_weights
    full = EngineConfig.bitnet_2b(max_seq=128, max_batch=1)
    w_all = make_weights(full, seed=1234)
    toks = np.random.default_rng(0).integers(0, full.vocab, size=args.steps)
    for L in [int(x) for x in args.layers.split(",")]:
        cfg = EngineConfig.bitnet_2b(max_seq=128, max_batch=1, layers=L)
        w = dict(w_all)
        w["layers"] = w_all["layers"][:L]
        oc = OM.ModelConfig(layers=L, max_seq=128)
        om = OM.OracleModel(oc, w, use_ref=True)
        om32 = OM.OracleModel(oc, w, use_ref=True, attn_fp32=True)
        eng = Engine(cfg)
        eng.load_weights(w)
        rows = []
        for pos, t in enumerate(toks):
            lo, ho = om.forward(int(t), pos, return_hidden=True)
            l32 = om32.forward(int(t), pos).float()
            got = int(eng.decode([int(t)])[0])
            lg = eng.buffer("logits", (1, cfg.vocab), torch.float32)[0].cpu()
            hg = eng.buffer("normed", (1, cfg.hidden), torch.bfloat16)[0].cpu()
            lo = lo.float()
            rel = float((lg - lo).norm() / lo.norm())
            rel32 = float((l32 - lo).norm() / lo.norm())
            hid_eq = float((hg != ho).float().mean())
            top2 = torch.topk(lo, 2)
            rows.append((pos, rel, rel32, hid_eq, got == int(top2.indices[0]), int(l32.argmax()) == int(top2.indices[0]), float(top2.values[0] - top2.values[1])))
        eng.close()
        print(f"=== layers={L}")
        for r in rows:
            print("pos %3d  rel(engine,oracle)=%.5f  rel(oracle_fp32attn,oracle)=%.5f  hidden mismatch frac=%.4f  tok_eq=%d  tok_eq(fp32attn)=%d  gap=%.4f" % r)
        print("summary L=%d: median rel engine %.5f, oracle-self %.5f; token agreement engine %d/%d, oracle-self %d/%d; pos0 hidden mismatch %.4f" % (
            L, float(np.median([r[1] for r in rows])), float(np.median([r[2] for r in rows])), sum(r[4] for r in rows), len(rows),
            sum(r[5] for r in rows), len(rows), rows[0][3]))


if __name__ == "__main__":
    main()
