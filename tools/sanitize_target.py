#!/usr/bin/env python
"""Small target for compute-sanitizer (memcheck / racecheck / synccheck / initcheck, ONE tool per gpurun call):
the split-K GEMV with its self-cleaning workspace, the one-shot batched GEMM, the phase kernels, the persistent
layer-tail kernel with its grid barrier, the fused attention with its in-kernel chunk merge, the LM head — on toy shapes,
eager launches (no CUDA graph), a few steps each."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    import os
    from wf_bitnet import ops
    from wf_bitnet.engine import Engine, EngineConfig
    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda").manual_seed(0)
    # GEMV: even tile split with the split-K round (odd row-block count), whole-row-block slabs, mma and dp4a consumers
    for (n, k, m) in ((2576, 1024, 1), (2576, 1024, 4), (512, 2560, 8), (4096, 512, 2)):
        packed = torch.randint(0, 256, (n, k // 4), generator=g, device="cuda", dtype=torch.uint8)
        tile = ops.i2s_to_tile(packed)
        q = torch.randint(-127, 128, (m, k), generator=g, device="cuda", dtype=torch.int8)
        a = ops.ternary_linear(tile, n, k, q, mode="i32_plus1", algo=1)
        b = ops.ternary_linear(tile, n, k, q, mode="i32_plus1", algo=2)
        assert torch.equal(a, b)
    # engine steps: tail kernel (B <= 8), phase kernels (WF_TAIL=0), batched phases (B > 8), split attention chunks
    for B, env in ((1, {"WF_TAIL": "1"}), (3, {"WF_TAIL": "1"}), (3, {"WF_TAIL": "0"}), (12, {"WF_TAIL": "1"}), (12, {"WF_TAIL": "0"})):
        os.environ.update(env)
        cfg = EngineConfig(hidden=256, layers=2, heads=4, kv_heads=2, head_dim=128, ffn=384, vocab=512, max_seq=1088, page_size=16, max_batch=B)
        eng = Engine(cfg)
        eng.init_synthetic(3)
        for slot in range(B):
            eng.set_context(slot, 1030 + slot, fill_random=True)      # > 1024 tokens: two attention chunks + the in-kernel merge
        toks = np.arange(B, dtype=np.int32)
        for _ in range(2):
            toks = eng.decode(toks)
        for _ in range(2):
            eng.decode_device(B, use_graph=False)
        eng.sync()
        eng.close()
    torch.cuda.synchronize()
    print("sanitize target ok")


if __name__ == "__main__":
    main()
