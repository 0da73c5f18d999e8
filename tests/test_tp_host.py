"""Host-side tensor-parallel logic on CPU: weight sharding reassembles to the full matrices, shard arithmetic equals
the unsharded arithmetic on the oracle, and the handle exchange works over a world_size-2 gloo group."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

import oracle as O

ROOT = Path(__file__).resolve().parent.parent


class Cfg:
    hidden, layers, heads, kv_heads, head_dim, ffn, vocab, rope_theta, max_seq = 256, 2, 4, 2, 128, 512, 300, 5e5, 32


def test_shard_weights_reassemble_and_vocab_slices():
    from wf_bitnet import tp
    from wf_bitnet.synth import make_weights
    w = make_weights(Cfg, seed=3)
    for size in (2,):
        shards = [tp.shard_weights(w, r, size) for r in range(size)]
        for li, layer in enumerate(w["layers"]):
            for name in ("q", "k", "v", "gate", "up"):
                cat = np.concatenate([s["layers"][li][f"{name}_proj"] for s in shards], axis=0)
                assert np.array_equal(cat, layer[f"{name}_proj"])
            for name in ("o", "down"):
                cat = np.concatenate([s["layers"][li][f"{name}_proj"] for s in shards], axis=1)
                assert np.array_equal(cat, layer[f"{name}_proj"])
            assert shards[0]["layers"][li]["o_scale"] == layer["o_scale"]
        assert shards[1]["embed"] is w["embed"]
    for vocab, size in ((128256, 8), (1000, 2), (300, 2), (128256, 1)):
        sl = [tp.vocab_slice(vocab, r, size) for r in range(size)]
        assert sl[0][0] == 0 and sum(n for _, n in sl) == vocab
        assert all(sl[i][0] + sl[i][1] == sl[i + 1][0] for i in range(size - 1))
        assert all(lo % 128 == 0 for lo, _ in sl)
    with pytest.raises(ValueError):
        tp.shard_cols(np.zeros((4, 32), np.uint8), 0, 2)      # K = 128 cannot be cut into two 128-blocks


def test_column_shards_sum_to_the_full_accumulator():
    """o / down under TP: every rank's exact partial sum over its K slice adds up to the TP = 1 accumulator (what the
    peer REDs compute), given the SAME full-vector int8 codes on every rank."""
    from wf_bitnet import tp
    rng = np.random.default_rng(0)
    wt = rng.integers(-1, 2, size=(64, 1024)).astype(np.float32)
    packed = O.pack_ternary(wt)[0]
    q = rng.integers(-128, 128, size=(3, 1024)).astype(np.int8)
    full = O.gemm_i32_fast(packed, q, 0)
    for size in (2, 4, 8):
        s = 1024 // size
        parts = [O.gemm_i32_fast(tp.shard_cols(packed, r, size), q[:, r * s:(r + 1) * s], 0) for r in range(size)]
        assert np.array_equal(sum(parts), full)


_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import torch.distributed as dist
from wf_bitnet import tp
dist.init_process_group("gloo")
r = dist.get_rank()
hs = tp.exchange_handles(bytes([r]) * 64)
assert hs == [bytes([0]) * 64, bytes([1]) * 64], hs
dist.barrier()
dist.destroy_process_group()
"""


def test_handle_exchange_gloo_world2(tmp_path):
    import subprocess
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port", "29541",
           str(script), str(ROOT), str(ROOT / "wrinklefree-inference-engine_b200")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
