"""GPU parity of the checkpoint loader + cache converter: an HF-style snapshot with the projections stored
as HF-packed uint8 [N/4, K] (tests/test_bitnet_core.py:21-53), as ternary floats, and as a packed_bitnet_v1 / b200_tile_v1 cache
entry written by convert_and_save_bitnet must all load into the engine to the SAME logits as the in-memory weights."""
import json

import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu


def _hf_pack(w: np.ndarray) -> np.ndarray:
    """inverse of _unpack_ternary_weights: rows [i*n, (i+1)*n) go to bits 2i..2i+1 of packed row r (blocked along N)."""
    N, K = w.shape
    n = N // 4
    c = (w.astype(np.int16) + 1).astype(np.uint8)
    return (c[0:n] | (c[n:2 * n] << 2) | (c[2 * n:3 * n] << 4) | (c[3 * n:4 * n] << 6)).astype(np.uint8)


def _state_dict(cfg, weights, form):
    sd = {"model.embed_tokens.weight": weights["embed"], "model.norm.weight": weights["final_norm"]}
    for li, layer in enumerate(weights["layers"]):
        p = f"model.layers.{li}"
        sd[f"{p}.input_layernorm.weight"] = layer["input_layernorm"]
        sd[f"{p}.post_attention_layernorm.weight"] = layer["post_attention_layernorm"]
        sd[f"{p}.self_attn.attn_sub_norm.weight"] = layer["attn_sub_norm"]
        sd[f"{p}.mlp.ffn_sub_norm.weight"] = layer["ffn_sub_norm"]
        for name in ("q", "k", "v", "o", "gate", "up", "down"):
            base = f"{p}.{'self_attn' if name in 'qkvo' else 'mlp'}.{name}_proj"
            w = layer[f"{name}_ternary"]
            if form == "hf":
                sd[f"{base}.weight"] = torch.from_numpy(_hf_pack(w))
            else:
                sd[f"{base}.weight"] = torch.from_numpy(w.astype(np.float32)).to(torch.bfloat16)
            sd[f"{base}.weight_scale"] = torch.tensor([layer[f"{name}_scale"]], dtype=torch.float32)
    return sd


def _logits(eng, cfg, toks):
    out = []
    for t in toks:
        eng.decode([t])
        out.append(eng.buffer("logits", (1, cfg.vocab), torch.float32)[0].cpu())
    return out


def test_loader_every_on_disk_form(tmp_path):
    from safetensors.torch import save_file
    from wf_bitnet import loader
    from wf_bitnet.engine import Engine, EngineConfig
    from wf_bitnet.synth import make_weights
    from wrinklefree_inference.cache import convert_and_save_bitnet

    cfg = EngineConfig(hidden=256, layers=2, heads=4, kv_heads=2, head_dim=128, ffn=384, vocab=512, max_seq=64, page_size=16, max_batch=1)
    weights = make_weights(cfg, seed=21, keep_ternary=True)
    toks = [5, 77, 300, 9]
    ref = Engine(cfg)
    ref.load_weights(weights)
    want = _logits(ref, cfg, toks)
    ref.close()

    # the HF form really is what the reference unpacks: oracle restatement of _unpack_ternary_weights
    w0 = weights["layers"][0]["q_ternary"]
    assert np.array_equal(O.unpack_hf(_hf_pack(w0)), w0.astype(np.float32))

    dirs = {}
    for form in ("hf", "float"):
        d = tmp_path / form
        d.mkdir()
        sd = _state_dict(cfg, weights, form)
        keys = sorted(sd)
        half = len(keys) // 2   # two shards, like a real snapshot
        save_file({k: sd[k].contiguous() for k in keys[:half]}, str(d / "model-00001-of-00002.safetensors"))
        save_file({k: sd[k].contiguous() for k in keys[half:]}, str(d / "model-00002-of-00002.safetensors"))
        (d / "config.json").write_text(json.dumps({"hidden_size": cfg.hidden}))
        dirs[form] = d
    dirs["packed_v1"] = convert_and_save_bitnet(str(dirs["hf"]), tmp_path / "cache_v1")
    dirs["tile_v1"] = convert_and_save_bitnet(str(dirs["float"]), tmp_path / "cache_tile", pack_format="b200_tile_v1")
    meta = json.loads((dirs["packed_v1"] / "cache_metadata.json").read_text())
    assert meta["format_version"] == "packed_bitnet_v1" and meta["pack_format"] == "blocked_uint8_v1"
    assert (dirs["packed_v1"] / "config.json").exists()

    # converted shards hold the reference's names and exactly the reference's packed bytes
    from safetensors.torch import load_file
    conv = {}
    for f in sorted(dirs["packed_v1"].glob("*.safetensors")):
        conv.update(load_file(str(f)))
    q = conv["model.layers.0.self_attn.q_proj.qweight"].numpy()
    assert np.array_equal(q, O.pack_ternary(w0.astype(np.float32))[0])
    assert "model.layers.0.self_attn.q_proj.weight" not in conv and "model.layers.0.self_attn.q_proj.weight_scale" in conv

    for form, d in dirs.items():
        eng = Engine(cfg)
        m = loader.load_safetensors_dir(eng, d, expect_format="packed_bitnet_v1" if "v1" in form else None)
        got = _logits(eng, cfg, toks)
        eng.close()
        for a, b in zip(got, want):
            assert torch.equal(a, b), form
        if form == "tile_v1":
            assert m["pack_format"] == "b200_tile_v1"


def test_loader_errors():
    from wf_bitnet import loader
    from wf_bitnet.engine import Engine, EngineConfig
    cfg = EngineConfig(hidden=256, layers=1, heads=2, kv_heads=1, head_dim=128, ffn=384, vocab=512, max_seq=32, page_size=16, max_batch=1)
    eng = Engine(cfg)
    with pytest.raises(KeyError):
        loader.load_state_dict(eng, {})
    with pytest.raises(ValueError):
        loader.load_state_dict(eng, {"model.layers.0.self_attn.q_proj.qweight": torch.zeros(8, 8, dtype=torch.uint8),
                                     "model.layers.0.self_attn.q_proj.weight_scale": torch.ones(1)})
    # a float master-weight checkpoint (not ternary) must be rejected, not clipped to {-1,0,+1} silently
    # (the reference converter guards the same path with _is_ternary_float, cache/bitnet_converter.py:80-103,161)
    g = torch.Generator().manual_seed(3)
    with pytest.raises(ValueError, match="not ternary"):
        loader.load_state_dict(eng, {"model.layers.0.self_attn.q_proj.weight": (0.02 * torch.randn(256, 256, generator=g)).to(torch.bfloat16),
                                     "model.layers.0.self_attn.q_proj.weight_scale": torch.ones(1)})
    eng.close()
