"""Deterministic cache keys (reference: src/wrinklefree_inference/cache/cache_key.py:1-67), with the pack format as a
parameter: the reference hard-codes `"pack_format": "blocked_uint8_v1"` (:27); a cache entry that also stores the GPU tile
layout is keyed with `"b200_tile_v1"` so the two never collide.  For the default format the key is the reference's, byte
for byte."""
import hashlib
import json
from pathlib import Path
from typing import Optional


def compute_cache_key(model_path: str, revision: Optional[str] = None, pack_format: str = "blocked_uint8_v1") -> str:
    key_data = {"model_path": model_path, "revision": revision or "main", "pack_format": pack_format}
    local_path = Path(model_path)
    if local_path.exists():
        key_data["source_hash"] = _compute_local_model_hash(local_path)
    key_hash = hashlib.sha256(json.dumps(key_data, sort_keys=True).encode()).hexdigest()[:12]
    model_name_safe = model_path.replace("/", "_").replace(".", "_").replace("-", "_")
    return f"{model_name_safe}_{key_hash}"


def _compute_local_model_hash(model_path: Path) -> str:
    hasher = hashlib.sha256()
    for sf in sorted(model_path.glob("*.safetensors")):
        hasher.update(sf.name.encode())
        hasher.update(str(sf.stat().st_size).encode())
        hasher.update(str(sf.stat().st_mtime_ns).encode())
    config_file = model_path / "config.json"
    if config_file.exists():
        hasher.update(config_file.read_bytes())
    return hasher.hexdigest()[:16]
