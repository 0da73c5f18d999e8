"""Weight-cache format helpers (reference: src/wrinklefree_inference/cache/__init__.py; the GCS / cache-key
plumbing is out of scope)."""
from .bitnet_converter import BITNET_WEIGHT_PATTERNS, convert_state_dict, pack_state_for_b200  # noqa: F401
