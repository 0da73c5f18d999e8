"""Weight-cache format helpers (reference: src/wrinklefree_inference/cache/__init__.py).  The converter and the cache key
are mirrored (the first packs on the GPU); the GCS client and the download / upload loader stay with the reference."""
from .bitnet_converter import (BITNET_WEIGHT_PATTERNS, B200_PACK_FORMAT, CACHE_FORMAT_VERSION, convert_and_save_bitnet,  # noqa: F401
                               convert_state_dict, pack_state_for_b200)
from .cache_key import compute_cache_key  # noqa: F401
