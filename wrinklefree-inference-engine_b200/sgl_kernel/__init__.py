"""Drop-in `sgl_kernel` surface for the BitNet path, backed by libwf_bitnet.so (B200, sm_100a).

Only the BitNet pieces of sgl-kernel that the WrinkleFree engine calls are provided
(reference call sites: src/wrinklefree_inference/kernels/bitnet_patch.py:14-21,
scripts/direct_inference_full.py:22, tests/test_kv_cache_cpp.py:273).  Importing this package
registers the `torch.ops.sgl_kernel.bitnet_*` custom ops, the same import side effect the
reference relies on (docs/research/notebook.md:97-99).
"""
from . import _ops  # noqa: F401  (registers torch.ops.sgl_kernel.*)
from . import quantization  # noqa: F401
from . import kvcache  # noqa: F401

__all__ = ["quantization", "kvcache"]
