"""wf_bitnet — host side of the B200 BitNet b1.58 decode path (ctypes over libwf_bitnet.so)."""
from ._lib import BITNET_BLOCK_SIZE, WfError, available, lib, require_gpu  # noqa: F401
