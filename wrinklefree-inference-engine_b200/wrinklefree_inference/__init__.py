"""Host-side mirror of the reference's `wrinklefree_inference` wrappers for the BitNet hot path
(kernels, kv_cache, cache.bitnet_converter), running on the sm_100a library.  Same public names as
src/wrinklefree_inference/{kernels,kv_cache,cache}/__init__.py of the reference."""
