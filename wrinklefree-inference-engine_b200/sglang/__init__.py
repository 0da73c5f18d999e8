"""Stand-in for the `sglang` package of the reference's fork (DeepOpt-com/sglang-bitnet): only
`sglang.srt.models.bitnet` — the module the reference's tests and scripts import for the BitNet path
(tests/test_bitnet_weight_loading.py:22-135, tests/test_bitnet_comparison.py:16-74, tests/test_vllm_cpu_stub.py:237-354,
scripts/sglang_model_direct.py:20, scripts/direct_inference_full.py:21) — is provided, backed by libwf_bitnet.so."""
