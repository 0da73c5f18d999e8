/* wf_bitnet.h — C ABI of libwf_bitnet.so, the B200 (sm_100a) implementation of the
 * WrinkleFree / sgl-kernel BitNet b1.58 decode hot path.
 *
 * Drop-in boundary.  Every entry point takes plain pointers, sizes and a CUDA stream
 * (as void*; pass torch.cuda.current_stream().cuda_stream).  All data pointers are DEVICE
 * pointers unless a parameter is named host_*.  Functions are asynchronous on `stream`,
 * never allocate device memory behind the caller's back (workspaces are explicit), never
 * throw and never print: they return WF_OK (0) or a negative status, and wf_last_error()
 * gives the message.  No torch types appear here; the reference-side binding (torch custom
 * ops under the sgl_kernel names) is shown in INTEGRATION.md.
 *
 * Each function cites the reference interface it replaces (paths relative to the
 * reference repository root).
 */
#ifndef WF_BITNET_H
#define WF_BITNET_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WF_OK 0
#define WF_ERR_INVALID (-1)     /* bad argument (reference: Python ValueError / TORCH_CHECK) */
#define WF_ERR_UNSUPPORTED (-2) /* shape outside what the kernels implement */
#define WF_ERR_CUDA (-3)        /* CUDA runtime error; message in wf_last_error() */

#define WF_BITNET_BLOCK_SIZE 128 /* sgl_kernel.quantization.BITNET_BLOCK_SIZE
                                    (src/wrinklefree_inference/kernels/bitnet_patch.py:20) */

/* element types of floating-point buffers */
#define WF_F32 0
#define WF_BF16 1
#define WF_F16 2
#define WF_I8 3 /* KV pools only: int8 codes + one fp32 scale per (token, kv head) */

/* ---- library ------------------------------------------------------------------------ */

/* sgl_kernel.quantization.bitnet_check_kernel_available() (bitnet_patch.py:15,22):
 * 1 when a compute-capability-10.x device is usable, else 0. */
int wf_bitnet_available(void);
const char* wf_last_error(void);
int wf_sm_count(void);
const char* wf_version(void);
/* plain device-to-device copy on `stream` (debug taps of engine buffers) */
int wf_copy_d2d(void* dst, const void* src, size_t bytes, void* stream);

/* ---- weight formats (north-star item 1) ------------------------------ */

/* _pack_ternary_weights (src/wrinklefree_inference/cache/bitnet_converter.py:26-77;
 * == quantize_i2_s, extern/BitNet/src/ggml-bitnet-mad.cpp:46-92): fp32 [N,K] ternary-ish ->
 * blocked I2_S u8 [N,K/4]; *scale_out (device f32[1]) = max|w| (1.0 if < 1e-6). K % 4 == 0. */
int wf_pack_ternary(const void* w, int w_dtype, int N, int K, uint8_t* i2s, float* scale_out, void* stream);

/* _unpack_i2_to_ternary (tests/test_bitnet_core.py:109-136): I2_S -> fp32 [N,K] in {-1,0,1}. */
int wf_unpack_i2s(const uint8_t* i2s, int N, int K, float* w, void* stream);

/* _unpack_ternary_weights + _pack_ternary_weights fused (tests/test_bitnet_core.py:21-53,
 * scripts/direct_inference_full.py:113-117): HF checkpoint u8 [N/4,K] -> I2_S u8 [N,K/4]. */
int wf_hf_to_i2s(const uint8_t* hf, int N4, int K, uint8_t* i2s, void* stream);

/* GPU tile layout "b200_tile_v1": [ceil(N/16)][K/128][512 B]; one 512-B tile = 16 rows x 128 k,
 * lane l of a warp owns bytes [16l,16l+16) = the four u32 A-fragment source words of
 * mma.m16n8k32.  Rows N..ceil16(N) are filled with code 1 (weight 0).
 * K % 128 == 0. */
size_t wf_tile_bytes(int N, int K);
int wf_repack_i2s_to_tile(const uint8_t* i2s, int N, int K, uint8_t* tile, void* stream);
int wf_repack_tile_to_i2s(const uint8_t* tile, int N, int K, uint8_t* i2s, void* stream);
/* gate/up fusion helper: interleave two I2_S matrices [N,K/4] in 16-row blocks
 * (out rows 32i..32i+15 = a rows 16i.., out rows 32i+16..32i+31 = b rows 16i..). N % 16 == 0. */
int wf_interleave_rows16(const uint8_t* a, const uint8_t* b, int N, int K4, uint8_t* out, void* stream);

/* ---- activation quantisation (north-star item 2) -------------------------- */

/* sgl_kernel.quantization.bitnet_quantize_activations / torch.ops.sgl_kernel.
 * bitnet_quantize_activations_cpu (docs/research/sglang_optimization_plan.md:72-74; arithmetic:
 * tests/test_bitnet_core.py:139-146).  x [M,K] (row stride ldx elements) -> q int8 [M,K].
 * per_token == 0: ONE absmax over the whole tensor (the reference op), scale_out f32[1];
 * per_token == 1: one scale per row, scale_out f32[M] (what the reference sees at M == 1).
 * rowsum_out (optional, i32[M]) = sum_k q[m,k], the Sigma-a correction term
 * (scripts/direct_inference_full.py:204-206).  workspace: >= 4 bytes, only for per_token == 0. */
int wf_act_quant_i8(const void* x, int x_dtype, int M, int K, int ldx, int per_token, int8_t* q, float* scale_out,
                    int32_t* rowsum_out, void* workspace, void* stream);

/* ---- ternary linear (north-star items 3, 4) ---------------------------- */

/* weight layouts */
#define WF_W_I2S 0  /* reference layout u8 [N,K/4] (what sgl_kernel ops receive) */
#define WF_W_TILE 1 /* b200_tile_v1 (what the engine keeps resident) */

/* output modes */
#define WF_OUT_I32_PLUS1 0 /* int32 [M,N] = sum (w+1)*a     — the reference kernel's accumulator */
#define WF_OUT_I32 1       /* int32 [M,N] = sum w*a         — needs rowsum */
#define WF_OUT_COMPAT_F32 2 /* f32 = host_scale * (float)sum (w+1)*a — bitnet_gemv / bitnet_gemm */
#define WF_OUT_LINEAR 3    /* (sum w*a) * w_scale[seg] * act_scale[m] -> out_dtype — bitnet_linear */

/* epilogue flags (WF_OUT_LINEAR only) */
#define WF_EPI_RESIDUAL 1  /* out = bf16(residual + bf16(lin))  (hidden = hidden + proj, direct_inference_full.py:276,283) */
#define WF_EPI_RELU2_MUL 2 /* rows interleaved gate/up in 16-row blocks; out[M,N/2] = relu(g)^2 * u (…:259) */

typedef struct wf_linear_args {
    const uint8_t* w;     /* weights, layout per w_layout */
    int w_layout;         /* WF_W_I2S | WF_W_TILE */
    int N, K;             /* out_features, in_features (K % 128 == 0) */
    const int8_t* q;      /* int8 activations [M,K], row stride ldq bytes */
    int M, ldq;
    const int32_t* rowsum;  /* i32[M] sum_k q (required for WF_OUT_I32 / WF_OUT_LINEAR) */
    const float* act_scale; /* f32[M] (per token) or f32[1]; WF_OUT_LINEAR */
    int act_scale_per_token;
    const float* w_scale; /* f32[nseg] device; WF_OUT_LINEAR */
    int nseg;             /* 1..4 row segments sharing one weight scale each (fused q|k|v, gate|up) */
    int seg_end[4];       /* exclusive end row of each segment (in the N numbering of this matrix) */
    float host_scale;     /* WF_OUT_COMPAT_F32: the Python float `scale` argument */
    int out_mode;
    int epi_flags;
    int out_dtype;        /* WF_F32 | WF_BF16 for WF_OUT_LINEAR */
    void* out;            /* [M, N] (or [M, N/2] with RELU2_MUL), row stride ldo elements */
    int ldo;
    const void* residual; /* bf16 [M,N], row stride ldo; WF_EPI_RESIDUAL */
    void* workspace;      /* wf_linear_workspace_bytes(); must be zero on first use, kernels keep it zero */
    size_t workspace_bytes;
    int algo;             /* 0 auto, 1 stream (M<=8), 2 mma tile, 3 tcgen05 */
} wf_linear_args;

size_t wf_linear_workspace_bytes(int M, int N, int K);
int wf_ternary_linear(const wf_linear_args* args, void* stream);

/* torch.ops.sgl_kernel.bitnet_gemv_cpu(packed_weights, activations, scale) -> f32[N]
 * (docs/research/sglang_optimization_plan.md:61-64; call sites bitnet_patch.py:69,
 * scripts/direct_inference_full.py:202, scripts/benchmark_kernels.py:46).
 * y[n] = scale * sum_k (w[n,k]+1) * x[k]; packed is the reference's I2_S [N,K/4]. */
int wf_bitnet_gemv(const uint8_t* packed_i2s, const int8_t* x, float scale, int N, int K, float* y, void* workspace,
                   size_t workspace_bytes, void* stream);
/* torch.ops.sgl_kernel.bitnet_gemm_cpu(packed, X[M,K], scale) -> f32[M,N] (…plan.md:66-70; bitnet_patch.py:73). */
int wf_bitnet_gemm(const uint8_t* packed_i2s, const int8_t* x, float scale, int M, int N, int K, float* y,
                   void* workspace, size_t workspace_bytes, void* stream);

/* wf_bitnet_gemm with no caller workspace and no allocation of any kind (M = 1 is bitnet_gemv).
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The split-K accumulators and tile
 * counters live in a file-scope __device__ array of WF_STATIC_WS_BYTES inside the library (csrc/no_alloc.cu), held by every
 * GPU that loads it until exit.  Same kernel and same bits as wf_bitnet_gemm.  Shapes whose scratch
 * (wf_linear_workspace_bytes(M, N, K) = 4*M*N + 4*ceil(N/16)*ceil(M/8) + 256 bytes) exceeds the array return
 * WF_ERR_UNSUPPORTED with the sizes in wf_last_error().  All calls on a device share the array and each call clears
 * the part it uses on `stream` first: calls that overlap on the GPU (different streams) race — serialise them. */
#define WF_STATIC_WS_BYTES (4u << 20)
size_t wf_static_workspace_max_bytes(void);
int wf_bitnet_gemm_static(const uint8_t* packed_i2s, const int8_t* x, float scale, int M, int N, int K, float* y,
                          void* stream);

/* ---- paged KV cache + decode attention (north-star item 5) ------------ */

/* Pool layout: [max_pages][L][2][Hkv][page_size][D] bf16 — one (page, layer, K|V, head) block is
 * page_size*D contiguous elements.  Token t of a sequence lives at page block_table[t / page_size],
 * slot t % page_size. */
typedef struct wf_kv_layout {
    void* pool;
    int max_pages, layers, kv_heads, page_size, head_dim;
    int dtype; /* WF_BF16 | WF_I8 | WF_F32 (f32 only for the KVCacheManager scatter/gather API). WF_I8: the reference's
                  KVCacheDtype.INT8 (kv_cache/kv_cache.py:198-228) at (token, head) granularity; k / v rows handed to
                  scatter / gather are bf16; pool = int8 region followed by the fp32 scale region (wf_kv_pool_bytes) */
} wf_kv_layout;

size_t wf_kv_pool_bytes(int max_pages, int layers, int kv_heads, int page_size, int head_dim, int dtype);

/* sgl_kernel.kvcache.KVCacheManager.scatter_kv / gather_kv (tests/test_kv_cache_cpp.py:104-127,
 * 328-335): k,v [T, Hkv*D] rows <-> (pages[t], slots[t], layer). */
int wf_kv_scatter(const wf_kv_layout* kv, const void* k, const void* v, const int32_t* pages, const int32_t* slots,
                  int T, int layer, void* stream);
int wf_kv_gather(const wf_kv_layout* kv, void* k_out, void* v_out, const int32_t* pages, const int32_t* slots, int T,
                 int layer, void* stream);

/* RoPE (half-split, bf16 sin/cos tables; scripts/direct_inference_full.py:28-70) on q and k of a
 * fused qkv row [B, (H+2Hkv)*D] (bf16), then append k,v at positions[b] of each sequence
 * (direct_inference_full.py:222-227; kv_cache.py:230-272 update). q_out bf16 [B,H,D]. */
int wf_rope_kv_append(const void* qkv, int B, int H, const wf_kv_layout* kv, int layer, const int32_t* positions,
                      const int32_t* block_table, int max_blocks, const void* cos_tab, const void* sin_tab,
                      int max_pos, void* q_out, void* stream);

/* attention_with_kv_cache (src/wrinklefree_inference/kv_cache/kv_cache.py:316-369) in its GQA form
 * (scripts/direct_inference_full.py:229-244): out[b,h,:] = softmax(q[b,h]·K[b,kvh]^T * scale) V[b,kvh]
 * over tokens [0, seq_lens[b]).  q bf16 [B,H,D]; out bf16 [B,H,D] (or f32).  Split-KV flash decode;
 * workspace from wf_attn_workspace_bytes. */
size_t wf_attn_workspace_bytes(int B, int H, int head_dim, int max_seq);
int wf_paged_decode_attn(const void* q, const wf_kv_layout* kv, int layer, const int32_t* block_table,
                         int max_blocks, const int32_t* seq_lens, int B, int H, float scale, void* out,
                         int out_dtype, void* workspace, size_t workspace_bytes, void* stream);

/* ---- glue around the linears -------------------------------------------- */

/* rms_norm (direct_inference_full.py:181-184) with the reference's bf16 rounding points, fused
 * with the following per-token activation quantisation (…:196).  x bf16 [M,K]; gamma bf16 [K].
 * Optional y_out (bf16 [M,K]) receives the normalised activations. */
int wf_rmsnorm_quant(const void* x, const void* gamma, float eps, int M, int K, void* y_out, int8_t* q,
                     float* scale_out, int32_t* rowsum_out, void* stream);

/* embedding row gather: out[b,:] = table[tokens[b],:] (bf16) (direct_inference_full.py:290) */
int wf_embed(const void* table, const int32_t* tokens, int B, int H, void* out, void* stream);

/* tied bf16 LM head + greedy argmax (direct_inference_full.py:297-300, :305-306):
 * logits[b,v] = sum_k h[b,k]*E[v,k] (fp32 accumulate, bf16 round like F.linear on bf16);
 * next[b] = argmax_v (lowest index wins ties, as torch.argmax).  logits_out optional (f32 [B,V]). */
size_t wf_lm_head_workspace_bytes(int B, int V);
int wf_lm_head_argmax(const void* h, const void* E, int B, int V, int K, float* logits_out, int32_t* next_out,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ---- decode engine (the whole step of scripts/direct_inference_full.py:267-302) --- */

typedef struct wf_model_config {
    int hidden, layers, heads, kv_heads, head_dim, ffn, vocab;
    float rope_theta, eps;
    int max_seq;    /* rope table length / per-sequence token capacity */
    int page_size;  /* KV page size in tokens */
    int max_batch;
    int max_pages;  /* KV pool pages */
    int tp_rank, tp_size; /* tensor parallel (1 = off); see INTEGRATION.md */
    int kv_dtype;         /* 0 or WF_BF16: bf16 KV pool; WF_I8: int8 KV pool (half the attention bytes) */
} wf_model_config;

typedef struct wf_engine wf_engine;

/* Allocates the resident state (tile-layout weights, KV pool, activations, workspaces). */
int wf_engine_create(const wf_model_config* cfg, wf_engine** out);
void wf_engine_destroy(wf_engine* e);
size_t wf_engine_device_bytes(const wf_engine* e);

/* Weight upload.  name is one of q,k,v,o,gate,up,down; i2s is a HOST or DEVICE pointer to the
 * reference-format I2_S matrix [N,K/4] (packed_bitnet_v1 `qweight`, cache/bitnet_converter.py:165-176);
 * the engine repacks into b200_tile_v1 (fusing q|k|v and interleaving gate|up). */
int wf_engine_set_linear(wf_engine* e, int layer, const char* name, const uint8_t* i2s, float weight_scale,
                         int is_host);
/* name: input_layernorm, post_attention_layernorm, attn_sub_norm, ffn_sub_norm (per layer, bf16 [dim]),
 * final_norm (layer = -1), embed (layer = -1, bf16 [V,H]). */
int wf_engine_set_tensor(wf_engine* e, int layer, const char* name, const void* data_bf16, size_t count, int is_host);
/* Fill every weight with seeded synthetic data ON THE DEVICE (bench / smoke). */
int wf_engine_init_synthetic(wf_engine* e, uint64_t seed);
int wf_engine_finalize_weights(wf_engine* e);

/* Sequence slots: block tables are managed by the engine (free-list page allocator with the
 * KVCacheManager semantics, tests/test_kv_cache_cpp.py:44-70). */
int wf_engine_reset(wf_engine* e);
/* mark slot b as holding `len` tokens of (synthetic/garbage) context: allocates pages only (bench) */
int wf_engine_set_context(wf_engine* e, int slot, int len, int fill_random);

/* One decode step for B sequences (slots 0..B-1): HOST token ids in, HOST next-token ids out
 * (greedy).  Copies inputs H2D, runs the (CUDA-graph-captured) step, copies results D2H,
 * synchronises.  positions advance by one per call.  This is the e2e entry point. */
int wf_engine_decode(wf_engine* e, const int32_t* host_tokens, int B, int32_t* host_next_tokens);
/* Same step, device-resident in/out, asynchronous on the engine stream (value timing). */
int wf_engine_decode_device(wf_engine* e, int B, int use_graph);
int wf_engine_sync(wf_engine* e);
void* wf_engine_stream(wf_engine* e);
/* Tensor parallel engines (cfg.tp_size > 1; one process per GPU, SURVEY §8e — the reference has no multi-GPU path,
 * this is the north star's "TP sharding of QKV/O/MLP with an all-reduce over NVLink").  Each rank creates its engine
 * with its tp_rank, uploads ITS shard through wf_engine_set_linear (q,k,v,gate,up: this rank's output rows; o,down: this
 * rank's input columns; norms/embedding replicated), then the ranks exchange wf_engine_tp_export() handles (CUDA IPC,
 * wf_engine_tp_handle_bytes() each, e.g. with torch.distributed.all_gather) and call wf_engine_tp_attach with all
 * tp_size handles in rank order.  After that wf_engine_decode* must be called by all ranks in lockstep; the
 * exchanges are peer-memory stores / REDs fused into the attention and linear kernels (csrc/tp_comm.cuh). */
size_t wf_engine_tp_handle_bytes(void);
int wf_engine_tp_export(wf_engine* e, void* handle_out);
int wf_engine_tp_attach(wf_engine* e, const void* handles);
/* debug/parity taps (device pointers into engine state) */
void* wf_engine_buffer(wf_engine* e, const char* name);
int wf_engine_launches_per_step(const wf_engine* e);
int wf_engine_time_kernel(wf_engine* e, const char* which, int B, int iters, float* ms_out);
/* Kernel-level GEMV measurement in the style of the reference's scripts/benchmark_kernels.py:28-66 (a loop of
 * bitnet_gemv calls, bytes = N*K/4 + M*K + out), but rotating over > L2 worth of independent weight copies so the
 * figure is an HBM figure.  flags: 1 = programmatic dependent launch between the calls, 2 = stream only (no math),
 * 4 = fused dequant epilogue (WF_OUT_LINEAR, bf16) instead of raw int32, 8 = force the mma.sync path, 16 / 32 = force 16 / 8 consumer warps, bits 8..15 = TMA
 * ring depth in 8 KB stages (0 = default).  ms_out = average per launch. */
int wf_bench_gemv(int N, int K, int M, int iters, int flags, float* ms_out);
/* The engine's streaming batched GEMM on its own (csrc/gemm_stream.cu; 1 <= M <= 32 activation rows): acc[m][n] += sum_k (w[n,k]+1) * q[m,k]
 * as exact integers in fp32 — the accumulator convention of bitnet_gemm (docs/research/sglang_optimization_plan.md:66-70) before its scale.
 * w_tile is b200_tile_v1 (wf_repack_i2s_to_tile), q is int8 [M][ldq], acc is f32 [M][N] that the caller zeroes.  For the parity tests and
 * kernel timing; iters > 1 repeats the launch (acc then holds iters times the sums). */
int wf_gemm_stream_acc(const uint8_t* w_tile, int N, int K, const int8_t* q, int M, int ldq, float* acc, int iters, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WF_BITNET_H */
