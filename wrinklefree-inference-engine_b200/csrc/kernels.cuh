// Cross-file launcher declarations used by the decode engine.
#pragma once
#include "linear_common.cuh"
#include "tp_comm.cuh"
namespace wf {
// vocab-split LM head under tensor parallelism: where every rank's argmax candidate goes
struct TpLm {
    TpComm tp;
    void* cand_peer[TP_MAX];   // {f32 logit, i32 index} [MB][TP_MAX] on every rank
    int signal_flag;
};
int launch_lm_head_argmax_tp(const void* h, const void* E, int B, int V, int K, float* logits_out, int ldl, int v_off, int32_t* next_out,
                             void* workspace, const TpLm* tpl, cudaStream_t s);
int launch_advance_tp(int32_t* positions, int32_t* seq_lens, int32_t* tokens, int32_t* next, int B, int feed_back, int* step, const TpLm* tpl,
                      cudaStream_t s);
int launch_rmsnorm_quant(const void* x, const void* gamma, float eps, int M, int K, void* y_out, int8_t* q, float* scale_out,
                         int32_t* rowsum_out, cudaStream_t s, bool pdl);
int rope_append(const void* qkv, int B, int H, int Hkv, int D, void* pool, int layers, int layer, int page_size,
                const int32_t* positions, const int32_t* block_table, int max_blocks, const void* cos_tab, const void* sin_tab,
                int max_pos, void* q_out, cudaStream_t s, bool pdl, float* kv_scales = nullptr);   // kv_scales: int8 pool
int attn_decode(const void* q, const void* pool, int layers, int layer, int Hkv, int page_size, const int32_t* block_table,
                int max_blocks, const int32_t* seq_lens, int B, int H, float scale, void* out, int out_dtype, int max_seq_hint,
                void* ws, size_t ws_bytes, cudaStream_t s, bool pdl, float* kv_scales = nullptr);
int launch_gemm_tc_red(const uint8_t* w_tile, int N, int K, const int8_t* q, int M, int ldq, int k_off, float* acc, int acc_ld,
                       cudaStream_t s, bool pdl);
int launch_embed(const void* table, const int32_t* tokens, int B, int H, void* out, cudaStream_t s);
int launch_advance(int32_t* positions, int32_t* seq_lens, int32_t* tokens, const int32_t* next, int B, int feed_back, cudaStream_t s);
int launch_fill_ternary(void* out, size_t bytes, uint64_t seed, cudaStream_t s);
int launch_fill_bf16(void* out, size_t n, uint64_t seed, float mean, float std, cudaStream_t s);
int launch_fill_f32(float* out, size_t n, float value, cudaStream_t s);
int launch_fill_randint_f32(float* out, size_t n, uint64_t seed, int amp, cudaStream_t s);
int launch_lm_head_argmax(const void* h, const void* E, int B, int V, int K, float* logits_out, int32_t* next_out, void* workspace,
                          cudaStream_t s);
}  // namespace wf
namespace wf {
// set every opt-in shared-memory attribute up front (function attributes must not change during stream capture)
void gemm_mma_init_attrs();
void gemv_stream_init_attrs();
void attn_init_attrs();
void lm_init_attrs();
void quant_init_attrs();
void gemm_tc_init_attrs();
void lm_head_tc_init_attrs();
inline void init_all_kernel_attrs() { gemm_mma_init_attrs(); gemv_stream_init_attrs(); attn_init_attrs(); lm_init_attrs(); quant_init_attrs(); gemm_tc_init_attrs(); lm_head_tc_init_attrs(); }
// tcgen05 LM head for 9..64 tokens (lm_head_tc.cu); same outputs as the mma.sync kernel in glue.cu
bool lm_head_tc_usable(int B, int V, int K, const void* h, const void* E);
int launch_lm_head_tc(const void* h, const void* E, int B, int V, int K, float* logits_out, int ldl, int v_off, float* pv, int* pi, int nparts,
                      cudaStream_t s, bool pdl);
}
