// Glue kernels around the ternary linears: embedding gather, tied bf16 LM head with
// greedy argmax, step bookkeeping, and on-device synthetic data generators for bench / smoke.
//
// Reference: scripts/direct_inference_full.py:290 (embedding row), :297-300 (F.linear(hidden, embed) in
// bf16), :305-306 (argmax sampling at temperature 0).
#include "common.cuh"
#include "tp_comm.cuh"
#include "wf_bitnet.h"
#include "kernels.cuh"
#include <stdlib.h>

namespace wf {

// ---------------------------------------------------------------------------------------------
__global__ void embed_kernel(const uint4* __restrict__ table, const int32_t* __restrict__ tokens, int Hv, uint4* __restrict__ out) {
    const int b = blockIdx.x;
    pdl_wait_prior();
    const int tok = tokens[b];
    for (int i = threadIdx.x; i < Hv; i += blockDim.x) out[(size_t)b * Hv + i] = table[(size_t)tok * Hv + i];
    pdl_launch_dependents();
}

// ---------------------------------------------------------------------------------------------
// LM head: logits[b, v] = sum_k h[b,k] * E[v,k]; E bf16 [V,K] streamed once from HBM (657 MB for the 2B
// model — more bytes than all ternary layers together).  CTA = 8 warps x 16 vocab rows,
// 64-wide k stages through a 4-stage cp.async ring, XOR-swizzled smem, ldmatrix + mma.m16n8k16 bf16.
constexpr int LM_ROWS = 128;
constexpr int LM_KC = 64;                       // k per stage (128 B per row)
constexpr int LM_STAGES = 3;                    // 72 KB per CTA at M = 64: three CTAs (24 warps) per SM

template <int NT>
__global__ void __launch_bounds__(256) lm_head_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ E,
                                                      int B, int V, int K, float* __restrict__ logits_out, int ldl, int v_off,
                                                      float* __restrict__ part_val, int* __restrict__ part_idx) {
    constexpr int MT = NT * 8;
    constexpr int E_STAGE = LM_ROWS * 128;
    constexpr int H_STAGE = MT * 128;
    constexpr int STAGE = E_STAGE + H_STAGE;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int v0 = blockIdx.x * LM_ROWS;
    const int m0 = blockIdx.y * MT;
    const int nchunks = K / LM_KC;

    // per-thread source rows / swizzled destinations of the E tile are loop-invariant: only the k offset moves per stage
    constexpr int EJ = LM_ROWS * 8 / 256;
    const __nv_bfloat16* e_src[EJ];
    int e_dst[EJ], e_nb[EJ];
#pragma unroll
    for (int j = 0; j < EJ; ++j) {
        const int g = tid + 256 * j;
        const int r = g >> 3, c = g & 7;
        const int v = v0 + r;
        e_nb[j] = (v < V) ? 16 : 0;
        e_src[j] = E + (size_t)(v < V ? v : 0) * K + c * 8;
        e_dst[j] = r * 128 + ((c ^ (r & 7)) << 4);
    }
    constexpr int HJ = (MT * 8 + 255) / 256;
    const __nv_bfloat16* h_src[HJ];
    int h_dst[HJ], h_nb[HJ];
#pragma unroll
    for (int j = 0; j < HJ; ++j) {
        const int g = tid + 256 * j;
        const int r = (g >> 3) % MT, c = g & 7;
        const int m = m0 + r;
        h_nb[j] = (m < B) ? 16 : 0;
        h_src[j] = h + (size_t)(m < B ? m : 0) * K + c * 8;
        h_dst[j] = r * 128 + ((c ^ (r & 7)) << 4);
    }
    auto load_stage = [&](int chunk, int stage) {
        unsigned char* es = smem + stage * STAGE;
        unsigned char* hs = es + E_STAGE;
#pragma unroll
        for (int j = 0; j < EJ; ++j) cp_async16_zfill(es + e_dst[j], e_src[j] + chunk * LM_KC, e_nb[j]);
#pragma unroll
        for (int j = 0; j < HJ; ++j)
            if (HJ * 256 == MT * 8 || tid + 256 * j < MT * 8) cp_async16_zfill(hs + h_dst[j], h_src[j] + chunk * LM_KC, h_nb[j]);
    };

    float acc[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i) { acc[i][0] = 0.f; acc[i][1] = 0.f; acc[i][2] = 0.f; acc[i][3] = 0.f; }

    pdl_wait_prior();
#pragma unroll
    for (int s = 0; s < LM_STAGES - 1; ++s) {
        if (s < nchunks) load_stage(s, s);
        cp_async_commit();
    }
    for (int c = 0; c < nchunks; ++c) {
        cp_async_wait<LM_STAGES - 2>();
        __syncthreads();
        {
            const int nc = c + LM_STAGES - 1;
            if (nc < nchunks) load_stage(nc, nc % LM_STAGES);
            cp_async_commit();
        }
        const unsigned char* es = smem + (c % LM_STAGES) * STAGE;
        const unsigned char* hs = es + E_STAGE;
        // A fragments for the 4 k-steps of this stage: rows warp*16 + (lane&15), granule 2*ks + (lane>>4)
        uint32_t a[4][4];
        {
            const int r = warp * 16 + (lane & 15);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const int cg = 2 * ks + (lane >> 4);
                ldmatrix_x4(a[ks][0], a[ks][1], a[ks][2], a[ks][3], es + r * 128 + ((cg ^ (r & 7)) << 4));
            }
        }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int r = nt * 8 + (lane & 7);
#pragma unroll
            for (int kp = 0; kp < 2; ++kp) {   // pairs of k-steps
                const int cg = 4 * kp + (lane >> 3);
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4(b0, b1, b2, b3, hs + r * 128 + ((cg ^ (r & 7)) << 4));
                mma_bf16_16816(acc[nt], a[2 * kp][0], a[2 * kp][1], a[2 * kp][2], a[2 * kp][3], b0, b1);
                mma_bf16_16816(acc[nt], a[2 * kp + 1][0], a[2 * kp + 1][1], a[2 * kp + 1][2], a[2 * kp + 1][3], b2, b3);
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();

    // logits tile -> smem [MT][LM_ROWS+1], rounded to bf16 like F.linear on bf16 tensors
    float* St = reinterpret_cast<float*>(smem);
    constexpr int SS = LM_ROWS + 1;
    {
        const int g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int ml = nt * 8 + 2 * t, nl = warp * 16 + g;
            St[ml * SS + nl] = bf16_round(acc[nt][0]);
            St[(ml + 1) * SS + nl] = bf16_round(acc[nt][1]);
            St[ml * SS + nl + 8] = bf16_round(acc[nt][2]);
            St[(ml + 1) * SS + nl + 8] = bf16_round(acc[nt][3]);
        }
    }
    __syncthreads();
    const int mcount = min(MT, B - m0);
    if (logits_out) {
        for (int i = tid; i < mcount * LM_ROWS; i += 256) {
            const int ml = i / LM_ROWS, nl = i % LM_ROWS;
            if (v0 + nl < V) logits_out[(size_t)(m0 + ml) * ldl + v0 + nl] = St[ml * SS + nl];
        }
    }
    // per batch row: max / lowest index over this CTA's 128 vocab rows (warp w handles rows w, w+8, ...)
    for (int ml = warp; ml < mcount; ml += 8) {
        float best = -INFINITY;
        int bi = 0x7fffffff;
        for (int nl = lane; nl < LM_ROWS; nl += 32) {
            const int v = v0 + nl;
            if (v < V) {
                const float x = St[ml * SS + nl];
                if (x > best || (x == best && v < bi)) { best = x; bi = v; }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) {
            part_val[(size_t)(m0 + ml) * gridDim.x + blockIdx.x] = best;
            part_idx[(size_t)(m0 + ml) * gridDim.x + blockIdx.x] = bi + v_off;   // global vocab index (vocab-split LM head)
        }
    }
    pdl_launch_dependents();
}

__global__ void __launch_bounds__(256) argmax_reduce_kernel(const float* __restrict__ part_val, const int* __restrict__ part_idx,
                                                            int nparts, int32_t* __restrict__ next, const TpLm tpl) {
    __shared__ float sv[8];
    __shared__ int si[8];
    const int b = blockIdx.x;
    pdl_wait_prior();
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < nparts; i += 256) {
        const float x = part_val[(size_t)b * nparts + i];
        const int v = part_idx[(size_t)b * nparts + i];
        if (x > best || (x == best && v < bi)) { best = x; bi = v; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w)
            if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
        if (tpl.tp.size > 1) {
            // vocab-split LM head: publish this rank's candidate on every rank; advance_kernel picks the winner
            for (int r = 0; r < tpl.tp.size; ++r) {
                float2* c = reinterpret_cast<float2*>(tpl.cand_peer[r]) + (size_t)b * TP_MAX + tpl.tp.rank;
                *c = make_float2(best, __int_as_float(bi));
            }
        } else {
            next[b] = bi;
        }
    }
    if (tpl.tp.size > 1) tp_signal_grid_done(tpl.tp, tpl.signal_flag, gridDim.x);
    pdl_launch_dependents();
}

// positions / seq_lens advance + feed the sampled token back (device-resident decode loop)
// (one CTA: B <= 64).  Under TP it first waits for every rank's LM-head candidate and reduces them (highest logit,
// lowest index on ties — torch.argmax).  It also closes the step: the device step counter drives the TP flags.
__global__ void advance_kernel(int32_t* positions, int32_t* seq_lens, int32_t* tokens, int32_t* next, int B, int feed_back, int* step,
                               const TpLm tpl) {
    pdl_wait_prior();
    if (tpl.tp.size > 1) {
        if (threadIdx.x < 32) tp_wait_flags_warp(tpl.tp, tpl.signal_flag);
        __syncthreads();
    }
    const int b = threadIdx.x;
    if (b < B) {
        if (tpl.tp.size > 1) {
            float best = -INFINITY;
            int bi = 0x7fffffff;
            const float2* c = reinterpret_cast<const float2*>(tpl.cand_peer[tpl.tp.rank]) + (size_t)b * TP_MAX;
            for (int r = 0; r < tpl.tp.size; ++r) {
                float2 v;
                asm volatile("ld.global.cg.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(c + r));
                const int vi = __float_as_int(v.y);
                if (v.x > best || (v.x == best && vi < bi)) { best = v.x; bi = vi; }
            }
            next[b] = bi;
        }
        positions[b] += 1;
        seq_lens[b] += 1;
        if (feed_back) tokens[b] = next[b];
    }
    __syncthreads();
    if (threadIdx.x == 0 && step) *step += 1;
}

// ---------------------------------------------------------------------------------------------
// synthetic data (bench / smoke only): splitmix64 counter hash
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// uniformly random ternary codes {0,1,2}, 16 per u32
__global__ void fill_ternary_kernel(uint32_t* out, size_t nwords, uint64_t seed) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) {
        uint64_t r = mix64(seed ^ (i * 0x2545F4914F6CDD1Dull));
        uint64_t r2 = mix64(r);
        uint32_t w = 0;
#pragma unroll
        for (int f = 0; f < 16; ++f) {
            const uint32_t bits = (uint32_t)((f < 8 ? r : r2) >> (8 * (f & 7))) & 0xff;
            w |= ((bits * 3u) >> 8) << (2 * f);
        }
        out[i] = w;
    }
}
// bf16 values mean + std * approx-normal (sum of 4 uniforms)
__global__ void fill_bf16_kernel(__nv_bfloat16* out, size_t n, uint64_t seed, float mean, float std) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint64_t r = mix64(seed ^ (i * 0x9E3779B97F4A7C15ull));
        float u = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) u += (float)((r >> (16 * j)) & 0xffff) * (1.0f / 65536.0f);
        out[i] = __float2bfloat16_rn(mean + std * (u - 2.0f) * 1.7320508f);
    }
}

int launch_embed(const void* table, const int32_t* tokens, int B, int H, void* out, cudaStream_t s) {
    embed_kernel<<<B, 128, 0, s>>>((const uint4*)table, tokens, H / 8, (uint4*)out);
    return launch_check(cudaGetLastError(), "embed");
}

int launch_advance_tp(int32_t* positions, int32_t* seq_lens, int32_t* tokens, int32_t* next, int B, int feed_back, int* step, const TpLm* tpl,
                      cudaStream_t s) {
    if (B > 128) { set_error("advance: B > 128"); return WF_ERR_UNSUPPORTED; }
    TpLm t = {};
    if (tpl) t = *tpl; else { t.tp.size = 1; t.signal_flag = -1; }
    advance_kernel<<<1, 128, 0, s>>>(positions, seq_lens, tokens, next, B, feed_back, step, t);
    return launch_check(cudaGetLastError(), "advance");
}
int launch_advance(int32_t* positions, int32_t* seq_lens, int32_t* tokens, const int32_t* next, int B, int feed_back, cudaStream_t s) {
    return launch_advance_tp(positions, seq_lens, tokens, const_cast<int32_t*>(next), B, feed_back, nullptr, nullptr, s);
}

int launch_fill_ternary(void* out, size_t bytes, uint64_t seed, cudaStream_t s) {
    fill_ternary_kernel<<<sm_count() * 8, 256, 0, s>>>((uint32_t*)out, bytes / 4, seed);
    return launch_check(cudaGetLastError(), "fill_ternary");
}
__global__ void fill_f32_kernel(float* out, size_t n, float v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = v;
}
int launch_fill_f32(float* out, size_t n, float value, cudaStream_t s) {
    fill_f32_kernel<<<sm_count() * 4, 256, 0, s>>>(out, n, value);
    return launch_check(cudaGetLastError(), "fill_f32");
}
// pseudo-random INTEGERS in [-amp, amp] held in fp32: a stand-in for a linear's raw accumulator (bench / timing set-up only)
__global__ void fill_randint_f32_kernel(float* out, size_t n, uint64_t seed, int amp) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint64_t z = (i + 1) * 0x9E3779B97F4A7C15ull + seed;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        out[i] = (float)((int)(z % (uint64_t)(2 * amp + 1)) - amp);
    }
}
int launch_fill_randint_f32(float* out, size_t n, uint64_t seed, int amp, cudaStream_t s) {
    fill_randint_f32_kernel<<<sm_count() * 4, 256, 0, s>>>(out, n, seed, amp);
    return launch_check(cudaGetLastError(), "fill_randint_f32");
}
int launch_fill_bf16(void* out, size_t n, uint64_t seed, float mean, float std, cudaStream_t s) {
    fill_bf16_kernel<<<sm_count() * 8, 256, 0, s>>>((__nv_bfloat16*)out, n, seed, mean, std);
    return launch_check(cudaGetLastError(), "fill_bf16");
}

template <int NT>
static int launch_lm(const void* h, const void* E, int B, int V, int K, float* logits_out, int ldl, int v_off, float* pv, int* pi, int nparts, cudaStream_t s) {
    constexpr int MT = NT * 8;
    constexpr int STAGE = LM_ROWS * 128 + MT * 128;
    constexpr int SM_PIPE = STAGE * LM_STAGES;
    constexpr int SM_EPI = MT * (LM_ROWS + 1) * 4;
    constexpr int SMEM = SM_PIPE > SM_EPI ? SM_PIPE : SM_EPI;
    auto kern = lm_head_kernel<NT>;
    static DevOnce attr;
    if (attr.need()) {
        if (cuda_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM), "lm smem")) return WF_ERR_CUDA;
    }
    dim3 grid(nparts, (B + MT - 1) / MT);
    kern<<<grid, 256, SMEM, s>>>((const __nv_bfloat16*)h, (const __nv_bfloat16*)E, B, V, K, logits_out, ldl, v_off, pv, pi);
    return launch_check(cudaGetLastError(), "lm_head");
}

int launch_lm_head_argmax(const void* h, const void* E, int B, int V, int K, float* logits_out, int32_t* next_out, void* workspace,
                          cudaStream_t s) {
    return launch_lm_head_argmax_tp(h, E, B, V, K, logits_out, V, 0, next_out, workspace, nullptr, s);
}

// E points at this rank's first vocab row, V = its row count, v_off = that row's global index, ldl = logits row stride
int launch_lm_head_argmax_tp(const void* h, const void* E, int B, int V, int K, float* logits_out, int ldl, int v_off, int32_t* next_out,
                             void* workspace, const TpLm* tpl, cudaStream_t s) {
    const int nparts = (V + LM_ROWS - 1) / LM_ROWS;
    float* pv = (float*)workspace;
    int* pi = (int*)(pv + (size_t)B * nparts);
    int rc;
    // 17..64 tokens: tcgen05 kind::f16 kernel (lm_head_tc.cu) — operands straight from shared memory, accumulators in TMEM; the mma.sync
    // kernel below re-reads the 64-token activation tile with every warp and stops at 65 % of the HBM peak at batch 64 (WF_LM_TC=0: off)
    static const int lm_tc = getenv("WF_LM_TC") ? atoi(getenv("WF_LM_TC")) : 1;
    if (lm_tc && lm_head_tc_usable(B, V, K, h, E)) rc = launch_lm_head_tc(h, E, B, V, K, logits_out, ldl, v_off, pv, pi, nparts, s, false);
    else if (B <= 8) rc = launch_lm<1>(h, E, B, V, K, logits_out, ldl, v_off, pv, pi, nparts, s);
    else if (B <= 16) rc = launch_lm<2>(h, E, B, V, K, logits_out, ldl, v_off, pv, pi, nparts, s);
    else if (B <= 32) rc = launch_lm<4>(h, E, B, V, K, logits_out, ldl, v_off, pv, pi, nparts, s);
    else rc = launch_lm<8>(h, E, B, V, K, logits_out, ldl, v_off, pv, pi, nparts, s);
    if (rc) return rc;
    if (next_out) {
        TpLm t = {};
        if (tpl) t = *tpl; else { t.tp.size = 1; t.signal_flag = -1; }
        argmax_reduce_kernel<<<B, 256, 0, s>>>(pv, pi, nparts, next_out, t);
        return launch_check(cudaGetLastError(), "argmax_reduce");
    }
    return WF_OK;
}

}  // namespace wf

using namespace wf;

extern "C" int wf_embed(const void* table, const int32_t* tokens, int B, int H, void* out, void* stream) {
    if (!table || !tokens || !out || B <= 0 || H <= 0 || H % 8) { set_error("wf_embed: bad arguments (H %% 8 == 0)"); return WF_ERR_INVALID; }
    return launch_embed(table, tokens, B, H, out, (cudaStream_t)stream);
}

extern "C" size_t wf_lm_head_workspace_bytes(int B, int V) {
    const size_t nparts = (size_t)(V + LM_ROWS - 1) / LM_ROWS;
    return (size_t)(B > 0 ? B : 1) * nparts * 8 + 256;
}

extern "C" int wf_lm_head_argmax(const void* h, const void* E, int B, int V, int K, float* logits_out, int32_t* next_out,
                                 void* workspace, size_t workspace_bytes, void* stream) {
    if (!h || !E || !workspace) { set_error("wf_lm_head_argmax: null pointer"); return WF_ERR_INVALID; }
    if (B <= 0 || V <= 0 || K <= 0 || K % LM_KC) { set_error("wf_lm_head_argmax: bad shape (K %% 64 == 0)"); return WF_ERR_INVALID; }
    if (B > 64) { set_error("wf_lm_head_argmax: B > 64 not supported"); return WF_ERR_UNSUPPORTED; }
    if (workspace_bytes < wf_lm_head_workspace_bytes(B, V)) { set_error("wf_lm_head_argmax: workspace too small"); return WF_ERR_INVALID; }
    return launch_lm_head_argmax(h, E, B, V, K, logits_out, next_out, workspace, (cudaStream_t)stream);
}

namespace wf {
template <int NT> static void lm_set_attr() {
    constexpr int MT = NT * 8;
    constexpr int SM_PIPE = (LM_ROWS * 128 + MT * 128) * LM_STAGES;
    constexpr int SM_EPI = MT * (LM_ROWS + 1) * 4;
    cudaFuncSetAttribute(lm_head_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_PIPE > SM_EPI ? SM_PIPE : SM_EPI);
}
void lm_init_attrs() { lm_set_attr<1>(); lm_set_attr<2>(); lm_set_attr<4>(); lm_set_attr<8>(); }
}
