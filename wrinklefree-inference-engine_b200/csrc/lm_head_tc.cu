// Tied bf16 LM head for 17..64 tokens on the 5th-generation tensor cores (tcgen05 kind::f16, bf16 in, fp32 accumulate).
//
//   logits[m, v] = sum_k h[m, k] * E[v, k]          (F.linear(hidden, embed), scripts/direct_inference_full.py:297-300;
//                                                    greedy argmax :305-306)
// E is the largest matrix of the step (657 MB for the 2B model: more bytes than all ternary layers together), read once:
// the kernel is HBM-bound and everything here exists to keep the stream running.  The mma.sync kernel (glue.cu) reaches 93 % of
// the HBM peak at batch 1 but 65 % at batch 64: every warp re-reads the whole 64-token activation tile through ldmatrix (80 KB of
// shared-memory reads per 24 KB weight stage, profiles/r01c_full_lmhead_bs64).  Here the tensor core reads both operands straight
// from shared memory and nothing else touches them:
//   * persistent CTA per SM, 6 warps: 4 epilogue (one per TMEM lane quarter), 1 TMA producer, 1 MMA issuer;
//   * a stage = 64 k of 128 vocab rows (16 KB) + 64 k of 64 tokens (8 KB), both by cp.async.bulk.tensor.2d with the 128-byte
//     swizzle the UMMA descriptors expect; rows beyond V / tokens beyond B are zero-filled by the TMA unit; 8 stages = 192 KB;
//   * D = 128 vocab rows (TMEM lanes) x 64 tokens (columns), fp32, double-buffered (2 x 64 columns): the epilogue of tile i
//     (logit stores + per-token argmax over the tile's rows) overlaps the stream of tile i + 1;
//   * one thread chosen by elect.sync issues 4 x tcgen05.mma (M128 N64 K16) per stage and commits the stage back to the producer;
//   * epilogue: tcgen05.ld 32x32b.x32 twice -> a thread holds its vocab row's 64 logits; a warp's store of ONE token covers 32
//     consecutive v (one 128-byte line); the per-token argmax goes warp shuffle -> shared memory -> one candidate per (token, tile)
//     in the same [B][tiles] layout the mma.sync kernel writes, so argmax_reduce_kernel is shared.
#include "common.cuh"
#include "tp_comm.cuh"
#include "wf_bitnet.h"
#include "kernels.cuh"
#include <cuda.h>

namespace wf {

constexpr int LT2_ROWS = 128;                      // vocab rows per tile (UMMA M)
constexpr int LT2_TOK = 64;                        // tokens (UMMA N)
constexpr int LT2_KC = 64;                         // k per stage: 128 bytes of bf16 = one swizzle atom row
constexpr int LT2_STAGES = 8;
constexpr int LT2_E_BYTES = LT2_ROWS * 128;        // 16 KB
constexpr int LT2_H_BYTES = LT2_TOK * 128;         // 8 KB
constexpr int LT2_STAGE_BYTES = LT2_E_BYTES + LT2_H_BYTES;
constexpr int LT2_SMEM = LT2_STAGES * LT2_STAGE_BYTES + 1024 + 4 * LT2_TOK * 8;
constexpr int LT2_THREADS = 6 * 32;
constexpr int LT2_TMEM_COLS = 128;

struct LmTcParams {
    int B, V, K, tiles;
    float* logits;
    int ldl, v_off;
    float* part_val;
    int* part_idx;
};

__device__ __forceinline__ uint64_t lm_smem_desc_sw128(uint32_t saddr) {   // K-major tile, 128-byte rows, 8-row groups 1024 B apart
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffffu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// D f32, A bf16 (K-major), B bf16 (K-major), N, M
constexpr uint32_t lm_idesc(int m, int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void lm_mma_bf16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void lm_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
#define LM_LD32(taddr, v)                                                                                              \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"       \
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"                            \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),      \
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), \
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),           \
                   "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),           \
                   "=r"(v[30]), "=r"(v[31])                                                                             \
                 : "r"(taddr) : "memory")

__global__ void __launch_bounds__(LT2_THREADS, 1) lm_head_tc_kernel(const __grid_constant__ CUtensorMap tmap_e, const __grid_constant__ CUtensorMap tmap_h,
                                                                    const LmTcParams p) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[LT2_STAGES], empty_bar[LT2_STAGES], tfull_bar[2], tempty_bar[2];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
    float2* cand = reinterpret_cast<float2*>(smem_raw + (ring - smem_u32(smem_raw)) + LT2_STAGES * LT2_STAGE_BYTES);   // [4 warps][64 tokens] {value, index}
    const int my_tiles = (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int kchunks = p.K / LT2_KC;
    const int n_items = my_tiles * kchunks;

    if (tid == 0) {
        for (int s = 0; s < LT2_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 4); }
        fence_mbar_init();
    }
    if (warp == 5) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(LT2_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;

    if (warp == 4) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_e) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_h) : "memory");
            // the embedding rows do not depend on the previous kernel: the first stages' weights are in flight before the normalised
            // hidden state exists; the activation boxes of those stages follow once it does
            int it = 0, pre = 0;
            const int npre = n_items < LT2_STAGES ? n_items : LT2_STAGES;
            for (; pre < npre; ++pre) {
                const int tl = pre / kchunks, kc = pre - tl * kchunks;
                const int tile = blockIdx.x + tl * gridDim.x;
                mbar_expect_tx(&full_bar[pre], LT2_STAGE_BYTES);
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(ring + pre * LT2_STAGE_BYTES), "l"(&tmap_e), "r"(kc * LT2_KC), "r"(tile * LT2_ROWS), "r"(smem_u32(&full_bar[pre])) : "memory");
            }
            pdl_wait_prior();
            for (int i = 0; i < npre; ++i) {
                const int tl = i / kchunks, kc = i - tl * kchunks;
                (void)tl;
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(ring + i * LT2_STAGE_BYTES + LT2_E_BYTES), "l"(&tmap_h), "r"(kc * LT2_KC), "r"(0), "r"(smem_u32(&full_bar[i])) : "memory");
            }
            for (it = npre; it < n_items; ++it) {
                const int s = it % LT2_STAGES;
                const int tl = it / kchunks, kc = it - tl * kchunks;
                const int tile = blockIdx.x + tl * gridDim.x;
                lm_wait(smem_u32(&empty_bar[s]), (uint32_t)(((it / LT2_STAGES) & 1) ^ 1));
                mbar_expect_tx(&full_bar[s], LT2_STAGE_BYTES);
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(ring + s * LT2_STAGE_BYTES), "l"(&tmap_e), "r"(kc * LT2_KC), "r"(tile * LT2_ROWS), "r"(smem_u32(&full_bar[s])) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(ring + s * LT2_STAGE_BYTES + LT2_E_BYTES), "l"(&tmap_h), "r"(kc * LT2_KC), "r"(0), "r"(smem_u32(&full_bar[s])) : "memory");
            }
        }
    } else if (warp == 5) {
        // ------------------------------------------------------------ MMA issuer: one elected thread runs the whole loop
        constexpr uint32_t idesc = lm_idesc(LT2_ROWS, LT2_TOK);
        const uint64_t adesc0 = lm_smem_desc_sw128(ring);
        const uint64_t bdesc0 = lm_smem_desc_sw128(ring + LT2_E_BYTES);
        const uint32_t full0 = smem_u32(&full_bar[0]), empty0 = smem_u32(&empty_bar[0]), tfull0 = smem_u32(&tfull_bar[0]);
        uint32_t elected;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(elected));
        if (elected) {
            uint32_t ph = 0;
            int kc = 0, tl = 0;
            uint32_t d_tmem = tmem;
            for (int i0 = 0; i0 < n_items; i0 += LT2_STAGES) {
#pragma unroll
                for (int s = 0; s < LT2_STAGES; ++s) {
                    if (i0 + s < n_items) {
                        if (kc == 0) {
                            const int buf = tl & 1;
                            lm_wait(smem_u32(&tempty_bar[buf]), (uint32_t)(((tl >> 1) & 1) ^ 1));
                            d_tmem = tmem + buf * LT2_TOK;
                        }
                        lm_wait(full0 + s * 8, ph);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint64_t ad = adesc0 + (uint64_t)(s * (LT2_STAGE_BYTES >> 4));
                        const uint64_t bd = bdesc0 + (uint64_t)(s * (LT2_STAGE_BYTES >> 4));
                        lm_mma_bf16_ss(d_tmem, ad, bd, idesc, kc != 0 ? 1u : 0u);          // 16 k per instruction = 32 bytes = +2 in the descriptor
                        lm_mma_bf16_ss(d_tmem, ad + 2, bd + 2, idesc, 1u);
                        lm_mma_bf16_ss(d_tmem, ad + 4, bd + 4, idesc, 1u);
                        lm_mma_bf16_ss(d_tmem, ad + 6, bd + 6, idesc, 1u);
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(empty0 + s * 8) : "memory");
                        if (++kc == kchunks) {
                            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tfull0 + (tl & 1) * 8) : "memory");
                            kc = 0; ++tl;
                        }
                    }
                }
                ph ^= 1u;
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------ epilogue warps: TMEM -> logits + per-token argmax candidates
        const int q = warp;
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        for (int tl = 0; tl < my_tiles; ++tl) {
            const int tile = blockIdx.x + tl * gridDim.x;
            const int buf = tl & 1;
            const int v = tile * LT2_ROWS + q * 32 + lane;                 // this thread's vocab row (inside this rank's slice)
            const bool v_ok = v < p.V;
            lm_wait(smem_u32(&tfull_bar[buf]), (uint32_t)((tl >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
            for (int c = 0; c < LT2_TOK / 32; ++c) {
                uint32_t r[32];
                LM_LD32(tmem + lane_base + buf * LT2_TOK + c * 32, r);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (c == LT2_TOK / 32 - 1) {
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[buf]);
                }
                const int mcnt = min(32, p.B - c * 32);                    // tokens of this half that exist (warp-uniform)
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    if (j < mcnt) {
                        const float x = bf16_round(__uint_as_float(r[j]));   // rounded to bf16 like F.linear on bf16 tensors (and like glue.cu)
                        const int m = c * 32 + j;
                        if (v_ok && p.logits) p.logits[(size_t)m * p.ldl + v] = x;
                        // argmax over the warp's 32 rows: highest logit, lowest index on ties (torch.argmax)
                        float best = v_ok ? x : -INFINITY;
                        int bi = v;
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
                        }
                        if (lane == 0) cand[q * LT2_TOK + m] = make_float2(best, __int_as_float(bi));
                    }
                }
            }
            // the four warps' candidates of a token meet in shared memory
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (tid < p.B) {
                float best = -INFINITY;
                int bi = 0x7fffffff;
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const float2 cv = cand[w * LT2_TOK + tid];
                    const int ci = __float_as_int(cv.y);
                    if (cv.x > best || (cv.x == best && ci < bi)) { best = cv.x; bi = ci; }
                }
                p.part_val[(size_t)tid * p.tiles + tile] = best;
                p.part_idx[(size_t)tid * p.tiles + tile] = bi + p.v_off;   // global vocab index (vocab-split LM head)
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");                 // cand is rewritten by the next tile
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    pdl_launch_dependents();
    if (warp == 5) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(LT2_TMEM_COLS) : "memory");
    }
}

typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn2 g_lm_encode = nullptr;
static DevOnce g_lm_attr;

void lm_head_tc_init_attrs() {
    if (g_lm_attr.need()) cudaFuncSetAttribute(lm_head_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LT2_SMEM);
    if (!g_lm_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            g_lm_encode = (EncodeTiledFn2)fn;
    }
}

bool lm_head_tc_usable(int B, int V, int K, const void* h, const void* E) {
    // from 17 tokens: measured 134.7 -> 108.1 us at batch 32 and 156.0 -> 117.4 us at batch 64 (2B: 657 MB), but 110.8 -> 118.0 us at batch 16,
    // where the mma.sync kernel's 16-token tile is cheap enough (profiles/r02aa_summary.txt)
    return B > 16 && B <= LT2_TOK && K % LT2_KC == 0 && ((uintptr_t)h & 15) == 0 && ((uintptr_t)E & 15) == 0 && V >= 1;
}

// same contract as launch_lm<NT> in glue.cu: logits (optional) + one argmax candidate per (token, 128-row tile)
int launch_lm_head_tc(const void* h, const void* E, int B, int V, int K, float* logits_out, int ldl, int v_off, float* pv, int* pi, int nparts,
                      cudaStream_t s, bool pdl) {
    lm_head_tc_init_attrs();
    if (!g_lm_encode) { set_error("lm_head_tc: cuTensorMapEncodeTiled unavailable"); return WF_ERR_CUDA; }
    CUtensorMap te, th;
    {
        const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)V};
        const cuuint64_t gstride[1] = {(cuuint64_t)K * 2};
        const cuuint32_t box[2] = {LT2_KC, LT2_ROWS};
        const cuuint32_t estr[2] = {1, 1};
        const CUresult cr = g_lm_encode(&te, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)E, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) { set_error("lm_head_tc: tensor map (E) failed (%d)", (int)cr); return WF_ERR_CUDA; }
    }
    {
        const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)B};
        const cuuint64_t gstride[1] = {(cuuint64_t)K * 2};
        const cuuint32_t box[2] = {LT2_KC, LT2_TOK};
        const cuuint32_t estr[2] = {1, 1};
        const CUresult cr = g_lm_encode(&th, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)h, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) { set_error("lm_head_tc: tensor map (h) failed (%d)", (int)cr); return WF_ERR_CUDA; }
    }
    LmTcParams p;
    p.B = B; p.V = V; p.K = K; p.tiles = nparts;
    p.logits = logits_out; p.ldl = ldl; p.v_off = v_off; p.part_val = pv; p.part_idx = pi;
    cudaLaunchConfig_t cfg = {};
    const int sms = sm_count();
    cfg.gridDim = dim3(nparts < sms ? nparts : sms);
    cfg.blockDim = dim3(LT2_THREADS);
    cfg.dynamicSmemBytes = LT2_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, lm_head_tc_kernel, te, th, p), "lm_head_tc");
}

}  // namespace wf
