// Prefill / large-batch ternary GEMM on the 5th-generation tensor cores (tcgen05, kind::i8) — north-star item 4.
//
//   S[m,n] = sum_k code[n,k] * a[m,k],  code = w+1 in {0,1,2} (u8), a int8, exact int32 accumulation
//   (what bitnet_gemm computes row by row with ggml_vec_dot_i2_i8_s, extern/BitNet/src/ggml-bitnet-mad.cpp:94-186;
//    signature src/wrinklefree_inference/kernels/bitnet_patch.py:73).  Used where the shape is a real dense
//   contraction (M >= 128 tokens: prefill, 8 x 2048 tokens => 6.8e13 int8 MACs for the 2B shape).
//
// B200 mapping (one persistent CTA per SM, 10 warps, warp-specialised):
//   * D tile = 128 weight rows (UMMA M = TMEM lanes) x 128 tokens (UMMA N), int32, double-buffered in TMEM
//     (2 x 128 columns) so the epilogue of tile i overlaps the main loop of tile i+1;
//   * the WEIGHT operand never touches shared memory: four "unpack" warps (one per TMEM lane quarter) read
//     their row's 32 packed bytes per 128-k block straight from global/L2 into registers (prefetched 4 blocks
//     ahead), expand 2-bit -> u8 with shift+mask and write the 128 codes into a TMEM operand ring
//     (tcgen05.st 32x32b.x32, 8 stages x 32 columns).  tcgen05.mma reads A from TMEM ("TS" form), so shared
//     memory only carries the activation operand: 64 B/clk/SM at the int8 peak instead of >128 B/clk for an
//     SS kernel that would also have to write the unpacked weights through shared memory;
//   * int8 activations [M,K] are the B operand: TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into an
//     8-stage x 16 KB ring; one mbarrier per stage collects the TMA transaction bytes AND the four unpack
//     warps' arrivals, one tcgen05.commit per stage releases both halves;
//   * one elected thread issues 4 x tcgen05.mma.cta_group::1.kind::i8 (M128 N128 K32) per stage;
//   * epilogue warps: tcgen05.ld 32x32b.x32 (lane = weight row, columns = tokens) -> the shared linear
//     epilogue (Sigma-a correction, scales, residual, relu^2*up) -> stores that are coalesced along n.
#include "linear_common.cuh"
#include <cuda.h>

namespace wf {

constexpr int TC_BM = 128;                         // weight rows per tile
constexpr int TC_BN = 128;                         // tokens per tile
constexpr int TC_BK = 128;                         // k per stage (one I2_S block)
constexpr int TC_STAGES = 8;
constexpr int TC_STAGE_BYTES = TC_BN * TC_BK;      // 16 KB of int8 activations
constexpr int TC_A_COLS = TC_BK / 4;               // TMEM columns per weight stage (4 codes per 32-bit cell)
constexpr int TC_A_COL0 = 2 * TC_BN;               // weight ring starts after the two accumulator buffers
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_PF = 4;                           // weight prefetch distance (stages, in registers)
constexpr int TC_THREADS = 320;                    // warps 0-3 epilogue, 4-7 unpack, 8 TMA, 9 MMA
constexpr int TC_SMEM = TC_STAGES * TC_STAGE_BYTES + 1024;

struct TcParams {
    const uint8_t* w;
    int w_layout;
    int M, N, K, KB;
    int tiles_n, tiles_m;
    EpiParams ep;
};

// ---- tcgen05 / TMA wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]; u8 x s8 -> s32
__device__ __forceinline__ void tc_mma_i8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// K-major operand tile with 128-byte rows, 128B swizzle: 8-row groups are 1024 B apart
__device__ __forceinline__ uint64_t tc_smem_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffffu) >> 4);        // start address
    d |= (uint64_t)1 << 16;                           // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                 // stride byte offset
    d |= (uint64_t)1 << 46;                           // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                           // SWIZZLE_128B
    return d;
}
// instruction descriptor: D s32, A u8 (K-major), B s8 (K-major), N, M
constexpr uint32_t tc_idesc(int m, int n) {
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

#define TC_ST32(taddr, v)                                                                                              \
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16," \
                 "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"                                  \
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),   \
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),          \
                 "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),        \
                 "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory")

#define TC_LD32(taddr, v)                                                                                              \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"       \
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"                            \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),      \
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), \
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),           \
                   "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),           \
                   "=r"(v[30]), "=r"(v[31])                                                                             \
                 : "r"(taddr) : "memory")

__global__ void __launch_bounds__(TC_THREADS, 1) ternary_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const TcParams p) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[TC_STAGES];    // TMA bytes + 4 unpack-warp arrivals
    __shared__ __align__(8) uint64_t empty_bar[TC_STAGES];   // tcgen05.commit: MMAs of the stage retired
    __shared__ __align__(8) uint64_t tfull_bar[2];           // accumulator buffer complete
    __shared__ __align__(8) uint64_t tempty_bar[2];          // accumulator buffer drained by the 4 epilogue warps
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;   // 128B-swizzle atoms need 1024-B alignment
    const int total_tiles = p.tiles_n * p.tiles_m;
    const int my_tiles = (total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const long long n_items = (long long)my_tiles * p.KB;

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { mbar_init(&full_bar[s], 5); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 4); }
        fence_mbar_init();
    }
    if (warp == 9) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(TC_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;

    if (warp == 8) {
        // ------------------------------------------------------------ TMA producer: int8 activations
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_q) : "memory");
            long long it = 0;
            for (int tl = 0; tl < my_tiles; ++tl) {
                const int tile = blockIdx.x + tl * gridDim.x;
                const int m_blk = tile / p.tiles_n;
                for (int kb = 0; kb < p.KB; ++kb, ++it) {
                    const int s = (int)(it % TC_STAGES);
                    mbar_wait(&empty_bar[s], (uint32_t)(((it / TC_STAGES) & 1) ^ 1));
                    mbar_expect_tx(&full_bar[s], TC_STAGE_BYTES);
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                                 ::"r"(ring + s * TC_STAGE_BYTES), "l"(&tmap_q), "r"(kb * TC_BK), "r"(m_blk * TC_BN), "r"(smem_u32(&full_bar[s])) : "memory");
                }
            }
        }
    } else if (warp == 9) {
        // ------------------------------------------------------------ MMA issuer (one thread)
        if (lane == 0) {
            constexpr uint32_t idesc = tc_idesc(TC_BM, TC_BN);
            long long it = 0;
            for (int tl = 0; tl < my_tiles; ++tl) {
                const int buf = tl & 1;
                mbar_wait(&tempty_bar[buf], (uint32_t)(((tl >> 1) & 1) ^ 1));
                tc_fence_after();
                const uint32_t d_tmem = tmem + buf * TC_BN;
                for (int kb = 0; kb < p.KB; ++kb, ++it) {
                    const int s = (int)(it % TC_STAGES);
                    mbar_wait(&full_bar[s], (uint32_t)((it / TC_STAGES) & 1));
                    tc_fence_after();
                    const uint32_t a_tmem = tmem + TC_A_COL0 + s * TC_A_COLS;
                    const uint32_t b_smem = ring + s * TC_STAGE_BYTES;
#pragma unroll
                    for (int k4 = 0; k4 < TC_BK / 32; ++k4)
                        tc_mma_i8_ts(d_tmem, a_tmem + k4 * 8, tc_smem_desc_sw128(b_smem + k4 * 32), idesc, (kb | k4) != 0 ? 1u : 0u);
                    tc_commit(&empty_bar[s]);
                }
                tc_commit(&tfull_bar[buf]);
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------ unpack warps: 2-bit weights -> u8 codes -> TMEM
        const int q = warp & 3;
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        const int K4 = p.K / 4;
        uint32_t pre[TC_PF][8];
        auto load_item = [&](long long item, uint32_t (&r)[8]) {
            if (item >= n_items) return;
            const int tl = (int)(item / p.KB), kb = (int)(item - (long long)tl * p.KB);
            const int tile = blockIdx.x + tl * gridDim.x;
            const int n_blk = tile % p.tiles_n;
            int row = n_blk * TC_BM + q * 32 + lane;
            if (row >= p.N) row = p.N - 1;   // padded rows compute garbage that the epilogue masks
            if (p.w_layout == WF_W_I2S) {
                const uint4* src = reinterpret_cast<const uint4*>(p.w + (size_t)row * K4 + (size_t)kb * 32);
                const uint4 a = __ldg(src), b = __ldg(src + 1);
                r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
            } else {
                // b200_tile_v1: word (16*(r&7) + 4*(j&3) + 2*(j>>2) + (r>>3)) of tile (row/16, kb) holds packed bytes [4j,4j+4) of row r
                const int r16 = row & 15;
                const uint32_t* src = reinterpret_cast<const uint32_t*>(p.w) + ((size_t)(row >> 4) * p.KB + kb) * 128 + 16 * (r16 & 7) + (r16 >> 3);
#pragma unroll
                for (int j = 0; j < 8; ++j) r[j] = __ldg(src + 4 * (j & 3) + 2 * (j >> 2));
            }
        };
#pragma unroll
        for (int u = 0; u < TC_PF; ++u) load_item(u, pre[u]);
        for (long long item0 = 0; item0 < n_items; item0 += TC_PF) {
#pragma unroll
            for (int u = 0; u < TC_PF; ++u) {
                const long long item = item0 + u;
                if (item < n_items) {
                    uint32_t v[32];
#pragma unroll
                    for (int g = 0; g < 4; ++g)
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[8 * g + j] = (pre[u][j] >> (6 - 2 * g)) & 0x03030303u;   // k = 32g + 4j .. +3
                    load_item(item + TC_PF, pre[u]);
                    const int s = (int)(item % TC_STAGES);
                    mbar_wait(&empty_bar[s], (uint32_t)(((item / TC_STAGES) & 1) ^ 1));
                    tc_fence_after();
                    const uint32_t taddr = tmem + lane_base + TC_A_COL0 + s * TC_A_COLS;
                    TC_ST32(taddr, v);
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&full_bar[s]);
                }
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue warps: TMEM -> registers -> global
        const int q = warp;
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        const bool relu2 = (p.ep.out_mode == WF_OUT_LINEAR) && (p.ep.epi_flags & WF_EPI_RELU2_MUL);
        for (int tl = 0; tl < my_tiles; ++tl) {
            const int tile = blockIdx.x + tl * gridDim.x;
            const int m_blk = tile / p.tiles_n, n_blk = tile % p.tiles_n;
            const int buf = tl & 1;
            mbar_wait(&tfull_bar[buf], (uint32_t)((tl >> 1) & 1));
            tc_fence_after();
            const int n = n_blk * TC_BM + q * 32 + lane;
#pragma unroll 1
            for (int c = 0; c < TC_BN / 32; ++c) {
                uint32_t v[32];
                TC_LD32(tmem + lane_base + buf * TC_BN + c * 32, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (c == TC_BN / 32 - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[buf]);
                }
                const int m0 = m_blk * TC_BN + c * 32;
                if (relu2) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int su = __shfl_down_sync(0xffffffffu, (int)v[j], 16);
                        if (lane < 16 && n < p.N && m0 + j < p.M) epilogue_store_relu2(p.ep, m0 + j, n, (int)v[j], su);
                    }
                } else if (n < p.N) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (m0 + j < p.M) epilogue_store(p.ep, m0 + j, n, (int)v[j]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TC_TMEM_COLS) : "memory");
    }
}

// ---- host side --------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static bool g_tc_attr = false;

void gemm_tc_init_attrs() {
    if (!g_tc_attr && cudaFuncSetAttribute(ternary_gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) == cudaSuccess) g_tc_attr = true;
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            g_encode = (EncodeTiledFn)fn;
    }
}

int launch_gemm_tc(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s) {
    gemm_tc_init_attrs();
    if (!g_tc_attr || !g_encode) { set_error("gemm_tc: tcgen05 path unavailable (driver entry point / smem attribute)"); return WF_ERR_CUDA; }
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)a->K, (cuuint64_t)a->M};
    const cuuint64_t gstride[1] = {(cuuint64_t)a->ldq};
    const cuuint32_t box[2] = {TC_BK, TC_BN};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = g_encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void*)a->q, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { set_error("gemm_tc: cuTensorMapEncodeTiled failed (%d)", (int)cr); return WF_ERR_CUDA; }
    TcParams p;
    p.w = a->w; p.w_layout = a->w_layout;
    p.M = a->M; p.N = a->N; p.K = a->K; p.KB = a->K / WF_QK;
    p.tiles_n = (a->N + TC_BM - 1) / TC_BM;
    p.tiles_m = (a->M + TC_BN - 1) / TC_BN;
    p.ep = ep;
    const int total = p.tiles_n * p.tiles_m;
    const int grid = total < sm_count() ? total : sm_count();
    ternary_gemm_tc_kernel<<<grid, TC_THREADS, TC_SMEM, s>>>(tmap, p);
    return launch_check(cudaGetLastError(), "ternary_gemm_tc");
}

}  // namespace wf
