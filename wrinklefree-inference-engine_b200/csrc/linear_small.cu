// Small-batch (M <= 8) decode linear: ONE launch = glue prologue + bandwidth-bound ternary GEMV.
//
//   prologue : x -> (residual | relu^2*up) -> RMSNorm -> per-token int8   (fused_common.cuh, every CTA redundantly:
//              M*K <= 8*28672 elements, cheaper than a launch + a trip through HBM)
//   main     : S[m,n] += sum_k code[n,k] * a[m,k] over this CTA's contiguous slab of b200_tile_v1 tiles
//              (reference row dot: extern/BitNet/src/ggml-bitnet-mad.cpp:94-186)
//   epilogue : exact fp32 RED of the partial sums into the global accumulator; the NEXT phase dequantises.
//
// B200 mapping:
//   * persistent grid, one CTA per SM, the T tiles of the matrix split evenly (split-K falls out of it);
//   * a dedicated producer warp streams the slab with the TMA engine (cp.async.bulk 1-D, 8 KB stages,
//     mbarrier full/empty ring).  The slab does not depend on the previous kernel, so the producer never
//     executes griddepcontrol.wait, and every CTA signals griddepcontrol.launch_dependents FIRST: under
//     programmatic dependent launch the NEXT linear's CTAs become co-resident and fill their rings while this
//     one is still computing — the weight stream runs ahead of the layer's dependency chain;
//   * 8 consumer warps: 16-byte LDS of a tile = the four A-fragment source words; 2-bit -> u8 is one AND per
//     fragment (fields stay at their bit position, four int32 accumulator sets are shifted back at flush);
//     int8 activations via ldmatrix from a skewed smem copy; mma.sync.m16n8k32.u8.s8.
#include "fused_common.cuh"

namespace wf {

constexpr int LS_CW = 12;                                 // consumer warps (3 per scheduler; 2 CTAs/SM must still fit)
constexpr int LS_TPW = 2;                                 // adjacent tiles per consumer warp per stage (both weight loads issued before the math)
constexpr int LS_STAGE_TILES = LS_TPW * LS_CW;            // 12 KB per stage
constexpr int LS_STAGE_BYTES = LS_STAGE_TILES * WF_TILE_BYTES;
constexpr int LS_STAGES = 5;                              // 60 KB ring
constexpr int LS_CT = LS_CW * 32;                         // consumer threads
constexpr int LS_THREADS = LS_CT + 32;

struct SmallLinParams {
    GlueParams g;
    const uint8_t* w;
    int N, K, KB;         // K / KB: columns of this rank's weight slice
    int k_off;            // first activation column of the slice
    long long T;
    int act_stride;       // bytes per activation row in smem (g.K + 16: 16-B skew => conflict-free ldmatrix)
    int sacc_rbs;         // row-blocks of smem accumulators
    TpDest dest;
    int acc_ld;
    float* lin_ascale;
    int32_t* lin_rowsum;
};

__device__ __forceinline__ int ls_dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// MR = 1: one activation row, CUDA-core path (dp4a on the tile words, int8 codes staged in the permuted order — see
// gemv_stream.cu); MR = 0: up to 8 rows on mma.sync.m16n8k32 (the batch is the n dimension).
template <int MR>
__global__ void __launch_bounds__(LS_THREADS, 2) linear_small_kernel(const SmallLinParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;                                         // LS_STAGES * LS_STAGE_BYTES
    unsigned char* acts = ring + LS_STAGES * LS_STAGE_BYTES;            // M * act_stride
    __nv_bfloat16* ys = reinterpret_cast<__nv_bfloat16*>(acts + ((p.g.M * p.act_stride + 127) & ~127));   // g.K bf16
    int* sacc = reinterpret_cast<int*>(reinterpret_cast<unsigned char*>(ys) + ((p.g.K * 2 + 127) & ~127));   // [rbl][8][16]
    __shared__ __align__(8) uint64_t full_bar[LS_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[LS_STAGES];
    __shared__ float fs[32];
    __shared__ int is[32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int M = p.g.M;
    const long long t0 = (long long)blockIdx.x * p.T / gridDim.x;
    const long long t1 = (long long)(blockIdx.x + 1) * p.T / gridDim.x;
    const int ntiles = (int)(t1 - t0);
    const int nstages = (ntiles + LS_STAGE_TILES - 1) / LS_STAGE_TILES;
    const int rb_first = (int)(t0 / p.KB);

    pdl_launch_dependents();   // let the next phase's CTAs in: their weight stream overlaps this phase
    const bool tr = p.g.trace != nullptr && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && tid == 0;
    unsigned long long* trp = tr ? p.g.trace + (blockIdx.x == 0 ? 0 : 16) : nullptr;
    if (tr) trp[0] = gtimer();
    if (tid == 0) {
        for (int s = 0; s < LS_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], LS_CW);
        }
        fence_mbar_init();
    }
    for (int i = tid; i < p.sacc_rbs * 128; i += LS_THREADS) sacc[i] = 0;
    __syncthreads();

    if (warp == LS_CW) {
        // ------------------------------------------------ producer warp (one elected lane); never waits on the prior grid
        if (lane == 0) {
            for (int it = 0; it < nstages; ++it) {
                const int s = it % LS_STAGES;
                mbar_wait(&empty_bar[s], ((it / LS_STAGES) & 1) ^ 1);
                const int tiles = min(LS_STAGE_TILES, ntiles - it * LS_STAGE_TILES);
                const uint32_t bytes = (uint32_t)tiles * WF_TILE_BYTES;
                mbar_expect_tx(&full_bar[s], bytes);
                bulk_g2s(ring + s * LS_STAGE_BYTES, p.w + (size_t)(t0 + (long long)it * LS_STAGE_TILES) * WF_TILE_BYTES, bytes,
                         &full_bar[s]);
            }
        }
    } else {
        // ------------------------------------------------ consumer warps
        glue_prefetch_gamma<LS_CT>(p.g);   // weights: no dependency on the previous phase
        pdl_wait_prior();          // everything below reads what the previous phase wrote
        if (tr) trp[1] = gtimer();
        if (p.g.wait_flag >= 0) {  // ... and, under tensor parallelism, what the peers wrote into this rank's buffers
            if (warp == 0) tp_wait_flags_warp(p.g.tp, p.g.wait_flag);
            group_bar<1, LS_CT>();
        }
        if (tr) trp[2] = gtimer();
        glue_zero_duty<LS_CT>(p.g);
        for (int m = 0; m < M; ++m) {
            float sc; int rsum;
            glue_row<1, LS_CT>(p.g, m, ys, reinterpret_cast<int8_t*>(acts + m * p.act_stride), fs, is, blockIdx.x == 0, sc, rsum, (tr && m == 0) ? trp + 8 : nullptr,
                               MR == 1);
            if (blockIdx.x == 0 && tid == 0) { p.lin_ascale[m] = sc; p.lin_rowsum[m] = rsum; }
        }
        group_bar<1, LS_CT>();     // activations complete
        if (tr) trp[3] = gtimer();

        constexpr int NACC = MR == 0 ? 16 : 8;
        int acc[NACC];
#pragma unroll
        for (int j = 0; j < NACC; ++j) acc[j] = 0;
        int cur_rb = -1;

        auto flush = [&](int rb) {
            if (rb < 0) return;
            int* dst = sacc + (rb - rb_first) * 128;
            if (MR == 0) {
                const int g = lane >> 2, t = lane & 3;
                int v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    v[j] = (acc[0 + j] >> 6) + (acc[4 + j] >> 4) + (acc[8 + j] >> 2) + acc[12 + j];
#pragma unroll
                    for (int gg = 0; gg < 4; ++gg) acc[gg * 4 + j] = 0;
                }
                const int m = 2 * t;
                if (m < M) {
                    if (v[0]) atomicAdd(dst + m * 16 + g, v[0]);
                    if (v[2]) atomicAdd(dst + m * 16 + g + 8, v[2]);
                }
                if (m + 1 < M) {
                    if (v[1]) atomicAdd(dst + (m + 1) * 16 + g, v[1]);
                    if (v[3]) atomicAdd(dst + (m + 1) * 16 + g + 8, v[3]);
                }
            } else {
                // acc[r*4 + g]: row (lane/4) + 8r, k-group g; the four lanes of a row hold disjoint k
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    int v = (acc[r * 4 + 0] >> 6) + (acc[r * 4 + 1] >> 4) + (acc[r * 4 + 2] >> 2) + acc[r * 4 + 3];
#pragma unroll
                    for (int g = 0; g < 4; ++g) acc[r * 4 + g] = 0;
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    if ((lane & 3) == 0 && v) atomicAdd(dst + (lane >> 2) + 8 * r, v);
                }
            }
        };

        const int KB = p.KB;
        const int arow = min(lane & 7, M - 1);
        const unsigned char* abase = MR == 0 ? acts + arow * p.act_stride + p.k_off + (lane >> 3) * 16 : acts + p.k_off + (lane & 3) * 32;
        auto tile_math = [&](const uint4& Wv, int kb) {
            if (MR == 0) {
                uint32_t b[8];
                const unsigned char* ap = abase + kb * 128;
                ldmatrix_x4(b[0], b[1], b[2], b[3], ap);
                ldmatrix_x4(b[4], b[5], b[6], b[7], ap + 64);
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint32_t mask = 0x03030303u << (6 - 2 * g);   // k-group g sits in bits (7-2g .. 6-2g)
                    int(&d)[4] = *reinterpret_cast<int(*)[4]>(&acc[g * 4]);
                    mma_u8s8_16832(d, Wv.x & mask, Wv.y & mask, Wv.z & mask, Wv.w & mask, b[2 * g], b[2 * g + 1]);
                }
            } else {
                const unsigned char* ap = abase + kb * 128;
                const uint4 a0 = *reinterpret_cast<const uint4*>(ap);        // (g0,h0) (g0,h1) (g1,h0) (g1,h1)
                const uint4 a1 = *reinterpret_cast<const uint4*>(ap + 16);   // (g2,h0) (g2,h1) (g3,h0) (g3,h1)
                const uint32_t A[4][2] = {{a0.x, a0.y}, {a0.z, a0.w}, {a1.x, a1.y}, {a1.z, a1.w}};
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint32_t mask = 0x03030303u << (6 - 2 * g);
                    acc[g] = ls_dp4a_us(Wv.x & mask, A[g][0], acc[g]);
                    acc[4 + g] = ls_dp4a_us(Wv.y & mask, A[g][0], acc[4 + g]);
                    acc[g] = ls_dp4a_us(Wv.z & mask, A[g][1], acc[g]);
                    acc[4 + g] = ls_dp4a_us(Wv.w & mask, A[g][1], acc[4 + g]);
                }
            }
        };

        // Tile -> warp mapping: a stage holds LS_TPW*LS_CW tiles, warp w owns the LS_TPW adjacent tiles starting at LS_TPW*w;
        // (row-block, k-block) are tracked incrementally (no division on the per-tile path).
        int kb0 = (int)(t0 - (long long)rb_first * KB) + LS_TPW * warp, rb0 = rb_first;
        while (kb0 >= KB) { kb0 -= KB; ++rb0; }
        const unsigned char* wbase = ring + warp * (LS_TPW * WF_TILE_BYTES) + lane * 16;
        int s = 0, ph = 0;
        for (int it = 0; it < nstages; ++it) {
            mbar_wait(&full_bar[s], ph);
            const int i0 = it * LS_STAGE_TILES + LS_TPW * warp;
            if (i0 < ntiles) {
                const unsigned char* wp = wbase + s * LS_STAGE_BYTES;
                const bool two = i0 + 1 < ntiles;
                const uint4 W0 = *reinterpret_cast<const uint4*>(wp);
                uint4 W1 = make_uint4(0, 0, 0, 0);
                if (two) W1 = *reinterpret_cast<const uint4*>(wp + WF_TILE_BYTES);
                int kb1 = kb0 + 1, rb1 = rb0;
                if (kb1 >= KB) { kb1 = 0; ++rb1; }
                if (rb0 != cur_rb) { flush(cur_rb); cur_rb = rb0; }
                tile_math(W0, kb0);
                if (two) {
                    if (rb1 != cur_rb) { flush(cur_rb); cur_rb = rb1; }
                    tile_math(W1, kb1);
                }
            }
            kb0 += LS_STAGE_TILES;
            while (kb0 >= KB) { kb0 -= KB; ++rb0; }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
            if (++s == LS_STAGES) { s = 0; ph ^= 1; }
        }
        flush(cur_rb);
        group_bar<1, LS_CT>();
        if (tr) trp[4] = gtimer();
        // ------------------------------------------------ exact fp32 RED of this CTA's partial sums
        if (ntiles > 0) {
            const int rb_last = (int)((t1 - 1) / p.KB);
            const int n_rbs = rb_last - rb_first + 1;
            if ((p.N & 3) == 0 && (p.acc_ld & 3) == 0) {
                // four consecutive rows of one token per RED.v4: a quarter of the NVLink packets under tensor parallelism
                for (int i = tid; i < n_rbs * 32; i += LS_CT) {
                    const int rbl = i >> 5, m = (i >> 2) & 7, r = (i & 3) * 4;
                    const int n = (rb_first + rbl) * 16 + r;
                    const int4 v = *reinterpret_cast<const int4*>(sacc + rbl * 128 + m * 16 + r);
                    if (m < M && n < p.N && (v.x | v.y | v.z | v.w)) {
                        // row-parallel linears under TP: the RED into every rank's accumulator IS the all-reduce
                        for (int d = 0; d < p.dest.n; ++d)
                            red_add_v4_f32(p.dest.acc[d] + (size_t)m * p.acc_ld + n, (float)v.x, (float)v.y, (float)v.z, (float)v.w);
                    }
                }
            } else {
                for (int i = tid; i < n_rbs * 128; i += LS_CT) {
                    const int rbl = i >> 7, m = (i >> 4) & 7, r = i & 15;
                    const int n = (rb_first + rbl) * 16 + r;
                    const int v = sacc[i];
                    if (m < M && n < p.N && v)
                        for (int d = 0; d < p.dest.n; ++d) red_add_f32(p.dest.acc[d] + (size_t)m * p.acc_ld + n, (float)v);
                }
            }
        }
    }
    if (tr) trp[5] = gtimer();
    if (p.dest.signal_flag >= 0) tp_signal_grid_done(p.g.tp, p.dest.signal_flag, gridDim.x);
    if (tr) trp[6] = gtimer();
}

// ---------------------------------------------------------------------------------------------------------------
// Standalone glue for large batch: grid = M rows, 256 threads; int8 codes / scale / rowsum go to global.
constexpr int GR_T = 512;   // one pass iteration for K <= 4096, two for the 2B ffn width
__global__ void __launch_bounds__(GR_T) glue_rows_kernel(const GlueParams p) {
    extern __shared__ __align__(16) unsigned char gsm[];
    __nv_bfloat16* ys = reinterpret_cast<__nv_bfloat16*>(gsm);
    __shared__ float fs[32];
    __shared__ int is[32];
    pdl_launch_dependents();
    pdl_wait_prior();
    if (p.wait_flag >= 0) {
        if (threadIdx.x < 32) tp_wait_flags_warp(p.tp, p.wait_flag);
        __syncthreads();
    }
    glue_zero_duty<GR_T>(p);
    const int m = blockIdx.x;
    float sc; int rsum;
    glue_row<1, GR_T>(p, m, ys, p.q_out + (size_t)m * p.K, fs, is, true, sc, rsum);
    if (threadIdx.x == 0) {
        if (p.scale_out) p.scale_out[m] = sc;
        if (p.rowsum_out) p.rowsum_out[m] = rsum;
    }
}

static void set_pdl(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* at, bool pdl) {
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
}

bool linear_small_fits(int M, int K) {   // K = length of the full activation vector
    if (M < 1 || M > 8) return false;
    const size_t smem = (size_t)LS_STAGES * LS_STAGE_BYTES + (((size_t)M * (K + 16) + 127) & ~(size_t)127) + (((size_t)K * 2 + 127) & ~(size_t)127) + 16 * 1024;
    return smem <= 220 * 1024;
}

int launch_linear_small(const SmallLinArgs& a, cudaStream_t s, bool pdl) {
    if (a.g.M < 1 || a.g.M > 8) { set_error("linear_small: M=%d outside 1..8", a.g.M); return WF_ERR_UNSUPPORTED; }
    if (a.K % WF_QK || a.g.K % WF_QK || a.k_off % WF_QK || a.k_off + a.K > a.g.K) { set_error("linear_small: bad K / k_off"); return WF_ERR_INVALID; }
    SmallLinParams p;
    p.g = a.g;
    p.w = a.w; p.N = a.N; p.K = a.K; p.KB = a.K / WF_QK; p.k_off = a.k_off;
    const int RB = (a.N + 15) / 16;
    p.T = (long long)p.KB * RB;
    p.act_stride = a.g.K + 16;
    p.dest = a.dest; p.acc_ld = a.acc_ld;
    p.lin_ascale = a.lin_ascale; p.lin_rowsum = a.lin_rowsum;
    int grid = sm_count();
    const long long min_tiles = 4;
    if (p.T / min_tiles < grid) grid = (int)((p.T + min_tiles - 1) / min_tiles);
    if (grid < 1) grid = 1;
    const long long max_tiles = (p.T + grid - 1) / grid + 1;
    p.sacc_rbs = (int)(max_tiles / p.KB) + 3;
    const size_t smem = (size_t)LS_STAGES * LS_STAGE_BYTES + (((size_t)a.g.M * p.act_stride + 127) & ~(size_t)127) +
                        (((size_t)a.g.K * 2 + 127) & ~(size_t)127) + (size_t)p.sacc_rbs * 512;
    if (smem > 220 * 1024) { set_error("linear_small: M*K=%d*%d does not fit in shared memory", a.g.M, a.g.K); return WF_ERR_UNSUPPORTED; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(LS_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    set_pdl(cfg, at, pdl);
    if (a.g.M == 1) return launch_check(cudaLaunchKernelEx(&cfg, linear_small_kernel<1>, p), "linear_small");
    return launch_check(cudaLaunchKernelEx(&cfg, linear_small_kernel<0>, p), "linear_small");
}

// Long rows at small batch: a thread-block cluster of C CTAs per row, each a contiguous slice; the three row reductions are finished over DSMEM
// (ClusterXR).  grid = (C, M), cluster = (C, 1, 1): blockIdx.x is the rank.
__global__ void __launch_bounds__(GR_T) glue_rows_cluster_kernel(const GlueParams p) {
    extern __shared__ __align__(16) unsigned char gsm[];
    __nv_bfloat16* ys = reinterpret_cast<__nv_bfloat16*>(gsm);
    __shared__ float fs[32];
    __shared__ int is[32];
    __shared__ float xs[4];
    pdl_launch_dependents();
    pdl_wait_prior();
    if (p.wait_flag >= 0) {
        if (threadIdx.x < 32) tp_wait_flags_warp(p.tp, p.wait_flag);
        __syncthreads();
    }
    {   // accumulator clearing duty, strided over the whole 2-D grid
        const int gtid = (blockIdx.y * gridDim.x + blockIdx.x) * GR_T + threadIdx.x, gstride = gridDim.x * gridDim.y * GR_T;
        if (p.zero0)
            for (int i = gtid * 4; i < p.zero0_n; i += gstride * 4) *reinterpret_cast<float4*>(p.zero0 + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p.zero1)
            for (int i = gtid * 4; i < p.zero1_n; i += gstride * 4) *reinterpret_cast<float4*>(p.zero1 + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const int C = gridDim.x, rank = blockIdx.x, m = blockIdx.y;
    const int gran = p.K / 8;
    const int kb = (int)((long long)gran * rank / C) * 8, ke = (int)((long long)gran * (rank + 1) / C) * 8;
    ClusterXR xr;
    xr.slot = xs; xr.n = C;
    float sc; int rsum;
    glue_row<1, GR_T, ClusterXR>(p, m, ys, p.q_out + (size_t)m * p.K, fs, is, true, sc, rsum, nullptr, QMAP_PLAIN, kb, ke, xr);
    if (threadIdx.x == 0 && rank == 0) {
        if (p.scale_out) p.scale_out[m] = sc;
        if (p.rowsum_out) p.rowsum_out[m] = rsum;
    }
    xr.sync();   // nobody leaves while a peer may still read its partials
}

int launch_glue_rows(const GlueParams& g, cudaStream_t s, bool pdl) {
    if (g.M < 1 || g.K % 8) { set_error("glue_rows: bad shape"); return WF_ERR_INVALID; }
    // clusters for long rows when the rows alone leave most SMs idle: OPT-IN (WF_GLUE_CLUSTER=1), measured (profiles/r02ax_summary.txt): 70B bs16/8K
    // 19.84 -> 19.73 ms, bs8/4K 10.74 -> 10.56; forced on shorter rows it loses — 2B bs64/4K 5.10 -> 5.39 ms, 8B bs16/2K 3.05 -> 3.27: a cluster
    // launch and its three cluster barriers cost more than the second half of a 4096..6912-element row
    static const int cl_env = getenv("WF_GLUE_CLUSTER") ? atoi(getenv("WF_GLUE_CLUSTER")) : 0;
    // default: only rows of >= 8192 elements (the 70B widths).  The slices cut the fp32 sum of squares differently from the one-CTA kernels, so the
    // bit-identity with the one-kernel-per-op path that the tests hold at the 2B / 8B widths (where 1/rms rounds to the same bf16 in all but
    // ~5e-4 of the rows) is not claimed there; WF_GLUE_CLUSTER=2..8 forces that cluster size for any row of >= 2048 elements (measurement).
    int C = g.K >= 16384 ? 8 : (g.K >= 8192 ? 4 : 1);
    if (cl_env > 1 && g.K >= 2048) C = cl_env > 8 ? 8 : cl_env;
    while (C > 1 && g.M * C > 2 * sm_count()) C >>= 1;
    if (cl_env && C > 1 && g.q_out) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(C, g.M);
        cfg.blockDim = dim3(GR_T);
        cfg.dynamicSmemBytes = ((size_t)(g.K / 8 / C) + 1) * 16;
        cfg.stream = s;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = pdl ? 2 : 1;
        return launch_check(cudaLaunchKernelEx(&cfg, glue_rows_cluster_kernel, g), "glue_rows_cluster");
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(g.M);
    cfg.blockDim = dim3(GR_T);
    cfg.dynamicSmemBytes = (size_t)g.K * 2;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    set_pdl(cfg, at, pdl);
    return launch_check(cudaLaunchKernelEx(&cfg, glue_rows_kernel, g), "glue_rows");
}

// ---------------------------------------------------------------------------------------------------------------
// Tensor parallel phase D': h = relu^2(gate) * up of this rank's ffn slice (direct_inference_full.py:259, with the
// reference's bf16 rounding points), stored into h_full of EVERY rank — the all-gather is the store itself.
__global__ void __launch_bounds__(256) h_exchange_kernel(const HExchangeArgs a) {
    pdl_launch_dependents();
    pdl_wait_prior();
    const int m = blockIdx.y;
    const int k = (blockIdx.x * 256 + threadIdx.x) * 8;
    if (k < a.ffn_local) {
        const int rs = a.rowsum[m];
        const float sa = a.ascale[m], wg = a.wscale[0], wu = a.wscale[1];
        const float* arow = a.acc + (size_t)m * 2 * a.ffn_local;
        const int ng = (k >> 4) * 32 + (k & 15);
        const float4 g0 = ldcg_f4(arow + ng), g1 = ldcg_f4(arow + ng + 4);
        const float4 u0 = ldcg_f4(arow + ng + 16), u1 = ldcg_f4(arow + ng + 20);
        const float G[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float U[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float g = bf16_round(lin_value(G[i], rs, wg, sa));
            const float u = bf16_round(lin_value(U[i], rs, wu, sa));
            const float r = fmaxf(g, 0.f);
            x[i] = bf16_round(bf16_round(r * r) * u);
        }
        const uint4 v = make_uint4(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]), pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
        for (int d = 0; d < a.n_dst; ++d) *reinterpret_cast<uint4*>(a.dst[d] + (size_t)m * a.ffn_full + a.col0 + k) = v;
    }
    if (a.signal_flag >= 0) tp_signal_grid_done(a.tp, a.signal_flag, gridDim.x * gridDim.y);
}

int launch_h_exchange(const HExchangeArgs& a, cudaStream_t s, bool pdl) {
    if (a.M < 1 || a.ffn_local % 16 || a.ffn_full % 8 || a.col0 % 8) { set_error("h_exchange: bad shape"); return WF_ERR_INVALID; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((a.ffn_local / 8 + 255) / 256, a.M);
    cfg.blockDim = dim3(256);
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    set_pdl(cfg, at, pdl);
    return launch_check(cudaLaunchKernelEx(&cfg, h_exchange_kernel, a), "h_exchange");
}

void linear_small_init_attrs() {
    cudaFuncSetAttribute(linear_small_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(linear_small_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(glue_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
}

}  // namespace wf
