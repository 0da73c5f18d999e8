// Bandwidth-bound ternary GEMV for decode (M <= 8 activation rows) — north-star item 3.
//
//   S[m,n] = sum_k code[n,k] * a[m,k]   (exact int32; reference: ggml_vec_dot_i2_i8_s,
//   extern/BitNet/src/ggml-bitnet-mad.cpp:94-186, one call per output row under OpenMP)
//
// B200 design:
//   * weights live in HBM in b200_tile_v1: [row-block][k-block] tiles of 512 B, so ANY contiguous range of
//     tiles is one contiguous byte range.  The T = (N/16)*(K/128) tiles are split evenly over a persistent
//     grid of one CTA per SM; each CTA's slab is streamed by the TMA engine (cp.async.bulk 1-D, 8 KB stages,
//     mbarrier full/empty ring, dedicated producer warp) — no tensor map, no per-thread address math, up to
//     64 KB in flight per SM;
//   * the slab does not depend on the previous kernel, so the producer starts BEFORE griddepcontrol.wait:
//     under programmatic dependent launch the weight stream of linear i+1 overlaps the tail of kernel i;
//   * 8 consumer warps: one 16-byte LDS gives a lane the four u32 source words of its mma A fragments;
//     2-bit -> u8 "unpack" is a single AND per fragment (fields are left at their bit position and the four
//     k-groups accumulate into four separate int32 accumulators that are shifted back at flush time — exact);
//     int8 activations come from a skewed smem copy via ldmatrix; mma.sync.m16n8k32.u8.s8 does 16 rows x 8
//     batch columns x 32 k per issue slot;
//   * split-K falls out of the even tile split: row-blocks cut by a CTA boundary are combined with exact
//     int32 atomics in a self-cleaning workspace; the CTA that completes a row-block runs the fused epilogue
//     (Sigma-a correction, scales, bf16, residual add or relu^2*up).
#include "linear_common.cuh"

namespace wf {

extern int g_gemv_debug;

constexpr int SV_MAX_STAGES = 24;                        // ring depth is a launch parameter (p.stages), default 8 = 64 KB

// unsigned (weight field, left at its bit position) x signed (int8 activation) 4-way dot product
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

struct StreamParams {
    const uint8_t* w;
    const int8_t* q;
    int M, N, K, ldq;
    int KB, RB;
    long long T;          // total tiles
    int grp;              // row-blocks per finalisation group (2 with RELU2_MUL)
    int act_stride;       // bytes per activation row in smem
    int sacc_rbs;         // row-blocks of smem accumulators
    int* ws_acc;
    int* ws_cnt;
    int stages;           // ring depth (<= SV_MAX_STAGES)
    int debug;            // bit0: stream-only (wf_bench_gemv: the ceiling of the TMA ring without the math)
    EpiParams ep;
};

// MR = 0: tensor path, mma.sync.m16n8k32.u8.s8 with the batch (<= 8 rows) in the n dimension.
// MR = 1, 2, 4: CUDA-core path for <= MR activation rows: dp4a.u32.s32 on the SAME tile words.  A lane's four u32 words
//   are rows (l/4, l/4+8) x I2_S byte columns (4t, 16+4t), t = l%4; masking k-group g leaves four consecutive k
//   (32g + 4t [+16] ..+3) of that row in the four bytes, which is exactly one dp4a against the activation word of the
//   same four k.  16 AND + 16*MR dp4a per lane per tile, every chain independent and fully pipelined — mma.sync int8
//   on sm_100 is a long-latency legacy path (measured: ~700 cycles per tile per warp), dp4a is not.  The int8
//   activations are staged permuted so that a lane's eight words of a k-block are 32 contiguous bytes (two LDS.128,
//   broadcast across the 8 lanes that share t).
template <int MR, int CW>
__global__ void __launch_bounds__((CW + 1) * 32, CW <= 8 ? 2 : 1) ternary_gemv_stream_kernel(const StreamParams p) {
    constexpr int NT = (CW + 1) * 32, CT = CW * 32;
    constexpr int SV_STAGE_TILES = 2 * CW, SV_STAGE_BYTES = SV_STAGE_TILES * WF_TILE_BYTES;   // a stage = one tile pair per consumer warp
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;                                        // stages * SV_STAGE_BYTES
    unsigned char* acts = ring + p.stages * SV_STAGE_BYTES;            // M * act_stride
    int* sacc = reinterpret_cast<int*>(acts + ((p.M * p.act_stride + 127) & ~127));   // [rbl][8][16]
    __shared__ __align__(8) uint64_t full_bar[SV_MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[SV_MAX_STAGES];
    __shared__ int s_last;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long t0 = (long long)blockIdx.x * p.T / gridDim.x;
    const long long t1 = (long long)(blockIdx.x + 1) * p.T / gridDim.x;
    const int ntiles = (int)(t1 - t0);
    const int nstages = (ntiles + SV_STAGE_TILES - 1) / SV_STAGE_TILES;
    const int rb_first = (int)(t0 / p.KB);
    const int rb_base = (rb_first / p.grp) * p.grp;   // sacc row-block origin (group aligned)
    const int S = p.stages;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], CW);
        }
        fence_mbar_init();
    }
    pdl_launch_dependents();   // dependents still wait (griddepcontrol.wait) for this grid to finish; this only lets their weight streams start
    for (int i = tid; i < p.sacc_rbs * 128; i += NT) sacc[i] = 0;
    __syncthreads();

    if (warp == CW) {
        // ------------------------------------------------ producer warp (one elected lane); never waits on the prior grid
        if (lane == 0) {
            int s = 0, ph = 1;
            for (int it = 0; it < nstages; ++it) {
                mbar_wait(&empty_bar[s], ph);
                const int tiles = min(SV_STAGE_TILES, ntiles - it * SV_STAGE_TILES);
                const uint32_t bytes = (uint32_t)tiles * WF_TILE_BYTES;
                mbar_expect_tx(&full_bar[s], bytes);
                bulk_g2s(ring + s * SV_STAGE_BYTES, p.w + (size_t)(t0 + (long long)it * SV_STAGE_TILES) * WF_TILE_BYTES, bytes,
                         &full_bar[s]);
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
    } else {
        // ------------------------------------------------ consumer warps
        // activations are produced by the preceding kernel: wait for it, then stage them
        pdl_wait_prior();
        if (MR == 0) {
            const int gran_per_row = p.K / 16;
            for (int g = tid; g < p.M * gran_per_row; g += CT) {
                const int m = g / gran_per_row, c = g % gran_per_row;
                cp_async16(acts + m * p.act_stride + c * 16, p.q + (size_t)m * p.ldq + c * 16);
            }
            cp_async_commit();
            cp_async_wait<0>();
        } else {
            // permuted staging: word (8g + t + 4h) of a k-block goes to word (8t + 2g + h)
            const int wpr = p.K / 4;
            for (int m = 0; m < p.M; ++m) {
                const uint32_t* src = reinterpret_cast<const uint32_t*>(p.q + (size_t)m * p.ldq);
                uint32_t* dst = reinterpret_cast<uint32_t*>(acts + m * p.act_stride);
                for (int w = tid; w < wpr; w += CT) {
                    const int wi = w & 31, g = wi >> 3, h = (wi >> 2) & 1, t = wi & 3;
                    dst[(w & ~31) + 8 * t + 2 * g + h] = src[w];
                }
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(CT) : "memory");

        constexpr int NACC = MR == 0 ? 16 : 8 * MR;
        int acc[NACC];
#pragma unroll
        for (int j = 0; j < NACC; ++j) acc[j] = 0;
        int cur_rb = -1;

        auto flush = [&](int rb) {
            if (rb < 0) return;
            int* dst = sacc + (rb - rb_base) * 128;
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
                if (m < p.M) {
                    if (v[0]) atomicAdd(dst + m * 16 + g, v[0]);
                    if (v[2]) atomicAdd(dst + m * 16 + g + 8, v[2]);
                }
                if (m + 1 < p.M) {
                    if (v[1]) atomicAdd(dst + (m + 1) * 16 + g, v[1]);
                    if (v[3]) atomicAdd(dst + (m + 1) * 16 + g + 8, v[3]);
                }
            } else {
                // acc[(m*2 + r)*4 + g]: row r0 + 8r of activation row m, k-group g
#pragma unroll
                for (int mr = 0; mr < 2 * MR; ++mr) {
                    int v = (acc[mr * 4 + 0] >> 6) + (acc[mr * 4 + 1] >> 4) + (acc[mr * 4 + 2] >> 2) + acc[mr * 4 + 3];
#pragma unroll
                    for (int g = 0; g < 4; ++g) acc[mr * 4 + g] = 0;
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    const int m = mr >> 1, r = (lane >> 2) + 8 * (mr & 1);
                    if ((lane & 3) == 0 && m < p.M && v) atomicAdd(dst + m * 16 + r, v);
                }
            }
        };

        // Tile -> warp mapping: a stage holds 2*CW tiles and warp w owns the adjacent pair (2w, 2w+1) of every stage — both
        // weight loads are issued before any math, the pair almost always shares a row-block (one flush per KB tiles), and
        // (row-block, k-block) are tracked incrementally (no division on the per-tile path).
        constexpr int ST = 2 * CW;
        int kb0 = (int)(t0 - (long long)rb_first * p.KB) + 2 * warp, rb0 = rb_first;
        while (kb0 >= p.KB) { kb0 -= p.KB; ++rb0; }
        const int arow = min(lane & 7, p.M - 1);
        const unsigned char* abase = MR == 0 ? acts + arow * p.act_stride + (lane >> 3) * 16 : acts + (lane & 3) * 32;

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
                uint32_t wm[4][4];   // [word][g]
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint32_t mask = 0x03030303u << (6 - 2 * g);
                    wm[0][g] = Wv.x & mask; wm[1][g] = Wv.y & mask; wm[2][g] = Wv.z & mask; wm[3][g] = Wv.w & mask;
                }
#pragma unroll
                for (int m = 0; m < MR; ++m) {
                    const unsigned char* ap = abase + min(m, p.M - 1) * p.act_stride + kb * 128;
                    const uint4 a0 = *reinterpret_cast<const uint4*>(ap);        // (g0,h0) (g0,h1) (g1,h0) (g1,h1)
                    const uint4 a1 = *reinterpret_cast<const uint4*>(ap + 16);   // (g2,h0) (g2,h1) (g3,h0) (g3,h1)
                    const uint32_t A[4][2] = {{a0.x, a0.y}, {a0.z, a0.w}, {a1.x, a1.y}, {a1.z, a1.w}};
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        int& lo = acc[(m * 2 + 0) * 4 + g];
                        int& hi = acc[(m * 2 + 1) * 4 + g];
                        lo = dp4a_us(wm[0][g], A[g][0], lo);
                        hi = dp4a_us(wm[1][g], A[g][0], hi);
                        lo = dp4a_us(wm[2][g], A[g][1], lo);
                        hi = dp4a_us(wm[3][g], A[g][1], hi);
                    }
                }
            }
        };

        int s = 0, ph = 0;
        const unsigned char* wbase = ring + warp * (2 * WF_TILE_BYTES) + lane * 16;
        for (int it = 0; it < nstages; ++it) {
            mbar_wait(&full_bar[s], ph);
            const int i0 = it * ST + 2 * warp;
            if (i0 < ntiles && !(p.debug & 1)) {
                const unsigned char* wp = wbase + s * (ST * WF_TILE_BYTES);
                const bool two = i0 + 1 < ntiles;
                const uint4 W0 = *reinterpret_cast<const uint4*>(wp);
                uint4 W1 = make_uint4(0, 0, 0, 0);
                if (two) W1 = *reinterpret_cast<const uint4*>(wp + WF_TILE_BYTES);
                int kb1 = kb0 + 1, rb1 = rb0;
                if (kb1 >= p.KB) { kb1 = 0; ++rb1; }
                if (rb0 != cur_rb) { flush(cur_rb); cur_rb = rb0; }
                tile_math(W0, kb0);
                if (two) {
                    if (rb1 != cur_rb) { flush(cur_rb); cur_rb = rb1; }
                    tile_math(W1, kb1);
                }
            }
            kb0 += ST;
            while (kb0 >= p.KB) { kb0 -= p.KB; ++rb0; }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
            if (++s == S) { s = 0; ph ^= 1; }
        }
        flush(cur_rb);
    }
    __syncthreads();

    // ------------------------------------------------ combine + fused epilogue, one finalisation group at a time
    if (ntiles > 0) {
        const int rb_last = (int)((t1 - 1) / p.KB);
        const int grp_tiles = p.grp * p.KB;
        const int rows_per_grp = p.grp * 16;
        const bool relu2 = (p.ep.out_mode == WF_OUT_LINEAR) && (p.ep.epi_flags & WF_EPI_RELU2_MUL);
        for (int grp = rb_first / p.grp; grp <= rb_last / p.grp; ++grp) {
            const long long g0 = (long long)grp * grp_tiles, g1 = g0 + grp_tiles;
            const int contrib = (int)((t1 < g1 ? t1 : g1) - (t0 > g0 ? t0 : g0));
            int* sg = sacc + (grp * p.grp - rb_base) * 128;   // [grp rb][m][16]
            const int nbase = grp * rows_per_grp;
            bool fin = true;
            if (contrib != grp_tiles) {
                for (int i = tid; i < p.grp * 128; i += NT) {
                    const int rbl = i >> 7, m = (i >> 4) & 7, r = i & 15;
                    const int n = nbase + rbl * 16 + r;
                    const int v = sg[i];
                    if (m < p.M && n < p.N && v) atomicAdd(p.ws_acc + (size_t)m * p.N + n, v);
                }
                __threadfence();
                __syncthreads();
                if (tid == 0) {
                    const int prev = atomicAdd(p.ws_cnt + grp, contrib);
                    s_last = (prev + contrib == grp_tiles);
                    if (s_last) p.ws_cnt[grp] = 0;
                }
                __syncthreads();
                fin = s_last != 0;
                if (fin) {
                    __threadfence();
                    for (int i = tid; i < p.grp * 128; i += NT) {
                        const int rbl = i >> 7, m = (i >> 4) & 7, r = i & 15;
                        const int n = nbase + rbl * 16 + r;
                        if (m < p.M && n < p.N) {
                            int* a = p.ws_acc + (size_t)m * p.N + n;
                            sg[i] = ldcg_i32(a);
                            *a = 0;
                        }
                    }
                }
                __syncthreads();
            }
            if (fin) {
                if (relu2) {
                    for (int i = tid; i < p.M * 16; i += NT) {
                        const int m = i >> 4, r = i & 15;
                        epilogue_store_relu2(p.ep, m, nbase + r, sg[m * 16 + r], sg[128 + m * 16 + r]);
                    }
                } else {
                    for (int i = tid; i < p.M * rows_per_grp; i += NT) {
                        const int m = i / rows_per_grp, rr = i % rows_per_grp;
                        const int n = nbase + rr;
                        if (n < p.N) epilogue_store(p.ep, m, n, sg[(rr >> 4) * 128 + m * 16 + (rr & 15)]);
                    }
                }
            }
        }
    }
}

int g_gemv_stages = 8;    // ring depth override (wf_bench_gemv)
int g_gemv_path = -1;     // -1 auto, 0 force the mma path
int g_gemv_cw = 0;        // 0 auto, 8 / 16 force the consumer-warp count

typedef void (*stream_kernel_t)(const StreamParams);
template <int CW>
static stream_kernel_t pick_mr(int M) {
    if (g_gemv_path == 0 || M > 4) return ternary_gemv_stream_kernel<0, CW>;
    if (M == 1) return ternary_gemv_stream_kernel<1, CW>;
    if (M == 2) return ternary_gemv_stream_kernel<2, CW>;
    return ternary_gemv_stream_kernel<4, CW>;
}
// small slabs (a few KB per SM) are latency-bound: 8 consumer warps, 8 KB stages, two CTAs per SM can co-reside so that
// the next launch's weight stream overlaps under PDL; large slabs are issue-bound: 16 consumer warps, 16 KB stages.
static stream_kernel_t pick_kernel(int M, long long tiles_per_cta, int* cw) {
    int c = g_gemv_cw ? g_gemv_cw : (tiles_per_cta >= 256 ? 16 : 8);
    *cw = c;
    return c == 16 ? pick_mr<16>(M) : pick_mr<8>(M);
}

static size_t g_stream_attr_smem = 0;
void gemv_stream_init_attrs() {
    bool ok = true;
    const stream_kernel_t all[8] = {ternary_gemv_stream_kernel<0, 8>, ternary_gemv_stream_kernel<1, 8>, ternary_gemv_stream_kernel<2, 8>, ternary_gemv_stream_kernel<4, 8>,
                                    ternary_gemv_stream_kernel<0, 16>, ternary_gemv_stream_kernel<1, 16>, ternary_gemv_stream_kernel<2, 16>, ternary_gemv_stream_kernel<4, 16>};
    for (stream_kernel_t k : all) ok &= cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) == cudaSuccess;
    if (ok) g_stream_attr_smem = 220 * 1024;
}

int launch_gemv_stream(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s, bool pdl) {
    if (a->w_layout != WF_W_TILE) { set_error("gemv_stream: needs the b200_tile_v1 layout"); return WF_ERR_UNSUPPORTED; }
    if (a->M < 1 || a->M > 8) { set_error("gemv_stream: M=%d outside 1..8", a->M); return WF_ERR_UNSUPPORTED; }
    StreamParams p;
    p.w = a->w; p.q = a->q;
    p.M = a->M; p.N = a->N; p.K = a->K; p.ldq = a->ldq;
    p.KB = a->K / WF_QK;
    p.RB = (a->N + 15) / 16;
    p.T = (long long)p.KB * p.RB;
    p.grp = (a->epi_flags & WF_EPI_RELU2_MUL) ? 2 : 1;
    p.act_stride = a->K + 16;
    p.ep = ep;
    p.debug = g_gemv_debug;
    int grid = sm_count();
    const long long min_tiles = 4;
    if (p.T / min_tiles < grid) grid = (int)((p.T + min_tiles - 1) / min_tiles);
    if (grid < 1) grid = 1;
    const long long max_tiles = (p.T + grid - 1) / grid + 1;
    p.sacc_rbs = (int)(max_tiles / p.KB) + 2 + 2 * p.grp;
    const size_t need_ws = (size_t)a->M * a->N * 4 + (size_t)p.RB * 4;
    if (!a->workspace || a->workspace_bytes < need_ws) { set_error("gemv_stream: workspace too small (%zu < %zu)", a->workspace_bytes, need_ws); return WF_ERR_INVALID; }
    p.ws_acc = (int*)a->workspace;
    p.ws_cnt = (int*)((char*)a->workspace + (size_t)a->M * a->N * 4);
    p.stages = g_gemv_stages < 2 ? 2 : (g_gemv_stages > SV_MAX_STAGES ? SV_MAX_STAGES : g_gemv_stages);
    int cw = 8;
    stream_kernel_t kern = pick_kernel(a->M, max_tiles, &cw);
    const int threads = (cw + 1) * 32;
    const size_t stage_bytes = (size_t)2 * cw * WF_TILE_BYTES;
    auto smem_for = [&](int st) { return (size_t)st * stage_bytes + (((size_t)p.M * p.act_stride + 127) & ~(size_t)127) + (size_t)p.sacc_rbs * 512; };
    while (p.stages > 4 && smem_for(p.stages) > 220 * 1024) --p.stages;
    const size_t smem = smem_for(p.stages);
    if (smem > 220 * 1024) { set_error("gemv_stream: M*K=%d*%d does not fit in shared memory", a->M, a->K); return WF_ERR_UNSUPPORTED; }
    if (smem > g_stream_attr_smem) {
        gemv_stream_init_attrs();
        if (smem > g_stream_attr_smem) { set_error("gemv_stream: cannot opt in to %zu bytes of shared memory", smem); return WF_ERR_CUDA; }
    }

    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, kern, p), "ternary_gemv_stream");
}

}  // namespace wf
