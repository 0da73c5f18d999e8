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

constexpr int SV_STAGE_TILES = 16;                       // 8 KB per stage
constexpr int SV_STAGE_BYTES = SV_STAGE_TILES * WF_TILE_BYTES;
constexpr int SV_STAGES = 8;                             // 64 KB ring
constexpr int SV_CONSUMER_WARPS = 8;
constexpr int SV_THREADS = (SV_CONSUMER_WARPS + 1) * 32;

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
    EpiParams ep;
};

__global__ void __launch_bounds__(SV_THREADS, 1) ternary_gemv_stream_kernel(const StreamParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;                                        // SV_STAGES * SV_STAGE_BYTES
    unsigned char* acts = ring + SV_STAGES * SV_STAGE_BYTES;           // M * act_stride
    int* sacc = reinterpret_cast<int*>(acts + ((p.M * p.act_stride + 127) & ~127));   // [rbl][8][16]
    __shared__ __align__(8) uint64_t full_bar[SV_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[SV_STAGES];
    __shared__ int s_last;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long t0 = (long long)blockIdx.x * p.T / gridDim.x;
    const long long t1 = (long long)(blockIdx.x + 1) * p.T / gridDim.x;
    const int ntiles = (int)(t1 - t0);
    const int nstages = (ntiles + SV_STAGE_TILES - 1) / SV_STAGE_TILES;
    const int rb_first = (int)(t0 / p.KB);
    const int rb_base = (rb_first / p.grp) * p.grp;   // sacc row-block origin (group aligned)

    if (tid == 0) {
        for (int s = 0; s < SV_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], SV_CONSUMER_WARPS);
        }
        fence_mbar_init();
    }
    for (int i = tid; i < p.sacc_rbs * 128; i += SV_THREADS) sacc[i] = 0;
    __syncthreads();

    if (warp == SV_CONSUMER_WARPS) {
        // ------------------------------------------------ producer warp (one elected lane)
        if (lane == 0) {
            for (int it = 0; it < nstages; ++it) {
                const int s = it % SV_STAGES;
                mbar_wait(&empty_bar[s], ((it / SV_STAGES) & 1) ^ 1);
                const int tiles = min(SV_STAGE_TILES, ntiles - it * SV_STAGE_TILES);
                const uint32_t bytes = (uint32_t)tiles * WF_TILE_BYTES;
                mbar_expect_tx(&full_bar[s], bytes);
                bulk_g2s(ring + s * SV_STAGE_BYTES, p.w + (size_t)(t0 + (long long)it * SV_STAGE_TILES) * WF_TILE_BYTES, bytes,
                         &full_bar[s]);
            }
        }
        pdl_wait_prior();
    } else {
        // ------------------------------------------------ consumer warps
        // activations are produced by the preceding kernel: wait for it, then stage them
        pdl_wait_prior();
        {
            const int gran_per_row = p.K / 16;
            const int ctid = tid;  // 0..255
            for (int g = ctid; g < p.M * gran_per_row; g += SV_CONSUMER_WARPS * 32) {
                const int m = g / gran_per_row, c = g % gran_per_row;
                cp_async16(acts + m * p.act_stride + c * 16, p.q + (size_t)m * p.ldq + c * 16);
            }
            cp_async_commit();
            cp_async_wait<0>();
            asm volatile("bar.sync 1, %0;" ::"n"(SV_CONSUMER_WARPS * 32) : "memory");
        }
        int acc[4][4];
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[g][j] = 0;
        int cur_rb = -1;
        const int arow = min(lane & 7, p.M - 1);
        const unsigned char* abase = acts + arow * p.act_stride + (lane >> 3) * 16;

        auto flush = [&](int rb) {
            if (rb < 0) return;
            const int g = lane >> 2, t = lane & 3;
            int v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[j] = (acc[0][j] >> 6) + (acc[1][j] >> 4) + (acc[2][j] >> 2) + acc[3][j];
#pragma unroll
                for (int gg = 0; gg < 4; ++gg) acc[gg][j] = 0;
            }
            int* dst = sacc + (rb - rb_base) * 128;
            const int m = 2 * t;
            if (m < p.M) {
                if (v[0]) atomicAdd(dst + m * 16 + g, v[0]);
                if (v[2]) atomicAdd(dst + m * 16 + g + 8, v[2]);
            }
            if (m + 1 < p.M) {
                if (v[1]) atomicAdd(dst + (m + 1) * 16 + g, v[1]);
                if (v[3]) atomicAdd(dst + (m + 1) * 16 + g + 8, v[3]);
            }
        };

        for (int it = 0; it < nstages; ++it) {
            const int s = it % SV_STAGES;
            mbar_wait(&full_bar[s], (it / SV_STAGES) & 1);
            const int tiles = min(SV_STAGE_TILES, ntiles - it * SV_STAGE_TILES);
            for (int tl = warp; tl < tiles; tl += SV_CONSUMER_WARPS) {
                const long long tg = t0 + (long long)it * SV_STAGE_TILES + tl;
                const int rb = (int)(tg / p.KB), kb = (int)(tg % p.KB);
                if (rb != cur_rb) { flush(cur_rb); cur_rb = rb; }
                const uint4 Wv = *reinterpret_cast<const uint4*>(ring + s * SV_STAGE_BYTES + tl * WF_TILE_BYTES + lane * 16);
                uint32_t b[8];
                const unsigned char* ap = abase + kb * 128;
                ldmatrix_x4(b[0], b[1], b[2], b[3], ap);
                ldmatrix_x4(b[4], b[5], b[6], b[7], ap + 64);
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint32_t mask = 0x03030303u << (6 - 2 * g);   // k-group g sits in bits (7-2g .. 6-2g)
                    mma_u8s8_16832(acc[g], Wv.x & mask, Wv.y & mask, Wv.z & mask, Wv.w & mask, b[2 * g], b[2 * g + 1]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
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
                for (int i = tid; i < p.grp * 128; i += SV_THREADS) {
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
                    for (int i = tid; i < p.grp * 128; i += SV_THREADS) {
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
                    for (int i = tid; i < p.M * 16; i += SV_THREADS) {
                        const int m = i >> 4, r = i & 15;
                        epilogue_store_relu2(p.ep, m, nbase + r, sg[m * 16 + r], sg[128 + m * 16 + r]);
                    }
                } else {
                    for (int i = tid; i < p.M * rows_per_grp; i += SV_THREADS) {
                        const int m = i / rows_per_grp, rr = i % rows_per_grp;
                        const int n = nbase + rr;
                        if (n < p.N) epilogue_store(p.ep, m, n, sg[(rr >> 4) * 128 + m * 16 + (rr & 15)]);
                    }
                }
            }
        }
    }
    pdl_launch_dependents();
}

static size_t g_stream_attr_smem = 0;
void gemv_stream_init_attrs() {
    if (cudaFuncSetAttribute(ternary_gemv_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) == cudaSuccess)
        g_stream_attr_smem = 220 * 1024;
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
    const size_t smem = (size_t)SV_STAGES * SV_STAGE_BYTES + (((size_t)p.M * p.act_stride + 127) & ~(size_t)127) + (size_t)p.sacc_rbs * 512;
    if (smem > 220 * 1024) { set_error("gemv_stream: M*K=%d*%d does not fit in shared memory", a->M, a->K); return WF_ERR_UNSUPPORTED; }
    if (smem > g_stream_attr_smem) {
        if (cuda_check(cudaFuncSetAttribute(ternary_gemv_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024), "smem attr")) return WF_ERR_CUDA;
        g_stream_attr_smem = 220 * 1024;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(SV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, ternary_gemv_stream_kernel, p), "ternary_gemv_stream");
}

}  // namespace wf
