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

constexpr int SV_TPW = 4;                                // tiles per consumer warp per stage
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
    int aligned;          // 1: CTA slabs are whole finalisation groups (no split-K, no workspace traffic)
    int stages;           // ring depth (<= SV_MAX_STAGES)
    int l2_ahead;         // producer prefetches this many stages beyond the ring into L2 (0 = off)
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
    constexpr int SV_STAGE_TILES = SV_TPW * CW, SV_STAGE_BYTES = SV_STAGE_TILES * WF_TILE_BYTES;   // a stage = SV_TPW adjacent tiles per consumer warp
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;                                        // stages * SV_STAGE_BYTES
    unsigned char* acts = ring + p.stages * SV_STAGE_BYTES;            // M * act_stride
    int* sacc = reinterpret_cast<int*>(acts + ((p.M * p.act_stride + 127) & ~127));   // [rbl][8][16]
    __shared__ __align__(8) uint64_t full_bar[SV_MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[SV_MAX_STAGES];
    __shared__ int s_fin[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    long long t0, t1;
    if (p.aligned) {   // whole groups of row-blocks per CTA
        const int G = (p.RB + p.grp - 1) / p.grp;
        t0 = ((long long)blockIdx.x * G / gridDim.x) * p.grp * p.KB;
        t1 = ((long long)(blockIdx.x + 1) * G / gridDim.x) * p.grp * p.KB;
        if (t1 > p.T) t1 = p.T;
    } else {           // even tile split (split-K across CTA boundaries)
        t0 = (long long)blockIdx.x * p.T / gridDim.x;
        t1 = (long long)(blockIdx.x + 1) * p.T / gridDim.x;
    }
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
                if (p.l2_ahead > 0) {
                    const int pf = it + S + p.l2_ahead - 1;   // first pass also covers [S, S + l2_ahead)
                    for (int j = (it == 0 ? S : pf); j <= pf && j < nstages; ++j) {
                        const int tl = min(SV_STAGE_TILES, ntiles - j * SV_STAGE_TILES);
                        bulk_prefetch_l2(p.w + (size_t)(t0 + (long long)j * SV_STAGE_TILES) * WF_TILE_BYTES, (uint32_t)tl * WF_TILE_BYTES);
                    }
                }
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

        // Tile -> warp mapping: a stage holds SV_TPW*CW tiles and warp w owns the SV_TPW adjacent tiles starting at
        // SV_TPW*w of every stage — all weight loads are issued before any math, the run almost always stays inside one
        // row-block (one flush per KB tiles), the per-stage bookkeeping is amortised over SV_TPW tiles and
        // (row-block, k-block) are tracked incrementally (no division on the per-tile path).
        constexpr int ST = SV_TPW * CW;
        int kb0 = (int)(t0 - (long long)rb_first * p.KB) + SV_TPW * warp, rb0 = rb_first;
        while (kb0 >= p.KB) { kb0 -= p.KB; ++rb0; }
        const int arow = min(lane & 7, p.M - 1);
        const unsigned char* abase = MR == 0 ? acts + arow * p.act_stride + (lane >> 3) * 16 : acts + (lane & 3) * 32;
        const int KB = p.KB;

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
                    const unsigned char* ap = abase + (m < p.M ? m : 0) * p.act_stride + kb * 128;
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
        const unsigned char* wbase = ring + warp * (SV_TPW * WF_TILE_BYTES) + lane * 16;
        for (int it = 0; it < nstages; ++it) {
            mbar_wait(&full_bar[s], ph);
            const int i0 = it * ST + SV_TPW * warp;
            if (i0 < ntiles && !(p.debug & 1)) {
                const unsigned char* wp = wbase + s * (ST * WF_TILE_BYTES);
                uint4 W[SV_TPW];
                if (i0 + SV_TPW <= ntiles) {
#pragma unroll
                    for (int j = 0; j < SV_TPW; ++j) W[j] = *reinterpret_cast<const uint4*>(wp + j * WF_TILE_BYTES);
                    if (kb0 + SV_TPW <= KB && rb0 == cur_rb) {
                        // fast path: the whole run sits in the current row-block
#pragma unroll
                        for (int j = 0; j < SV_TPW; ++j) tile_math(W[j], kb0 + j);
                    } else {
                        int kb = kb0, rb = rb0;
#pragma unroll
                        for (int j = 0; j < SV_TPW; ++j) {
                            if (rb != cur_rb) { flush(cur_rb); cur_rb = rb; }
                            tile_math(W[j], kb);
                            if (++kb == KB) { kb = 0; ++rb; }
                        }
                    }
                } else {
                    int kb = kb0, rb = rb0;
                    for (int j = 0; j < SV_TPW && i0 + j < ntiles; ++j) {
                        const uint4 Wv = *reinterpret_cast<const uint4*>(wp + j * WF_TILE_BYTES);
                        if (rb != cur_rb) { flush(cur_rb); cur_rb = rb; }
                        tile_math(Wv, kb);
                        if (++kb == KB) { kb = 0; ++rb; }
                    }
                }
            }
            kb0 += ST;
            while (kb0 >= KB) { kb0 -= KB; ++rb0; }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
            if (++s == S) { s = 0; ph ^= 1; }
        }
        flush(cur_rb);
    }
    __syncthreads();

    // ------------------------------------------------ combine + fused epilogue
    // Row-block groups cut by a CTA boundary (at most the first and the last of the slab) are combined through the
    // self-cleaning int32 workspace; whoever completes a group finalises it.  All groups this CTA finalises then go
    // through the fused epilogue in ONE parallel pass.
    if (ntiles > 0) {
        const int rb_last = (int)((t1 - 1) / p.KB);
        const int grp_tiles = p.grp * p.KB;
        const int rows_per_grp = p.grp * 16;
        const bool relu2 = (p.ep.out_mode == WF_OUT_LINEAR) && (p.ep.epi_flags & WF_EPI_RELU2_MUL);
        const int gfirst = rb_first / p.grp, glast = rb_last / p.grp;
        bool fin_first = true, fin_last = true;
        if (!p.aligned) {
            // both boundary groups go through ONE round of (partial sums -> fence -> arrival counter -> read back)
            int contrib[2], gid[2];
            bool part[2];
            for (int e = 0; e < 2; ++e) {
                gid[e] = e == 0 ? gfirst : glast;
                const long long g0 = (long long)gid[e] * grp_tiles, g1 = g0 + grp_tiles;
                contrib[e] = (int)((t1 < g1 ? t1 : g1) - (t0 > g0 ? t0 : g0));
                const long long gt = (g1 < p.T ? g1 : p.T) - g0;          // tiles the group really has (last group may be short)
                part[e] = contrib[e] != (int)gt && !(e == 1 && glast == gfirst);
            }
            if (part[0] || part[1]) {
                for (int e = 0; e < 2; ++e) {
                    if (!part[e]) continue;
                    const int* sg = sacc + (gid[e] * p.grp - rb_base) * 128;   // [grp rb][m][16]
                    const int nbase = gid[e] * rows_per_grp;
                    for (int i = tid; i < p.grp * 128; i += NT) {
                        const int rbl = i >> 7, m = (i >> 4) & 7, r = i & 15;
                        const int n = nbase + rbl * 16 + r;
                        const int v = sg[i];
                        if (m < p.M && n < p.N && v) atomicAdd(p.ws_acc + (size_t)m * p.N + n, v);
                    }
                }
                __threadfence();
                __syncthreads();
                if (tid < 2 && part[tid]) {
                    const long long g0 = (long long)gid[tid] * grp_tiles;
                    const int gt = (int)((g0 + grp_tiles < p.T ? g0 + grp_tiles : p.T) - g0);
                    const int prev = atomicAdd(p.ws_cnt + gid[tid], contrib[tid]);
                    s_fin[tid] = (prev + contrib[tid] == gt);
                    if (s_fin[tid]) p.ws_cnt[gid[tid]] = 0;
                }
                __syncthreads();
                if (part[0]) fin_first = s_fin[0] != 0;
                if (part[1]) fin_last = s_fin[1] != 0;
                if ((part[0] && fin_first) || (part[1] && fin_last)) __threadfence();
                for (int e = 0; e < 2; ++e) {
                    if (!part[e] || !(e == 0 ? fin_first : fin_last)) continue;
                    int* sg = sacc + (gid[e] * p.grp - rb_base) * 128;
                    const int nbase = gid[e] * rows_per_grp;
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
        }
        if (glast == gfirst) fin_last = fin_first;
        const int ngrp = glast - gfirst + 1;
        if (relu2) {
            for (int i = tid; i < ngrp * p.M * 16; i += NT) {
                const int gl = i / (p.M * 16), rem = i - gl * (p.M * 16);
                const int m = rem >> 4, r = rem & 15;
                const int grp = gfirst + gl;
                if ((grp == gfirst && !fin_first) || (grp == glast && !fin_last)) continue;
                const int* sg = sacc + (grp * 2 - rb_base) * 128;
                epilogue_store_relu2(p.ep, m, grp * 32 + r, sg[m * 16 + r], sg[128 + m * 16 + r]);
            }
        } else {   // grp == 1: groups are row-blocks
            for (int i = tid; i < ngrp * p.M * 16; i += NT) {
                const int gl = i / (p.M * 16), rem = i - gl * (p.M * 16);
                const int m = rem >> 4, r = rem & 15;
                const int grp = gfirst + gl;
                if ((grp == gfirst && !fin_first) || (grp == glast && !fin_last)) continue;
                const int n = grp * 16 + r;
                if (n < p.N) epilogue_store(p.ep, m, n, sacc[(grp - rb_base) * 128 + m * 16 + r]);
            }
        }
    }
}

int g_gemv_stages = 5;    // ring depth (16 KB stages) — wf_bench_gemv can override
int g_gemv_path = -1;     // -1 auto, 0 force the mma path
int g_gemv_l2_ahead = 0;  // L2 prefetch distance in stages
int g_gemv_grid_mult = 0; // CTAs per SM, 0 = auto
int g_gemv_aligned = 1;   // 0: always use the even tile split

typedef void (*stream_kernel_t)(const StreamParams);
constexpr int SV_CW = 8;   // consumer warps per CTA; large matrices run two CTAs per SM (two independent TMA streams)
static stream_kernel_t pick_kernel(int M) {
    // one activation row: CUDA cores (dp4a); 2..8 rows: the tensor path costs the same for any M <= 8
    if (g_gemv_path == 0 || M > 1) return ternary_gemv_stream_kernel<0, SV_CW>;
    return ternary_gemv_stream_kernel<1, SV_CW>;
}

static size_t g_stream_attr_smem = 0;
void gemv_stream_init_attrs() {
    bool ok = true;
    const stream_kernel_t all[2] = {ternary_gemv_stream_kernel<0, SV_CW>, ternary_gemv_stream_kernel<1, SV_CW>};
    for (stream_kernel_t k : all) ok &= cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) == cudaSuccess;
    if (ok) g_stream_attr_smem = 220 * 1024;
}

// the int8 activations of all M rows stay resident in shared memory next to the weight ring (>= 3 stages) and the accumulators
bool gemv_stream_fits(int M, int K) {
    if (M < 1 || M > 8) return false;
    const size_t act_bytes = (((size_t)M * (K + 16) + 127) & ~(size_t)127);
    return act_bytes + 3 * (size_t)SV_TPW * SV_CW * WF_TILE_BYTES + 24 * 1024 <= 220 * 1024;
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
    p.l2_ahead = g_gemv_l2_ahead;
    // Grid: two CTAs per SM (two independent TMA streams per SM: +15 % stream rate, measured) once every CTA still gets a few
    // stages and both fit in shared memory; one per SM otherwise.  Slabs: whole row-block groups per CTA when there are
    // enough of them and the matrix is small (no split-K: saves three dependent L2 round trips), an even tile split else.
    const int sms = sm_count();
    const size_t act_bytes = (((size_t)p.M * p.act_stride + 127) & ~(size_t)127);
    const size_t stage_bytes = (size_t)SV_TPW * SV_CW * WF_TILE_BYTES;
    int mult = g_gemv_grid_mult > 0 ? g_gemv_grid_mult : (p.T >= (long long)2 * sms * 3 * SV_TPW * SV_CW ? 2 : 1);
    p.stages = g_gemv_stages < 2 ? 2 : (g_gemv_stages > SV_MAX_STAGES ? SV_MAX_STAGES : g_gemv_stages);
    const int G = (p.RB + p.grp - 1) / p.grp;
    int grid = 0;
    size_t smem = 0;
    for (;;) {
        grid = sms * mult;
        const long long min_tiles = 4;
        if (p.T / min_tiles < grid) grid = (int)((p.T + min_tiles - 1) / min_tiles);
        if (grid < 1) grid = 1;
        // whole groups per CTA when that costs <= 12.5 % balance (the split-K round costs ~3 us of dependent L2 round trips)
        if (g_gemv_aligned && G < grid && G >= sms) grid = G;   // one whole group per CTA
        p.aligned = (g_gemv_aligned != 0) && G >= grid &&
                    ((p.T / grid) < (long long)8 * SV_TPW * SV_CW || (long long)((G + grid - 1) / grid) * grid * 8 <= (long long)G * 9);
        const long long max_tiles = p.aligned ? (long long)((G + grid - 1) / grid + 1) * p.grp * p.KB : (p.T + grid - 1) / grid + 1;
        p.sacc_rbs = (int)(max_tiles / p.KB) + 2 + 2 * p.grp;
        const size_t fixed = act_bytes + (size_t)p.sacc_rbs * 512;
        if (mult == 2 && g_gemv_grid_mult == 0) {
            // both CTAs of an SM must fit: shrink the ring down to 3 stages before giving up the second CTA
            int st = p.stages;
            while (st > 3 && 2 * (st * stage_bytes + fixed + 1024) > 226 * 1024) --st;
            if (2 * (st * stage_bytes + fixed + 1024) > 226 * 1024) { mult = 1; p.stages = 8; continue; }
            p.stages = st;
        }
        while (p.stages > 2 && p.stages * stage_bytes + fixed > 220 * 1024) --p.stages;
        smem = p.stages * stage_bytes + fixed;
        break;
    }
    const size_t need_ws = (size_t)a->M * a->N * 4 + (size_t)p.RB * 4;
    if (!a->workspace || a->workspace_bytes < need_ws) { set_error("gemv_stream: workspace too small (%zu < %zu)", a->workspace_bytes, need_ws); return WF_ERR_INVALID; }
    p.ws_acc = (int*)a->workspace;
    p.ws_cnt = (int*)((char*)a->workspace + (size_t)a->M * a->N * 4);
    stream_kernel_t kern = pick_kernel(a->M);
    const int threads = (SV_CW + 1) * 32;
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
