// Persistent "layer tail" kernel, batched form (8 < M <= 64): the same four linear phases as layer_tail.cu
// (o -> gate|up -> down -> next layer's q|k|v) in ONE launch, with the per-row glue and the batched ternary GEMM inside.
//
// Per phase:   [grid barrier]  row glue: CTA r < M normalises / quantises activation row r once (not per consumer CTA)
//              [grid barrier]  GEMM: S[m, n] = sum_k code[n, k] * a[m, k] for this CTA's row-blocks, every k, all M tokens
//                              epilogue: plain stores of the raw accumulators, or h = relu(gate)^2 * up for gate|up
// (reference: one ggml_vec_dot_i2_i8_s per (row, token), extern/BitNet/src/ggml-bitnet-mad.cpp:94-186, under bitnet_gemm's
// OpenMP loop; glue order scripts/direct_inference_full.py:210-285.)
//
// B200 mapping:
//   * persistent CTA per SM, 18 warps: 16 consumers (4 token groups of 16 x 4 k-slots), one activation loader, one TMA
//     producer.  Whole 16-row blocks per CTA (no split-K across CTAs: every output is stored once by its owner, nothing is
//     zeroed, no fp32 RED round trip through L2 — the 4*M*N bytes per linear of the previous design are gone);
//   * K is walked in chunks of 4 k-blocks: the TMA producer lands the chunk's weight strips of ALL the CTA's row-blocks in one
//     ring slot (cp.async.bulk, mbarrier full/empty) and runs ahead across phase boundaries — weights do not depend on the
//     previous phase, so they stream while the consumers sit in a barrier or in the glue; the loader warp brings the 64 x 512 B
//     int8 activation chunk with cp.async (L2 only) into a 3-slot ring once the phase's rows are published;
//   * consumer warp (tg, ks): the B fragments (16 tokens x 128 k) of ITS k-block are loaded once per chunk with ldmatrix
//     and reused for every row-block of the CTA (up to 6): mma.sync.m16n8k32.u8.s8 with the 2-bit fields extracted by one
//     shift+AND per fragment; accumulators stay in registers over the whole K; the four k-slots meet in shared memory.
#include "fused_common.cuh"
#include <string.h>

namespace wf {

constexpr int LB_CW = 16;
constexpr int LB_CT = LB_CW * 32;            // 512 consumer threads
constexpr int LB_THREADS = LB_CT + 64;       // + loader warp (16) + producer warp (17)
constexpr int LB_KC = 4;                     // k-blocks per chunk = k-slots
constexpr int LB_RBMAX = 4;                  // row-blocks per pass (accumulators: 4 x 2 x 4 registers; 18 warps are allocated as 20 -> 96 registers per thread)
constexpr int LB_WSLOT = LB_RBMAX * LB_KC * WF_TILE_BYTES;   // 12 KB
constexpr int LB_WSLOTS = 10;
constexpr int LB_ASTRIDE = QROW_BYTES;                       // 512 + 16-byte skew: conflict-free ldmatrix
constexpr int LB_ASLOT = QCHUNK_BYTES;                       // 33 KB: one chunk of the glue's chunk-major image (QMAP_CHUNKED)
static_assert(LB_ASTRIDE == LB_KC * 128 + 16, "chunk = LB_KC k-blocks");
constexpr int LB_ASLOTS = 3;
constexpr int LB_SS = LB_RBMAX * 16 + 4;                     // reduction buffer row stride (words)
constexpr int LB_RED_BYTES = 64 * LB_SS * 4;                 // 25.6 KB (also the glue's bf16 row scratch)

struct BatchPhase {
    GlueParams g;            // q_out / scale_out / rowsum_out: the published int8 rows of this phase
    const uint8_t* w;
    int N, K, KB, RB, rb_unit, rot, epi;
    float* acc;
    int acc_ld;
    __nv_bfloat16* h_out;
    int h_ld;
    const float* h_wscale;
};
struct BatchParams {
    int nph, M;
    BatchPhase ph[LT_MAX_PHASES];
    unsigned long long* bar;     // [2 * LT_MAX_PHASES]
    int* epoch;
    int bump_epoch;
    int* err;
    unsigned long long* trace;   // WF_TRACE: globaltimer taps of CTA 0 (same slots as layer_tail.cu)
};

__device__ __forceinline__ void lb_slab(const BatchPhase& ph, int cta, int grid, int& rb_lo, int& rb_hi) {
    const int units = ph.RB / ph.rb_unit;
    int c = cta + ph.rot;
    if (c >= grid) c -= grid;
    rb_lo = (int)((long long)c * units / grid) * ph.rb_unit;
    rb_hi = (int)((long long)(c + 1) * units / grid) * ph.rb_unit;
}

__device__ __forceinline__ void lb_grid_barrier(unsigned long long* slot, unsigned long long target, int* err) {
    group_bar<1, LB_CT>();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(slot), "l"(1ull) : "memory");
        unsigned long long v;
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
        if (v < target && !(err && ld_relaxed_gpu(err))) {
            const unsigned long long t0 = spin_now();
            for (;;) {
                asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
                if (v >= target) break;
                if (spin_now() - t0 > WF_SPIN_BUDGET_NS) { if (err) atomicExch(err, 2); break; }
            }
        }
    }
    group_bar<1, LB_CT>();
}

__global__ void __launch_bounds__(LB_THREADS, 1) layer_tail_batch_kernel(const __grid_constant__ BatchParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* wring = smem;                                        // LB_WSLOTS * LB_WSLOT
    unsigned char* aring = wring + LB_WSLOTS * LB_WSLOT;                // LB_ASLOTS * LB_ASLOT
    int* red = reinterpret_cast<int*>(aring + LB_ASLOTS * LB_ASLOT);    // [64][LB_SS]
    __shared__ __align__(8) uint64_t wfull[LB_WSLOTS], wempty[LB_WSLOTS], afull[LB_ASLOTS], aempty[LB_ASLOTS], phase_bar;
    __shared__ float fs[32];
    __shared__ int is[32];
    __shared__ float s_scale[64];
    __shared__ int s_rsum[64];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int M = p.M, grid = gridDim.x, cta = blockIdx.x;

    if (tid == 0) {
        for (int s = 0; s < LB_WSLOTS; ++s) { mbar_init(&wfull[s], 1); mbar_init(&wempty[s], LB_CW); }
        for (int s = 0; s < LB_ASLOTS; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], LB_CW); }
        mbar_init(&phase_bar, 1);
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == LB_CW + 1) {
        // ------------------------------------------------ TMA producer: weight strips of every (phase, pass, chunk), running ahead
        if (lane == 0) {
            int s = 0, bit = 0;
            for (int pi = 0; pi < p.nph; ++pi) {
                const BatchPhase& ph = p.ph[pi];
                int rb_lo, rb_hi;
                lb_slab(ph, cta, grid, rb_lo, rb_hi);
                const int nchunks = (ph.KB + LB_KC - 1) / LB_KC;
                for (int rb0 = rb_lo; rb0 < rb_hi; rb0 += LB_RBMAX) {
                    const int nr = min(LB_RBMAX, rb_hi - rb0);
                    for (int ch = 0; ch < nchunks; ++ch) {
                        const int kcl = min(LB_KC, ph.KB - ch * LB_KC);
                        mbar_wait(&wempty[s], bit ^ 1);
                        mbar_expect_tx(&wfull[s], (uint32_t)(nr * kcl * WF_TILE_BYTES));
                        for (int i = 0; i < nr; ++i)
                            bulk_g2s(wring + s * LB_WSLOT + i * LB_KC * WF_TILE_BYTES,
                                     ph.w + ((size_t)(rb0 + i) * ph.KB + (size_t)ch * LB_KC) * WF_TILE_BYTES, (uint32_t)(kcl * WF_TILE_BYTES), &wfull[s]);
                        if (++s == LB_WSLOTS) { s = 0; bit ^= 1; }
                    }
                }
            }
        }
        return;
    }
    if (warp == LB_CW) {
        // ------------------------------------------------ activation producer: the phase's published int8 rows, one bulk copy per
        // chunk (the glue wrote them chunk-major with the ring's row skew, so rows 0..M-1 of a chunk are one contiguous range;
        // rows >= M of a slot keep whatever they held — their accumulators are never stored)
        if (lane == 0) {
            int s = 0, bit = 0;
            const uint32_t bytes = (uint32_t)M * LB_ASTRIDE;
            for (int pi = 0; pi < p.nph; ++pi) {
                const BatchPhase& ph = p.ph[pi];
                int rb_lo, rb_hi;
                lb_slab(ph, cta, grid, rb_lo, rb_hi);
                mbar_wait(&phase_bar, pi & 1);          // the rows of this phase are published (second grid barrier of the phase passed)
                fence_proxy_async_global();             // generic-proxy stores of the glue CTAs -> this thread's async-proxy reads
                const int nchunks = (ph.KB + LB_KC - 1) / LB_KC;
                const int8_t* q = ph.g.q_out;
                for (int rb0 = rb_lo; rb0 < rb_hi; rb0 += LB_RBMAX) {
                    for (int ch = 0; ch < nchunks; ++ch) {
                        mbar_wait(&aempty[s], bit ^ 1);
                        mbar_expect_tx(&afull[s], bytes);
                        bulk_g2s(aring + s * LB_ASLOT, q + (size_t)ch * LB_ASLOT, bytes, &afull[s]);
                        if (++s == LB_ASLOTS) { s = 0; bit ^= 1; }
                    }
                }
            }
        }
        return;
    }

    // ---------------------------------------------------- consumer warps
    const int tg = warp & 3, ks = warp >> 2;            // token group (16 tokens), k-slot
    const bool tg_live = tg * 16 < M;                   // a token group past the batch has nothing to compute
    unsigned long long* tr = (p.trace && cta == 0 && tid == 0) ? p.trace : nullptr;
    if (tr) tr[0] = gtimer();
    pdl_wait_prior();
    if (tr) tr[1] = gtimer();
    const unsigned long long bar_target = (unsigned long long)grid * (unsigned long long)(ld_relaxed_gpu(p.epoch) + (p.bump_epoch ? 1 : 0));
    int ws = 0, wbit = 0, as = 0, abit = 0;
    int nbar = 0;

    for (int pi = 0; pi < p.nph; ++pi) {
        const BatchPhase& ph = p.ph[pi];
        if (pi > 0) {
            lb_grid_barrier(p.bar + nbar, bar_target, p.err);   // the previous phase's outputs are complete
            ++nbar;
            if (pi == 1) pdl_launch_dependents();               // every CTA of this grid is resident: dependents can no longer starve the barriers
        }
        if (tr) tr[2 + 5 * pi] = gtimer();
        // ---- row glue, once per row: CTA r handles row r (residual / RMSNorm / absmax / int8 codes -> global)
        if (cta < M) {
            float sc; int rsum;
            if (ph.g.K <= LB_CT * 8 * 2 && ph.g.mode != GL_RELU2 && ph.g.acc_own_cols == 0)      // the row stays in registers between the passes
                glue_row_regs<1, LB_CT, 2>(ph.g, cta, ph.g.q_out + (size_t)cta * LB_ASTRIDE, fs, is, true, sc, rsum, QMAP_CHUNKED);
            else
                // bf16 row scratch: the activation ring is idle between a phase's two grid barriers (its loader waits for phase_bar), and
                // holds rows of up to 50 K elements (8B / 70B ffn vectors) where the reduction buffer holds 13 K
                glue_row<1, LB_CT>(ph.g, cta, reinterpret_cast<__nv_bfloat16*>(aring), ph.g.q_out + (size_t)cta * LB_ASTRIDE, fs, is, true, sc, rsum, nullptr, QMAP_CHUNKED);
            if (tid == 0) { ph.g.scale_out[cta] = sc; ph.g.rowsum_out[cta] = rsum; }
            fence_proxy_async_global();                         // the codes are read back by bulk copies (async proxy)
        }
        if (tr) tr[3 + 5 * pi] = gtimer();
        lb_grid_barrier(p.bar + nbar, bar_target, p.err);       // every row is published
        ++nbar;
        if (pi == 0 && p.nph == 1) pdl_launch_dependents();
        if (tid == 0) mbar_arrive(&phase_bar);                  // release the activation producer
        if (tr) tr[4 + 5 * pi] = gtimer();
        if (ph.epi == 1) {
            for (int m = tid; m < M; m += LB_CT) { s_scale[m] = ldcg_f(ph.g.scale_out + m); s_rsum[m] = ldcg_i32(ph.g.rowsum_out + m); }
        }

        int rb_lo, rb_hi;
        lb_slab(ph, cta, grid, rb_lo, rb_hi);
        const int nchunks = (ph.KB + LB_KC - 1) / LB_KC;
        for (int rb0 = rb_lo; rb0 < rb_hi; rb0 += LB_RBMAX) {
            const int nr = min(LB_RBMAX, rb_hi - rb0);
            int acc[LB_RBMAX][2][4];
#pragma unroll
            for (int i = 0; i < LB_RBMAX; ++i)
#pragma unroll
                for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][nt][j] = 0;
            for (int ch = 0; ch < nchunks; ++ch) {
                mbar_wait(&wfull[ws], wbit);
                mbar_wait(&afull[as], abit);
                if (tg_live && ch * LB_KC + ks < ph.KB) {
                    const unsigned char* ab = aring + as * LB_ASLOT;
                    uint32_t b[2][8];
#pragma unroll
                    for (int nt = 0; nt < 2; ++nt) {
                        const unsigned char* ap = ab + (size_t)((tg * 2 + nt) * 8 + (lane & 7)) * LB_ASTRIDE + ks * 128 + (lane >> 3) * 16;
                        ldmatrix_x4(b[nt][0], b[nt][1], b[nt][2], b[nt][3], ap);
                        ldmatrix_x4(b[nt][4], b[nt][5], b[nt][6], b[nt][7], ap + 64);
                    }
                    const unsigned char* wb = wring + ws * LB_WSLOT + ks * WF_TILE_BYTES + lane * 16;
#pragma unroll
                    for (int i = 0; i < LB_RBMAX; ++i) {
                        if (i < nr) {
                            const uint4 Wv = *reinterpret_cast<const uint4*>(wb + i * LB_KC * WF_TILE_BYTES);
#pragma unroll
                            for (int g = 0; g < 4; ++g) {
                                const int sh = 6 - 2 * g;   // field of k-group g (ggml-bitnet-mad.cpp:76-81)
                                const uint32_t a0 = (Wv.x >> sh) & 0x03030303u, a1 = (Wv.y >> sh) & 0x03030303u;
                                const uint32_t a2 = (Wv.z >> sh) & 0x03030303u, a3 = (Wv.w >> sh) & 0x03030303u;
                                mma_u8s8_16832(acc[i][0], a0, a1, a2, a3, b[0][2 * g], b[0][2 * g + 1]);
                                mma_u8s8_16832(acc[i][1], a0, a1, a2, a3, b[1][2 * g], b[1][2 * g + 1]);
                            }
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) { mbar_arrive(&wempty[ws]); mbar_arrive(&aempty[as]); }
                if (++ws == LB_WSLOTS) { ws = 0; wbit ^= 1; }
                if (++as == LB_ASLOTS) { as = 0; abit ^= 1; }
            }
            // ---- the four k-slots meet in shared memory: red[m][row] (row = local row of this pass)
            {
                const int g = lane >> 2, t = lane & 3;
#pragma unroll
                for (int r = 0; r < LB_KC; ++r) {
                    if (ks == r && tg_live) {
#pragma unroll
                        for (int i = 0; i < LB_RBMAX; ++i) {
                            if (i < nr) {
#pragma unroll
                                for (int nt = 0; nt < 2; ++nt) {
                                    const int m0 = (tg * 2 + nt) * 8 + 2 * t;
                                    int* d0 = red + m0 * LB_SS + i * 16 + g;
                                    int* d1 = d0 + LB_SS;
                                    if (r == 0) { d0[0] = acc[i][nt][0]; d1[0] = acc[i][nt][1]; d0[8] = acc[i][nt][2]; d1[8] = acc[i][nt][3]; }
                                    else { d0[0] += acc[i][nt][0]; d1[0] += acc[i][nt][1]; d0[8] += acc[i][nt][2]; d1[8] += acc[i][nt][3]; }
                                }
                            }
                        }
                    }
                    group_bar<1, LB_CT>();
                }
            }
            if (tr && rb0 + LB_RBMAX >= rb_hi) tr[5 + 5 * pi] = gtimer();
            // ---- epilogue: this CTA owns every k of these rows — plain stores
            if (ph.epi == 0) {
                const int quads = nr * 4;
                for (int idx = tid; idx < M * quads; idx += LB_CT) {
                    const int m = idx / quads, c = idx - m * quads;
                    const int n = rb0 * 16 + 4 * c;
                    if (n < ph.N) {
                        const int4 v = *reinterpret_cast<const int4*>(red + m * LB_SS + 4 * c);
                        float* dst = ph.acc + (size_t)m * ph.acc_ld + n;
                        if (n + 3 < ph.N && (ph.acc_ld & 3) == 0) *reinterpret_cast<float4*>(dst) = make_float4((float)v.x, (float)v.y, (float)v.z, (float)v.w);
                        else {
                            const int vv[4] = {v.x, v.y, v.z, v.w};
                            for (int e = 0; e < 4; ++e) if (n + e < ph.N) dst[e] = (float)vv[e];
                        }
                    }
                }
            } else {
                // gate|up pairs: h = bf16(bf16(relu(g)^2) * u)   (direct_inference_full.py:259)
                const float wg = ph.h_wscale[0], wu = ph.h_wscale[1];
                const int npairs = nr >> 1;
                for (int idx = tid; idx < M * npairs * 4; idx += LB_CT) {
                    const int m = idx / (npairs * 4), c = idx - m * (npairs * 4);
                    const int pr = c >> 2, r = (c & 3) * 4;
                    const int4 G = *reinterpret_cast<const int4*>(red + m * LB_SS + (2 * pr) * 16 + r);
                    const int4 U = *reinterpret_cast<const int4*>(red + m * LB_SS + (2 * pr + 1) * 16 + r);
                    const int gv[4] = {G.x, G.y, G.z, G.w}, uv[4] = {U.x, U.y, U.z, U.w};
                    const int rs = s_rsum[m];
                    const float sa = s_scale[m];
                    float x[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float gg = bf16_round(lin_value((float)gv[e], rs, wg, sa));
                        const float uu = bf16_round(lin_value((float)uv[e], rs, wu, sa));
                        const float tt = fmaxf(gg, 0.f);
                        x[e] = bf16_round(bf16_round(tt * tt) * uu);
                    }
                    const int col = ((rb0 >> 1) + pr) * 16 + r;
                    *reinterpret_cast<uint2*>(ph.h_out + (size_t)m * ph.h_ld + col) = make_uint2(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]));
                }
            }
            group_bar<1, LB_CT>();      // red is reused by the next pass / the next phase's glue
        }
        if (tr) tr[6 + 5 * pi] = gtimer();
    }
    if (p.bump_epoch && cta == 0 && tid == 0) atomicAdd(p.epoch, 1);   // after CTA 0's last barrier: every CTA has read the epoch
}

// ---------------------------------------------------------------------------------------------------------------
constexpr size_t LB_SMEM = (size_t)LB_WSLOTS * LB_WSLOT + (size_t)LB_ASLOTS * LB_ASLOT + LB_RED_BYTES;

bool layer_tail_batch_fits(int M, int Kmax) {
    return M >= 1 && M <= 64 && (size_t)Kmax * 2 <= (size_t)LB_ASLOTS * LB_ASLOT;
}

int launch_layer_tail_batch(const TailArgs& a, cudaStream_t s, bool pdl) {
    if (a.nph < 1 || a.nph > LT_MAX_PHASES || a.M < 1 || a.M > 64) { set_error("layer_tail_batch: bad phase count / batch"); return WF_ERR_INVALID; }
    BatchParams p;
    memset(&p, 0, sizeof p);
    p.nph = a.nph; p.M = a.M;
    const int grid = sm_count();
    if (a.M > grid) { set_error("layer_tail_batch: more rows than CTAs"); return WF_ERR_UNSUPPORTED; }
    long long extra = 0;
    for (int i = 0; i < a.nph; ++i) {
        const TailPhaseArgs& x = a.ph[i];
        if (x.K % WF_QK || x.g.K != x.K || (size_t)x.K * 2 > (size_t)LB_ASLOTS * LB_ASLOT) { set_error("layer_tail_batch: bad K"); return WF_ERR_UNSUPPORTED; }
        if (!x.g.q_out || !x.g.scale_out || !x.g.rowsum_out) { set_error("layer_tail_batch: the glue needs q_out / scale_out / rowsum_out"); return WF_ERR_INVALID; }
        BatchPhase& ph = p.ph[i];
        ph.g = x.g; ph.g.M = a.M;
        ph.w = x.w; ph.N = x.N; ph.K = x.K; ph.KB = x.K / WF_QK; ph.RB = (x.N + 15) / 16;
        ph.rb_unit = x.relu2 ? 2 : 1;
        if (ph.RB % ph.rb_unit) { set_error("layer_tail_batch: gate|up needs an even number of row blocks"); return WF_ERR_INVALID; }
        ph.epi = x.relu2 ? 1 : 0;
        ph.acc = x.acc; ph.acc_ld = x.acc_ld;
        ph.h_out = x.h_out; ph.h_ld = x.h_ld; ph.h_wscale = x.h_wscale;
        const int units = ph.RB / ph.rb_unit;
        ph.rot = (int)(extra % grid);
        extra += units % grid;
    }
    p.bar = a.bar; p.epoch = a.epoch; p.bump_epoch = a.bump_epoch; p.err = a.err; p.trace = a.trace;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(LB_THREADS);
    cfg.dynamicSmemBytes = LB_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, layer_tail_batch_kernel, p), "layer_tail_batch");
}

void layer_tail_batch_init_attrs() {
    cudaFuncSetAttribute(layer_tail_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LB_SMEM);
}

}  // namespace wf
