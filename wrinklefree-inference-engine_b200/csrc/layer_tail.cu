// Persistent "layer tail" kernel (M <= 8): every linear phase between two attention kernels in ONE launch.
//
//   phase 0  attn_sub_norm + quant              -> o-proj            -> raw fp32 accumulators acc_o
//   phase 1  hidden + o -> post_attention_norm   -> gate|up           -> h = relu(gate)^2 * up (bf16, epilogue)
//   phase 2  ffn_sub_norm + quant                -> down-proj         -> acc_down
//   phase 3  hidden + down -> input_layernorm    -> q|k|v of layer+1  -> acc_qkv          (absent after the last layer)
// (scripts/direct_inference_full.py:186-208 per linear, :210-264 / :267-285 for the order of the glue).  The step then is
// embed-phase, { attention, tail } x layers, final norm + LM head: 2 launches per layer instead of 5.
//
// B200 mapping:
//   * one persistent CTA per SM; phases are separated by a GRID BARRIER (release-add on a global counter + acquire spin,
//     bounded), not by a kernel boundary;
//   * one producer warp streams this CTA's slab of EVERY phase's matrix through one TMA ring (cp.async.bulk 1-D, 12 KB
//     stages, mbarrier full/empty) without ever stopping at a phase boundary: while the consumers sit in a barrier or in the
//     norm / quant prologue of phase p, the weights of phases p+1.. are already landing in shared memory.  The ring (up to
//     ~170 KB per SM, 25 MB per chip) holds more than a 2B layer's whole tail, so HBM streaming is decoupled from the
//     dependency chain; the producer starts BEFORE griddepcontrol.wait, i.e. while the attention kernel is still running;
//   * slabs are WHOLE 16-row blocks (pairs of blocks for the interleaved gate|up matrix): a CTA owns every k of its rows, so
//     outputs are written once by their owner — no split-K, no atomics, no zeroing duty — and gate|up can apply
//     relu(gate)^2 * up in its epilogue; the extra CTAs of a non-divisible phase are rotated so that the bytes a CTA streams
//     over the whole tail stay balanced;
//   * batch 1 (the north-star shape) goes one step further: NO grid barrier between the phases.  The accumulator / h rows are
//     hand-off buffers that hold a sentinel until their owner stores the data; the consumer's row glue polls the words it needs
//     (one L2 hop instead of release fence + atomic arrival + acquire poll + load), every CTA keeps the residual row and the
//     previous phase's activation scale / rowsum locally (it computed them itself), and owners re-arm the sentinels two phases
//     after the last reader is known to be done: acc_o rows in the down epilogue, h and acc_down rows in the o epilogue;
//   * the norm / quant prologue (fused_common.cuh glue_row) runs redundantly in every CTA (<= 8 rows), results stay in
//     shared memory; consumers: dp4a on the tile words (M = 1) or mma.sync.m16n8k32 with the batch in the n dimension.
#include "fused_common.cuh"
#include <string.h>

namespace wf {

constexpr int LT_CW = 12;                                 // consumer warps (+ the producer warp = 12 warps: the register file is handed out per 4 warps, so 13 warps are budgeted as 16 = 128 registers per thread, 12 as 168)
constexpr int LT_TPW = 2;                                 // adjacent tiles per consumer warp per stage
constexpr int LT_STAGE_TILES = LT_TPW * LT_CW;            // 24 tiles = 12 KB per stage
constexpr int LT_STAGE_BYTES = LT_STAGE_TILES * WF_TILE_BYTES;
constexpr int LT_CT = LT_CW * 32;
constexpr int LT_THREADS = LT_CT + 32;
constexpr int LT_MAX_STAGES = 16;

struct TailPhase {
    GlueParams g;            // input glue; g.K = length of the activation vector (== K of the matrix: single-GPU engines)
    const uint8_t* w;        // b200_tile_v1 [N][K]
    int N, K, KB, RB;
    int rb_unit;             // row-blocks are handed out in units of rb_unit (2 = gate|up pairs)
    int rot;                 // rotation of the CTA -> unit assignment (balances the bytes per CTA over the tail)
    int epi;                 // 0: raw fp32 accumulators [M][acc_ld]; 1: h = relu(gate)^2 * up, bf16 [M][h_ld]
    float* acc;
    int acc_ld;
    float* lin_ascale;       // [M] out: activation scale of this phase's input (read by the consumer of acc)
    int32_t* lin_rowsum;     // [M] out: Sigma a
    __nv_bfloat16* h_out;
    int h_ld;
    const float* h_wscale;   // device [2]: gate, up
};

struct TailParams {
    int nph, M, stages, act_stride, sacc_rbs, ys_elems, xs_elems;   // xs: flag-in-data form, the residual row kept from phase 1 to phase 3
    TailPhase ph[LT_MAX_PHASES];
    unsigned long long* bar;     // [LT_MAX_PHASES] monotonic arrival counters of this launch site
    int* epoch;                  // tail-path decode steps so far (device counter): barrier target = grid * epoch
    int bump_epoch;              // the first launch of a step (single phase, no barrier) increments it
    int lam;                     // 1 (M = 1, every K <= 9216): phases hand their outputs over flag-in-data (fused_common.cuh HandOff), no grid barrier
    int* err;
    unsigned long long* trace;   // WF_TRACE timeline taps (CTA 0): [0] entry, [1] dependency wait done, then per phase 5 stamps
};

__device__ __forceinline__ int lt_dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// this CTA's row-block range of a phase
__device__ __forceinline__ void lt_slab(const TailPhase& ph, int cta, int grid, int& rb_lo, int& rb_hi) {
    const int units = ph.RB / ph.rb_unit;
    int c = cta + ph.rot;
    if (c >= grid) c -= grid;
    rb_lo = (int)((long long)c * units / grid) * ph.rb_unit;
    rb_hi = (int)((long long)(c + 1) * units / grid) * ph.rb_unit;
}

// Grid barrier for the consumer threads of all CTAs (the producer warps keep streaming).  Arrival = release-add, wait =
// acquire spin; counters are monotonic over decode steps (no reset, CUDA-graph replay safe).  Bounded like every spin here.
__device__ __forceinline__ void lt_grid_barrier(unsigned long long* slot, unsigned long long target, int* err) {
    group_bar<1, LT_CT>();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(slot), "l"(1ull) : "memory");
        unsigned long long v;
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
        if (v < target && !(err && ld_relaxed_gpu(err))) {      // after one timeout every later barrier falls through: the step is lost anyway
            const unsigned long long t0 = spin_now();
            for (;;) {
                asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
                if (v >= target) break;
                if (spin_now() - t0 > WF_SPIN_BUDGET_NS) { if (err) atomicExch(err, 1); break; }
            }
        }
    }
    group_bar<1, LT_CT>();
}

template <int MR, bool LAM>
__global__ void __launch_bounds__(LT_THREADS, 1) layer_tail_kernel(const __grid_constant__ TailParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;                                                       // stages * LT_STAGE_BYTES
    unsigned char* acts = ring + p.stages * LT_STAGE_BYTES;                           // M * act_stride
    __nv_bfloat16* ys = reinterpret_cast<__nv_bfloat16*>(acts + ((p.M * p.act_stride + 127) & ~127));
    __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<unsigned char*>(ys) + ((p.ys_elems * 2 + 127) & ~127));
    int* sacc = reinterpret_cast<int*>(reinterpret_cast<unsigned char*>(xs) + ((p.xs_elems * 2 + 127) & ~127));   // [rbl][8][16]
    __shared__ __align__(8) uint64_t full_bar[LT_MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[LT_MAX_STAGES];
    __shared__ float fs[32];
    __shared__ int is[32];
    __shared__ float s_scale[8];
    __shared__ int s_rsum[8];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int M = p.M;
    const int grid = gridDim.x, cta = blockIdx.x;

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], LT_CW);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == LT_CW) {
        // ------------------------------------------------ producer warp (one elected lane): never waits on anything but ring space
        if (lane == 0) {
            int s = 0, ph_bit = 0;
            for (int pi = 0; pi < p.nph; ++pi) {
                const TailPhase& ph = p.ph[pi];
                int rb_lo, rb_hi;
                lt_slab(ph, cta, grid, rb_lo, rb_hi);
                const int ntiles = (rb_hi - rb_lo) * ph.KB;
                const unsigned char* src = ph.w + (size_t)rb_lo * ph.KB * WF_TILE_BYTES;
                for (int t = 0; t < ntiles; t += LT_STAGE_TILES) {
                    mbar_wait(&empty_bar[s], ph_bit ^ 1);
                    const uint32_t bytes = (uint32_t)min(LT_STAGE_TILES, ntiles - t) * WF_TILE_BYTES;
                    mbar_expect_tx(&full_bar[s], bytes);
                    bulk_g2s(ring + s * LT_STAGE_BYTES, src + (size_t)t * WF_TILE_BYTES, bytes, &full_bar[s]);
                    if (++s == p.stages) { s = 0; ph_bit ^= 1; }
                }
            }
        }
        return;
    }

    // ---------------------------------------------------- consumer warps
    const bool tr = p.trace != nullptr && cta == 0 && tid == 0;
    if (tr) p.trace[0] = gtimer();
    glue_prefetch_gamma<LT_CT>(p.ph[0].g);
    pdl_wait_prior();                                   // the attention kernel's output (or the previous step) is complete
    if (tr) p.trace[1] = gtimer();
    // Barrier counters are monotonic over the decode steps that take this path (no reset, CUDA-graph replay safe).  The epoch is
    // bumped by the step's first launch, which finished (transitively, through the attention kernel) before this one got here.
    if (p.bump_epoch && cta == 0 && tid == 0) atomicAdd(p.epoch, 1);
    const unsigned long long bar_target = (unsigned long long)grid * (unsigned long long)ld_relaxed_gpu(p.epoch);
    int s = 0, ph_bit = 0;                              // ring position, continues across phases
    // flag-in-data form: re-arm the sentinels of the rows THIS CTA owns in a hand-off buffer, once every reader of the buffer is known to
    // be done and (at least a phase) before this CTA stores the next data there
    auto arm = [&](const TailPhase& q, bool h_buf) {
        int lo, hi;
        lt_slab(q, cta, grid, lo, hi);
        if (h_buf) {
            uint32_t* dst = reinterpret_cast<uint32_t*>(q.h_out) + (lo >> 1) * 8;            // 16 bf16 = 8 words per gate|up pair
            for (int i = tid; i < ((hi - lo) >> 1) * 8; i += LT_CT) dst[i] = WF_HANDOFF_SENTINEL;
        } else {
            uint32_t* dst = reinterpret_cast<uint32_t*>(q.acc) + lo * 16;
            const int n = min(hi * 16, q.N) - lo * 16;
            for (int i = tid; i < n; i += LT_CT) dst[i] = WF_HANDOFF_SENTINEL;
        }
    };
    // h and acc_down: their readers were the previous launch (a kernel boundary ago); this CTA writes them in phases 1 and 2
    if (LAM) { arm(p.ph[1], true); arm(p.ph[2], false); }
    const int arow = min(lane & 7, M - 1);

    for (int pi = 0; pi < p.nph; ++pi) {
        const TailPhase& ph = p.ph[pi];
        constexpr bool lam = LAM;   // (its own instantiation: the hand-off code costs registers the barrier form does not have to pay)
        if (pi > 0) {
            glue_prefetch_gamma<LT_CT>(ph.g);
            if (!lam) {
                lt_grid_barrier(p.bar + (pi - 1), bar_target, p.err);
                // every CTA of this grid is resident now: letting the next kernel's CTAs in can no longer starve the barrier
                if (pi == 1) pdl_launch_dependents();
            }
        }
        if (tr) p.trace[2 + 5 * pi] = gtimer();          // barrier passed
        int rb_lo, rb_hi;
        lt_slab(ph, cta, grid, rb_lo, rb_hi);
        const int n_rbs = rb_hi - rb_lo;
        const int KB = ph.KB;
        const int ntiles = n_rbs * KB;
        const int nstages = (ntiles + LT_STAGE_TILES - 1) / LT_STAGE_TILES;

        // ---- prologue: x -> (residual) -> RMSNorm -> per-token int8, every row, into shared memory
        for (int i = tid; i < n_rbs * 128; i += LT_CT) sacc[i] = 0;
        for (int m = 0; m < M; ++m) {
            float sc; int rsum;
            // K <= 9216 (every 2B vector): the row stays in registers between the three passes
            if (lam) {
                // phase 0 reads the attention kernel's output (kernel boundary); 1..3 read what phase pi-1 is still writing.  The residual
                // row of phase 1 (hid + o) stays in shared memory for phase 3; scale / rowsum of the producing linear are this CTA's own.
                HandOff ho;
                ho.poll = pi > 0; ho.use_prev = pi > 0; ho.prev_scale = s_scale[0]; ho.prev_rowsum = s_rsum[0];
                ho.hid_sm_in = (pi == 3) ? xs : nullptr;
                ho.hid_sm_out = (pi == 1) ? xs : nullptr;
                ho.err = p.err; ho.tag = 300 + pi;
                glue_row_regs<1, LT_CT, 3>(ph.g, m, reinterpret_cast<int8_t*>(acts + m * p.act_stride), fs, is, cta == 0, sc, rsum, MR == 1, &ho);
            } else if (ph.g.K <= LT_CT * 8 * 3 && ph.g.mode != GL_RELU2 && ph.g.acc_own_cols == 0)
                glue_row_regs<1, LT_CT, 3>(ph.g, m, reinterpret_cast<int8_t*>(acts + m * p.act_stride), fs, is, cta == 0, sc, rsum, MR == 1);
            else
                glue_row<1, LT_CT>(ph.g, m, ys, reinterpret_cast<int8_t*>(acts + m * p.act_stride), fs, is, cta == 0, sc, rsum, nullptr, MR == 1);
            if (tid == 0) {
                s_scale[m] = sc; s_rsum[m] = rsum;
                if (cta == 0) { ph.lin_ascale[m] = sc; ph.lin_rowsum[m] = rsum; }
            }
        }
        group_bar<1, LT_CT>();     // activations (and the cleared accumulators) complete
        // flag-in-data form: having read every word of the previous phase's output means every CTA of the grid is resident (each wrote
        // its rows), so letting the next kernel's CTAs in can no longer starve this grid
        if (lam && pi == 1) pdl_launch_dependents();
        // acc_o: read by every CTA's phase-1 glue; all of them are past it, since this CTA has just read every word of h, which each
        // CTA stores after its own phase-1 glue.  Its owner writes it again in the next launch.
        if (lam && pi == 2) arm(p.ph[0], false);
        if (tr) p.trace[3 + 5 * pi] = gtimer();          // glue done

        // ---- main loop: S[m, n] += sum_k code[n, k] * a[m, k] over this CTA's row-blocks
        constexpr int NACC = MR == 0 ? 16 : 8;
        int acc[NACC];
#pragma unroll
        for (int j = 0; j < NACC; ++j) acc[j] = 0;
        int cur_rb = -1;
        auto flush = [&](int rbl) {
            if (rbl < 0) return;
            int* dst = sacc + rbl * 128;
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
                const unsigned char* ap = abase + kb * 128;
                const uint4 a0 = *reinterpret_cast<const uint4*>(ap);
                const uint4 a1 = *reinterpret_cast<const uint4*>(ap + 16);
                const uint32_t A[4][2] = {{a0.x, a0.y}, {a0.z, a0.w}, {a1.x, a1.y}, {a1.z, a1.w}};
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint32_t mask = 0x03030303u << (6 - 2 * g);
                    acc[g] = lt_dp4a_us(Wv.x & mask, A[g][0], acc[g]);
                    acc[4 + g] = lt_dp4a_us(Wv.y & mask, A[g][0], acc[4 + g]);
                    acc[g] = lt_dp4a_us(Wv.z & mask, A[g][1], acc[g]);
                    acc[4 + g] = lt_dp4a_us(Wv.w & mask, A[g][1], acc[4 + g]);
                }
            }
        };
        // stage = 24 tiles, warp w owns tiles 2w, 2w+1; (local row-block, k-block) tracked incrementally
        int kb0 = LT_TPW * warp, rbl0 = 0;
        while (kb0 >= KB) { kb0 -= KB; ++rbl0; }
        const unsigned char* wbase = ring + warp * (LT_TPW * WF_TILE_BYTES) + lane * 16;
        for (int it = 0; it < nstages; ++it) {
            mbar_wait(&full_bar[s], ph_bit);
            if (tr && it == 0) p.trace[4 + 5 * pi] = gtimer();   // first weight stage present
            const int i0 = it * LT_STAGE_TILES + LT_TPW * warp;
            if (i0 < ntiles) {
                const unsigned char* wp = wbase + s * LT_STAGE_BYTES;
                const bool two = i0 + 1 < ntiles;
                const uint4 W0 = *reinterpret_cast<const uint4*>(wp);
                uint4 W1 = make_uint4(0, 0, 0, 0);
                if (two) W1 = *reinterpret_cast<const uint4*>(wp + WF_TILE_BYTES);
                int kb1 = kb0 + 1, rbl1 = rbl0;
                if (kb1 >= KB) { kb1 = 0; ++rbl1; }
                if (rbl0 != cur_rb) { flush(cur_rb); cur_rb = rbl0; }
                tile_math(W0, kb0);
                if (two) {
                    if (rbl1 != cur_rb) { flush(cur_rb); cur_rb = rbl1; }
                    tile_math(W1, kb1);
                }
            }
            kb0 += LT_STAGE_TILES;
            while (kb0 >= KB) { kb0 -= KB; ++rbl0; }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
            if (++s == p.stages) { s = 0; ph_bit ^= 1; }
        }
        flush(cur_rb);
        group_bar<1, LT_CT>();
        if (tr) p.trace[5 + 5 * pi] = gtimer();          // main loop done

        // ---- epilogue: this CTA owns every k of its rows — plain stores
        if (ph.epi == 0) {
            for (int i = tid; i < n_rbs * 32; i += LT_CT) {
                const int rbl = i >> 5, m = (i >> 2) & 7, r = (i & 3) * 4;
                const int n = (rb_lo + rbl) * 16 + r;
                if (m < M && n < ph.N) {
                    const int4 v = *reinterpret_cast<const int4*>(sacc + rbl * 128 + m * 16 + r);
                    float* dst = ph.acc + (size_t)m * ph.acc_ld + n;
                    if (n + 3 < ph.N && (ph.acc_ld & 3) == 0) *reinterpret_cast<float4*>(dst) = make_float4((float)v.x, (float)v.y, (float)v.z, (float)v.w);
                    else {
                        const int vv[4] = {v.x, v.y, v.z, v.w};
                        for (int e = 0; e < 4; ++e) if (n + e < ph.N) dst[e] = (float)vv[e];
                    }
                }
            }
        } else {
            // gate|up pairs (interleaved in 16-row blocks): h = bf16(bf16(relu(g)^2) * u)   (direct_inference_full.py:259)
            const float wg = ph.h_wscale[0], wu = ph.h_wscale[1];
            const int npairs = n_rbs >> 1;
            for (int i = tid; i < npairs * 8 * 8; i += LT_CT) {      // (pair, m, 2 rows)
                const int pr = i >> 6, m = (i >> 3) & 7, r = (i & 7) * 2;
                if (m < M) {
                    const int rs = s_rsum[m];
                    const float sa = s_scale[m];
                    const int2 G = *reinterpret_cast<const int2*>(sacc + (2 * pr) * 128 + m * 16 + r);
                    const int2 U = *reinterpret_cast<const int2*>(sacc + (2 * pr + 1) * 128 + m * 16 + r);
                    float x[2];
                    const int gv[2] = {G.x, G.y}, uv[2] = {U.x, U.y};
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float g = bf16_round(lin_value((float)gv[e], rs, wg, sa));
                        const float u = bf16_round(lin_value((float)uv[e], rs, wu, sa));
                        const float t = fmaxf(g, 0.f);
                        x[e] = bf16_round(bf16_round(t * t) * u);
                    }
                    const int col = ((rb_lo >> 1) + pr) * 16 + r;
                    *reinterpret_cast<uint32_t*>(ph.h_out + (size_t)m * ph.h_ld + col) = pack_bf16x2(x[0], x[1]);
                }
            }
        }
        if (tr) p.trace[6 + 5 * pi] = gtimer();          // epilogue stores issued
    }
    if (p.nph == 1) pdl_launch_dependents();
}

// ---------------------------------------------------------------------------------------------------------------
static size_t lt_fixed_smem(int M, int Kmax, int sacc_rbs, int xs_elems = 0) {
    return (((size_t)M * (Kmax + 16) + 127) & ~(size_t)127) + (((size_t)Kmax * 2 + 127) & ~(size_t)127) + (((size_t)xs_elems * 2 + 127) & ~(size_t)127) + (size_t)sacc_rbs * 512 + 1024;
}

bool layer_tail_fits(int M, int Kmax, int max_rb_per_cta) {
    if (M < 1 || M > 8) return false;
    return lt_fixed_smem(M, Kmax, max_rb_per_cta + 2) + 4 * (size_t)LT_STAGE_BYTES <= 220 * 1024;
}

int launch_layer_tail(const TailArgs& a, cudaStream_t s, bool pdl) {
    if (a.nph < 1 || a.nph > LT_MAX_PHASES || a.M < 1 || a.M > 8) { set_error("layer_tail: bad phase count / batch"); return WF_ERR_INVALID; }
    TailParams p;
    memset(&p, 0, sizeof p);
    p.nph = a.nph; p.M = a.M;
    const int grid = sm_count();
    int Kmax = 0, max_rbs = 0;
    long long extra = 0;   // running rotation: CTAs that received an extra unit in earlier phases come last in later ones
    for (int i = 0; i < a.nph; ++i) {
        const TailPhaseArgs& x = a.ph[i];
        if (x.K % WF_QK || x.g.K != x.K) { set_error("layer_tail: K %% 128 != 0 or glue length != K"); return WF_ERR_INVALID; }
        TailPhase& ph = p.ph[i];
        ph.g = x.g; ph.g.M = a.M;
        ph.w = x.w; ph.N = x.N; ph.K = x.K; ph.KB = x.K / WF_QK; ph.RB = (x.N + 15) / 16;
        ph.rb_unit = x.relu2 ? 2 : 1;
        if (ph.RB % ph.rb_unit) { set_error("layer_tail: gate|up needs an even number of row blocks"); return WF_ERR_INVALID; }
        ph.epi = x.relu2 ? 1 : 0;
        ph.acc = x.acc; ph.acc_ld = x.acc_ld; ph.lin_ascale = x.lin_ascale; ph.lin_rowsum = x.lin_rowsum;
        ph.h_out = x.h_out; ph.h_ld = x.h_ld; ph.h_wscale = x.h_wscale;
        const int units = ph.RB / ph.rb_unit;
        ph.rot = (int)(extra % grid);
        extra += units % grid;
        const int per = ((units + grid - 1) / grid) * ph.rb_unit;
        if (per > max_rbs) max_rbs = per;
        if (x.K > Kmax) Kmax = x.K;
    }
    p.act_stride = Kmax + 16;
    p.ys_elems = Kmax;
    p.sacc_rbs = max_rbs + 2;
    // flag-in-data form: batch 1, the whole tail, every row in the register form of the glue (K <= 9216: the 2B shape).  Extending it to
    // the 8B / 70B ffn rows through the scratch form of the glue was measured: 1.95 -> 1.92 ms (8B), 9.25 -> 9.59 ms (70B), and the extra
    // code cost the 2B shape 8 % (profiles/r02v_summary.txt) — not kept.  Neither was an instantiation whose register glue holds the 8B ffn
    // row in 5 granules per thread: 316 bytes of spills, 1.95 -> 2.69 ms (profiles/r02x_summary.txt).
    const bool lam_ok = a.lam && a.M == 1 && a.nph >= 3 && Kmax <= LT_CT * 8 * 3;
    p.lam = lam_ok ? 1 : 0;
    p.xs_elems = lam_ok ? a.ph[1].K : 0;
    const size_t fixed = lt_fixed_smem(a.M, Kmax, p.sacc_rbs, p.xs_elems);
    int stages = (int)((220 * 1024 - fixed) / LT_STAGE_BYTES);
    if (stages > LT_MAX_STAGES) stages = LT_MAX_STAGES;
    if (stages < 3) { set_error("layer_tail: M*K=%d*%d does not fit in shared memory", a.M, Kmax); return WF_ERR_UNSUPPORTED; }
    p.stages = stages;
    p.bar = a.bar; p.epoch = a.epoch; p.bump_epoch = a.bump_epoch; p.err = a.err; p.trace = a.trace;
    if (a.bump_epoch && a.nph != 1) { set_error("layer_tail: the epoch-bumping launch must be single-phase"); return WF_ERR_INVALID; }
    const size_t smem = (size_t)stages * LT_STAGE_BYTES + fixed;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(LT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    if (a.M == 1 && p.lam) return launch_check(cudaLaunchKernelEx(&cfg, layer_tail_kernel<1, true>, p), "layer_tail");
    if (a.M == 1) return launch_check(cudaLaunchKernelEx(&cfg, layer_tail_kernel<1, false>, p), "layer_tail");
    return launch_check(cudaLaunchKernelEx(&cfg, layer_tail_kernel<0, false>, p), "layer_tail");
}

void layer_tail_init_attrs() {
    cudaFuncSetAttribute(layer_tail_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(layer_tail_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(layer_tail_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
}

}  // namespace wf
