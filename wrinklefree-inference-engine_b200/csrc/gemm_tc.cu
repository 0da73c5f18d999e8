// Prefill / large-batch ternary GEMM on the 5th-generation tensor cores (tcgen05, kind::i8) — north-star item 4.
//
//   S[m,n] = sum_k code[n,k] * a[m,k],  code = w+1 in {0,1,2} (u8), a int8, exact int32 accumulation
//   (what bitnet_gemm computes row by row with ggml_vec_dot_i2_i8_s, extern/BitNet/src/ggml-bitnet-mad.cpp:94-186;
//    signature src/wrinklefree_inference/kernels/bitnet_patch.py:73).  Used where the shape is a real dense
//   contraction (M >= 128 tokens: prefill, 8 x 2048 tokens => 6.8e13 int8 MACs for the 2B shape).
//
// B200 mapping (one persistent CTA per SM, 4 + 4*NU + 2 warps, warp-specialised):
//   * D tile = 128 weight rows (UMMA M = TMEM lanes) x 128 tokens (UMMA N), int32, double-buffered in TMEM
//     (2 x 128 columns) so the epilogue of tile i overlaps the main loop of tile i+1;
//   * the WEIGHT operand never touches shared memory: NU groups of four "unpack" warps (one per TMEM lane quarter)
//     read their row's 32 packed bytes per 128-k block straight from global/L2 into registers (prefetched 4 blocks
//     ahead), expand 2-bit -> u8 with shift+mask and write the 128 codes into a TMEM operand ring
//     (tcgen05.st 32x32b.x32, 8 stages x 32 columns).  tcgen05.mma reads A from TMEM ("TS" form), so shared
//     memory only carries the activation operand (16 KB written by TMA + 16 KB read by the MMAs per stage);
//   * int8 activations [M,K] are the B operand: TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into an
//     8-stage x 16 KB ring; one mbarrier per stage collects the TMA transaction bytes AND the four unpack
//     warps' arrivals, one tcgen05.commit per stage releases both halves;
//   * one elected thread issues 4 x tcgen05.mma.cta_group::1.kind::i8 (M128 N128 K32) per stage;
//   * epilogue warps: tcgen05.ld 32x32b.x32 (lane = weight row, columns = tokens) -> the linear epilogue (Sigma-a
//     correction, scales, residual, relu^2*up) in registers -> one store per token: a warp's 32 lanes are 32 consecutive n,
//     so every store instruction is a full coalesced line and nothing is transposed through shared memory.
// Shared memory is the binding resource of the main loop: TMA write + MMA read of the activation tile = 128 B/clk at the
// 256-clk/stage tensor floor, i.e. all of an SM's shared-memory bandwidth.  Round 2 measurements (profiles/r02_prefill_*):
// with the epilogue's math and stores switched off the kernel runs 2.15 PTOP/s (374 clk per stage) — the round-1 epilogue
// (smem transpose + dependent per-token global loads) cost 40 % on the short-K shapes.  The cta_group::2 form (each SM
// holds half of the activation tile) is what would lift the 374-clk stage.
#include "linear_common.cuh"
#include "fused_common.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

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
constexpr int TC_L2_AHEAD = 6;                     // decode form: L2 prefetch distance of the weight stream, in items of an unpack group
constexpr int TC_EPW = 8;                          // epilogue warps: two per TMEM lane quarter, 64 of the tile's 128 tokens each
constexpr int TC_EP_TOKS = TC_BN / (TC_EPW / 4);   // tokens per epilogue warp
constexpr int TC_EP_TOK = TC_EPW * TC_EP_TOKS * 8;  // per epilogue warp: {(float) Sigma a, activation scale} of its tokens
constexpr int tc_threads(int nu) { return (TC_EPW + 4 * nu + 2) * 32; }   // warps 0..7 epilogue, then 4*NU unpack (NU groups of 4), TMA, MMA
constexpr int TC_SMEM = TC_STAGES * TC_STAGE_BYTES + TC_EP_TOK + 1024;

struct TcParams {
    const uint8_t* w;
    int w_layout;
    int M, N, K, KB;
    int tiles_n, tiles_m;
    int dbg;   // timing experiments only (WF_TC_DBG, bit mask): 1 = epilogue skips math/stores, 2 = also the TMEM loads, 4 = unpack warps skip the TMEM stores, 8 = no activation TMA
    // batched-decode form (launch_gemm_tc_red): the K dimension is cut into `splitk` equal runs of KBs k-blocks, one CTA
    // tile each, and the epilogue adds the raw accumulators into a global fp32 buffer (exact: integers < 2^24) instead
    // of storing a finished output — the consumer phase of the decode step applies scales / residual / norm.
    int splitk, KBs;
    int l2_ahead;   // decode form: L2 prefetch distance of the weight stream in items of an unpack group (0 = off)
    float* red_acc;
    int red_ld;
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

// CL = thread-block-cluster size: the CL CTAs of a cluster work on the SAME 128 tokens and CL consecutive 128-row
// weight tiles; each CTA TMA-loads 128/CL tokens of the activation stage and multicasts them to all CL CTAs, so the
// L2 -> SM activation traffic (the limiter of this kernel: ~6300 B/clk chip-wide) drops by CL.
// NU = unpack warp groups (4 warps each, one per TMEM lane quarter); group g expands the stages with item % NU == g.
// One warp issues ~190 instructions per stage at IPC ~0.3, so a single group paces the kernel at ~850 clk per stage
// (measured) against the 256-clk tensor floor; three groups bring the unpack under the floor.
template <int CL, int NU>
__global__ void __launch_bounds__(tc_threads(NU), 1) ternary_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const TcParams p) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[TC_STAGES];    // TMA bytes + 4 unpack-warp arrivals
    __shared__ __align__(8) uint64_t empty_bar[TC_STAGES];   // tcgen05.commit: MMAs of the stage retired
    __shared__ __align__(8) uint64_t tfull_bar[2];           // accumulator buffer complete
    __shared__ __align__(8) uint64_t tempty_bar[2];          // accumulator buffer drained by the 4 epilogue warps
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;   // 128B-swizzle atoms need 1024-B alignment
    const int crank = (CL > 1) ? (int)(blockIdx.x % CL) : 0;          // 1-D cluster: rank == blockIdx.x % CL
    const int cluster_id = blockIdx.x / CL, n_clusters = gridDim.x / CL;
    const int ngroups = (p.tiles_n + CL - 1) / CL;                    // groups of CL weight-row tiles
    const int total_tiles = ngroups * p.tiles_m * p.splitk;           // cluster tiles (x K-splits in the decode form)
    const int my_tiles = (total_tiles - cluster_id + n_clusters - 1) / n_clusters;
    const int KBs = p.KBs;                                            // k-blocks per tile (== KB unless split)
    const int n_items = my_tiles * KBs;
    auto tile_coords = [&](int tl, int& m_blk, int& n_blk, int& ks) {
        const int ct0 = cluster_id + tl * n_clusters;
        ks = ct0 % p.splitk;                                          // splits of one tile run side by side
        const int ct = ct0 / p.splitk;
        m_blk = ct / ngroups;
        n_blk = (ct - m_blk * ngroups) * CL + crank;                  // may be >= tiles_n (padding tile: epilogue masks it)
    };
    pdl_launch_dependents();   // no-op unless launched with programmatic stream serialisation (decode step)

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { mbar_init(&full_bar[s], 5); mbar_init(&empty_bar[s], CL); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], TC_EPW); }
        fence_mbar_init();
    }
    constexpr int W_TMA = TC_EPW + 4 * NU, W_MMA = W_TMA + 1;
    if (warp == W_MMA) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(TC_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (CL > 1) {   // peers must see initialised barriers before the first multicast / remote commit
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;

    if (warp == W_TMA) {
        // ------------------------------------------------------------ TMA producer: int8 activations
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_q) : "memory");
            pdl_wait_prior();   // the activations are the previous kernel's output (the weights are not: unpack warps run ahead)
            int it = 0;
            constexpr int SLICE = TC_BN / CL;   // tokens this CTA fetches for the whole cluster
            for (int tl = 0; tl < my_tiles; ++tl) {
                int m_blk, n_blk, ks;
                tile_coords(tl, m_blk, n_blk, ks);
                for (int kb = 0; kb < KBs; ++kb, ++it) {
                    const int kg = ks * KBs + kb;
                    const int s = it % TC_STAGES;
                    mbar_wait(&empty_bar[s], (uint32_t)(((it / TC_STAGES) & 1) ^ 1));   // all CL CTAs retired this slot
                    if (CL == 1 && (p.dbg & 8)) { mbar_arrive(&full_bar[s]); continue; }  // timing experiment: no activation traffic
                    mbar_expect_tx(&full_bar[s], TC_STAGE_BYTES);                        // own barrier: all CL slices land here
                    const uint32_t dst = ring + s * TC_STAGE_BYTES + crank * SLICE * TC_BK;
                    if (CL == 1)
                        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                                     ::"r"(dst), "l"(&tmap_q), "r"(kg * TC_BK), "r"(m_blk * TC_BN), "r"(smem_u32(&full_bar[s])) : "memory");
                    else
                        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%2, %3}], [%4], %5;"
                                     ::"r"(dst), "l"(&tmap_q), "r"(kg * TC_BK), "r"(m_blk * TC_BN + crank * SLICE), "r"(smem_u32(&full_bar[s])),
                                       "h"((uint16_t)((1u << CL) - 1)) : "memory");
                }
            }
        }
    } else if (warp == W_MMA) {
        // ------------------------------------------------------------ MMA issuer (whole warp walks the loop, lane 0 issues)
        // The issue loop IS the pace of this kernel when it is long: ncu (profiles/r02_full_gemm_tc_noepi) showed ~75 dependent
        // SASS instructions per stage (R2UR moves, address arithmetic, one ELECT waterfall per tcgen05 instruction) = ~500 clk
        // against the 256-clk tensor time of a stage, with every other role waiting on this warp.  So the ring position is a
        // compile-time constant here (the loop walks items in groups of TC_STAGES): operand descriptors and barrier addresses
        // are base + immediate, and the only per-stage arithmetic left is the tile bookkeeping.
        constexpr uint32_t idesc = tc_idesc(TC_BM, TC_BN);
        const uint64_t bdesc0 = tc_smem_desc_sw128(ring);
        const uint32_t a_tmem0 = __shfl_sync(0xffffffffu, tmem, 0) + TC_A_COL0;
        const uint32_t full0 = smem_u32(&full_bar[0]), empty0 = smem_u32(&empty_bar[0]);
        const uint32_t tfull0 = smem_u32(&tfull_bar[0]);
        // One ELECTED thread runs the whole loop (elect.sync, not `lane == 0`: ptxas then knows the region is single-threaded
        // and issues the tcgen05 instructions straight from uniform registers instead of one ELECT waterfall loop each).
        uint32_t elected;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(elected));
        if (elected) {
            uint32_t ph = 0;
            int kb = 0, tl = 0;            // position inside the current tile, tile index
            uint32_t d_tmem = a_tmem0 - TC_A_COL0;
            for (int i0 = 0; i0 < n_items; i0 += TC_STAGES) {
#pragma unroll
                for (int s = 0; s < TC_STAGES; ++s) {
                    if (i0 + s < n_items) {
                        if (kb == 0) {   // a new tile: its accumulator buffer must have been drained by the epilogue warps
                            const int buf = tl & 1;
                            mbar_wait(&tempty_bar[buf], (uint32_t)(((tl >> 1) & 1) ^ 1));
                            d_tmem = a_tmem0 - TC_A_COL0 + buf * TC_BN;
                        }
                        {   // wait for the TMA bytes and the four unpack warps of this stage
                            uint32_t ok;
                            do {
                                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                                             : "=r"(ok) : "r"(full0 + s * 8), "r"(ph) : "memory");
                            } while (!ok);
                        }
                        tc_fence_after();
                        const uint64_t bd = bdesc0 + (uint64_t)(s * (TC_STAGE_BYTES >> 4));
                        const uint32_t at = a_tmem0 + s * TC_A_COLS;
                        tc_mma_i8_ts(d_tmem, at, bd, idesc, kb != 0 ? 1u : 0u);
                        tc_mma_i8_ts(d_tmem, at + 8, bd + 2, idesc, 1u);
                        tc_mma_i8_ts(d_tmem, at + 16, bd + 4, idesc, 1u);
                        tc_mma_i8_ts(d_tmem, at + 24, bd + 6, idesc, 1u);
                        if (CL == 1)
                            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(empty0 + s * 8) : "memory");
                        else   // the slot is refilled by every CTA of the cluster: tell all of them
                            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                                         ::"r"(empty0 + s * 8), "h"((uint16_t)((1u << CL) - 1)) : "memory");
                        if (++kb == KBs) {
                            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tfull0 + (tl & 1) * 8) : "memory");
                            kb = 0; ++tl;
                        }
                    }
                }
                ph ^= 1u;
            }
        }
        __syncwarp();
    } else if (warp >= TC_EPW) {
        // ------------------------------------------------------------ unpack warps: 2-bit weights -> u8 codes -> TMEM
        // The SM is ISSUE-bound in this kernel (unpack + epilogue + issue ~ 1100 warp instructions per 256-clk stage against
        // 4 issue slots per clk), so the per-item path is kept to the loads, the 7 x 8 shift/mask instructions, one tcgen05.st
        // and pointer / ring-slot increments: no multiplications, no divisions, no per-item tile arithmetic.
        const int q = warp & 3, grp = (warp - TC_EPW) >> 2;
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        const int K4 = p.K / 4;
        uint32_t pre[TC_PF][8];
        const bool i2s = p.w_layout == WF_W_I2S;
        const uint32_t kb_stride = i2s ? 32u : (uint32_t)WF_TILE_BYTES;
        // ---- load cursor: (tile, k-block, pointer) of the next item this group fetches; items of a group are NU apart
        int l_tl = 0, l_kb = grp;
        const uint8_t* l_src = p.w;
        bool l_valid = true;
        // Decode form on large matrices (70B shapes: 117 MB per linear, nothing L2-resident): the registers hold 4 items = 32 KB in flight
        // per SM, half of what HBM latency x bandwidth asks for (Little: ~44 KB), so every thread also pulls its 64-byte window of the
        // k-block TC_L2_AHEAD items ahead into L2 (rows r and r + 8 share a window: one sector each).
        const int l2_ahead = p.l2_ahead;
        uint32_t pf_off = 0;
        auto seek = [&]() {   // off the per-item path: tile boundary (or start)
            while (l_kb >= KBs) { l_kb -= KBs; ++l_tl; }
            l_valid = l_tl < my_tiles;
            if (!l_valid) return;
            int m_blk, n_blk, ks;
            tile_coords(l_tl, m_blk, n_blk, ks);
            int row = n_blk * TC_BM + q * 32 + lane;
            if (row >= p.N) row = p.N - 1;   // padded rows compute garbage that the epilogue masks
            // b200_tile_v1: word (16*(r&7) + 4*(j&3) + 2*(j>>2) + (r>>3)) of tile (row/16, kb) holds packed bytes [4j,4j+4) of row r
            const uint8_t* rowbase = i2s ? p.w + (size_t)row * K4
                                         : p.w + (size_t)(row >> 4) * p.KB * WF_TILE_BYTES + (size_t)(16 * (row & 7) + ((row >> 3) & 1)) * 4;
            l_src = rowbase + ((size_t)ks * KBs + l_kb) * kb_stride;   // this tile's K run, this group's first k-block in it
            pf_off = i2s ? 0u : (uint32_t)(((row >> 3) & 1) * 28);
        };
        seek();
        auto load_item = [&](uint32_t (&r)[8]) {
            if (!l_valid) return;
            if (i2s) {
                const uint4 a = __ldg(reinterpret_cast<const uint4*>(l_src)), b = __ldg(reinterpret_cast<const uint4*>(l_src) + 1);
                r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) r[j] = __ldg(reinterpret_cast<const uint32_t*>(l_src) + 4 * (j & 3) + 2 * (j >> 2));
            }
            if (l2_ahead && l_kb + l2_ahead * NU < KBs)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(l_src + (size_t)l2_ahead * NU * kb_stride + pf_off));
            l_kb += NU;
            l_src += NU * kb_stride;
            if (l_kb >= KBs) seek();
        };
#pragma unroll
        for (int u = 0; u < TC_PF; ++u) load_item(pre[u]);
        // ---- ring cursor of the item being written: slot = item % TC_STAGES, parity of its use count
        uint32_t slot = (uint32_t)grp, par = 1u;       // first pass over the ring waits on parity 1 (fresh barriers: returns at once)
        const uint32_t full0 = smem_u32(&full_bar[0]), empty0 = smem_u32(&empty_bar[0]);
        const uint32_t tbase = tmem + lane_base + TC_A_COL0;
        uint32_t pending = 0xffffffffu;                 // full-barrier address of the stage whose tcgen05.st is still in flight
        auto retire = [&]() {
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(pending) : "memory");
        };
        for (int item0 = grp; item0 < n_items; item0 += TC_PF * NU) {
#pragma unroll
            for (int u = 0; u < TC_PF; ++u) {
                if (item0 + u * NU < n_items) {
                    uint32_t v[32];
#pragma unroll
                    for (int g = 0; g < 4; ++g)
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[8 * g + j] = (pre[u][j] >> (6 - 2 * g)) & 0x03030303u;   // k = 32g + 4j .. +3
                    load_item(pre[u]);
                    // retire the PREVIOUS stage's tcgen05.st only now: its latency overlapped the unpack above
                    if (pending != 0xffffffffu) retire();
                    {
                        uint32_t ok;
                        do {
                            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                                         : "=r"(ok) : "r"(empty0 + slot * 8), "r"(par) : "memory");
                        } while (!ok);
                    }
                    tc_fence_after();
                    if (!(p.dbg & 4)) TC_ST32(tbase + slot * TC_A_COLS, v);              // (dbg 4: timing experiment without the TMEM stores)
                    pending = full0 + slot * 8;
                    slot += NU;
                    if (slot >= TC_STAGES) { slot -= TC_STAGES; par ^= 1u; }
                }
            }
        }
        if (pending != 0xffffffffu) retire();
    } else {
        // ------------------------------------------------------------ epilogue warps: TMEM -> registers -> global
        // A thread owns one weight row n (TMEM lane) and receives 32 tokens per tcgen05.ld.  The output is token-major, so a
        // warp's store of ONE token covers 32 consecutive n: a full 128-byte line in fp32 (64 B in bf16) — coalesced as it
        // stands, no transpose through shared memory.  Two warps share a lane quarter and take 64 tokens each: with four warps
        // and ~34 issued instructions per token the epilogue paced the short-K shapes (ncu r02_full_gemm_tc: the four warps
        // never idle, the MMA warp 27 % of its time waiting for a drained accumulator).  gate|up tiles (relu^2*up): the 16 up
        // rows of a 32-row group sit in lanes 16..31 and are handed to the gate lanes with one shuffle.
        const int q = warp & 3, th = warp >> 2;                                    // lane quarter, token half
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        // per-token {(float) Sigma a, activation scale} of this warp's 64 tokens, staged BEFORE the accumulator wait (dependent
        // global loads inside the store loop were ~10 k clk per tile); read back as one broadcast LDS.64 per token
        float2* tok = reinterpret_cast<float2*>(smem_raw + (ring - smem_u32(smem_raw)) + TC_STAGES * TC_STAGE_BYTES) + warp * TC_EP_TOKS;
        const int mode = p.ep.out_mode;
        const bool relu2 = (mode == WF_OUT_LINEAR) && (p.ep.epi_flags & WF_EPI_RELU2_MUL);
        const bool resid = (mode == WF_OUT_LINEAR) && (p.ep.epi_flags & WF_EPI_RESIDUAL);
        const bool obf16 = (mode == WF_OUT_LINEAR) && p.ep.out_dtype == WF_BF16;
        const int ncols = relu2 ? p.N / 2 : p.N;
        const bool red = p.red_acc != nullptr;
        const bool fast_mode = !red && mode == WF_OUT_LINEAR && !relu2 && !resid && obf16 && !p.dbg;
        const bool fold_ws = fast_mode && p.ep.nseg <= 1;
        const float ws_all = fold_ws ? p.ep.w_scale[0] : 1.f;
        if (red) pdl_wait_prior();   // the accumulator buffer was cleared by an earlier phase of the step
        for (int tl = 0; tl < my_tiles; ++tl) {
            int m_blk, n_blk, ks;
            tile_coords(tl, m_blk, n_blk, ks);
            const int buf = tl & 1;
            const int mw0 = m_blk * TC_BN + th * TC_EP_TOKS;                       // first token of this warp
            if (!red && mode != WF_OUT_I32_PLUS1 && mode != WF_OUT_COMPAT_F32) {
                __syncwarp();                                                       // the previous tile's readers are done
#pragma unroll
                for (int j = 0; j < TC_EP_TOKS / 32; ++j) {
                    const int m = min(mw0 + j * 32 + lane, p.M - 1);
                    const int rs = p.ep.rowsum ? p.ep.rowsum[m] : 0;
                    const float sa = (mode == WF_OUT_LINEAR) ? p.ep.act_scale[p.ep.act_scale_per_token ? m : 0] : 1.f;
                    // |Sigma a| <= 127 K < 2^24: exact.  One weight scale for the whole matrix: fl(w_s * Sigma a) is staged here, once
                    // per token, instead of being recomputed by every row (same product, same rounding)
                    tok[j * 32 + lane] = make_float2(fold_ws ? __fmul_rn(ws_all, (float)rs) : (float)rs, sa);
                }
                __syncwarp();
            }
            const int n = n_blk * TC_BM + q * 32 + lane;                           // this thread's weight row
            const int col = relu2 ? (n >> 5) * 16 + (n & 15) : n;                  // its output column (gate lanes under relu^2*up)
            const bool store_lane = relu2 ? (lane < 16 && col < ncols) : (n < p.N);
            float ws = 1.f;
            if (mode == WF_OUT_LINEAR) ws = relu2 ? p.ep.w_scale[(lane >= 16 && p.ep.nseg > 1) ? 1 : 0] : seg_scale(p.ep, min(n, p.N - 1));
            const bool fast = fast_mode && mw0 + TC_EP_TOKS <= p.M && n_blk * TC_BM + q * 32 + 32 <= p.N;   // warp-uniform
            mbar_wait(&tfull_bar[buf], (uint32_t)((tl >> 1) & 1));
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < TC_EP_TOKS / 32; ++c) {
                uint32_t v[32];
                if (!(p.dbg & 2)) {
                    TC_LD32(tmem + lane_base + buf * TC_BN + th * TC_EP_TOKS + c * 32, v);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                }
                if (c == TC_EP_TOKS / 32 - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[buf]);
                }
                if (p.dbg) continue;
                const int m0 = mw0 + c * 32;
                if (fast) {
                    // full tile, bf16 linear output: ~9 issued instructions per token (LDS.64, I2F, 3 FMUL, FADD, F2F, pointer add, STG);
                    // the reference's rounding sequence (linear_value in linear_common.cuh)
                    __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(p.ep.out) + (size_t)m0 * p.ep.ldo + col;
                    const size_t ldo = (size_t)p.ep.ldo;
                    const float2* tk = tok + c * 32;
#pragma unroll
                    if (fold_ws) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float2 t = tk[j];
                            *op = __float2bfloat16_rn(__fmul_rn(__fsub_rn(__fmul_rn((float)(int)v[j], ws), t.x), t.y));
                            op += ldo;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float2 t = tk[j];
                            *op = __float2bfloat16_rn(__fmul_rn(__fsub_rn(__fmul_rn((float)(int)v[j], ws), __fmul_rn(ws, t.x)), t.y));
                            op += ldo;
                        }
                    }
                    continue;
                }
                const int mcnt = min(32, p.M - m0);                                // tokens of this chunk that exist (warp-uniform)
                if (red) {
                    // decode form: exact fp32 RED of the raw accumulators (split-K partial sums meet in L2)
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < mcnt && store_lane) red_add_f32(p.red_acc + (size_t)(m0 + j) * p.red_ld + col, (float)(int)v[j]);
                    continue;
                }
                if (mode == WF_OUT_LINEAR) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (j < mcnt) {
                            const float2 t = tok[c * 32 + j];
                            // the reference's rounding sequence (linear_value in linear_common.cuh)
                            float y = __fmul_rn(__fsub_rn(__fmul_rn((float)(int)v[j], ws), __fmul_rn(ws, t.x)), t.y);
                            const size_t o = (size_t)(m0 + j) * p.ep.ldo + col;
                            if (relu2) {
                                const float g = bf16_round(y);
                                const float u = __shfl_down_sync(0xffffffffu, g, 16);   // the up row of this column: lane + 16
                                const float r = fmaxf(g, 0.f);
                                y = bf16_round(bf16_round(r * r) * u);
                            } else if (resid) {
                                if (store_lane) y = __bfloat162float(p.ep.residual[o]) + bf16_round(y);
                            }
                            if (store_lane) {
                                if (obf16) reinterpret_cast<__nv_bfloat16*>(p.ep.out)[o] = __float2bfloat16_rn(y);
                                else reinterpret_cast<float*>(p.ep.out)[o] = y;
                            }
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (j < mcnt && store_lane) {
                            const int S = (int)v[j];
                            uint32_t w;
                            if (mode == WF_OUT_I32_PLUS1) w = (uint32_t)S;
                            else if (mode == WF_OUT_I32) w = (uint32_t)(S - (int)tok[c * 32 + j].x);
                            else w = __float_as_uint(__fmul_rn((float)S, p.ep.host_scale));   // bitnet_gemv/gemm: (float)sumi * scale
                            reinterpret_cast<uint32_t*>(p.ep.out)[(size_t)(m0 + j) * p.ep.ldo + col] = w;
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CL > 1) {   // nobody leaves while a peer may still multicast into this CTA's ring or signal its barriers
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    if (warp == W_MMA) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TC_TMEM_COLS) : "memory");
    }
}

// ---- host side --------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static DevOnce g_tc_attr;
static int g_tc_cluster = 0, g_tc_nu = 0;

void gemm_tc_init_attrs() {
    if (g_tc_attr.need()) {
        bool ok = cudaFuncSetAttribute(ternary_gemm_tc_kernel<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) == cudaSuccess;
        ok = ok && cudaFuncSetAttribute(ternary_gemm_tc_kernel<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) == cudaSuccess;
        ok = ok && cudaFuncSetAttribute(ternary_gemm_tc_kernel<1, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) == cudaSuccess;
        ok = ok && cudaFuncSetAttribute(ternary_gemm_tc_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) == cudaSuccess;
        ok = ok && cudaFuncSetAttribute(ternary_gemm_tc_kernel<2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) == cudaSuccess;
        ok = ok && cudaFuncSetAttribute(ternary_gemm_tc_kernel<4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) == cudaSuccess;
        (void)ok;
    }
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            g_encode = (EncodeTiledFn)fn;
    }
    if (!g_tc_cluster) {
        const char* e = getenv("WF_TC_CLUSTER");
        // pairs of CTAs share one multicast activation tile: 2B prefill shape 2312 -> 2331 TOP/s, 8B 2146 -> 2221 (call r02ah); 4 is slower
        g_tc_cluster = e ? atoi(e) : 2;
        if (g_tc_cluster != 1 && g_tc_cluster != 2 && g_tc_cluster != 4) g_tc_cluster = 2;
        const char* u = getenv("WF_TC_NU");
        g_tc_nu = u ? atoi(u) : 2;     // two unpack groups: 18 warps at 112 registers (three groups = 22 warps at 88 spill)
        if (g_tc_nu < 1 || g_tc_nu > 3) g_tc_nu = 2;
    }
}

template <int CL, int NU>
static int launch_tc_cl(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s) {
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)a->K, (cuuint64_t)a->M};
    const cuuint64_t gstride[1] = {(cuuint64_t)a->ldq};
    const cuuint32_t box[2] = {TC_BK, TC_BN / CL};
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
    p.splitk = 1; p.KBs = p.KB; p.red_acc = nullptr; p.red_ld = 0; p.l2_ahead = 0;
    { const char* d = getenv("WF_TC_DBG"); p.dbg = d ? atoi(d) : 0; }
    const int total = ((p.tiles_n + CL - 1) / CL) * p.tiles_m;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(tc_threads(NU));
    cfg.dynamicSmemBytes = TC_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = CL > 1 ? 1 : 0;
    static int max_clusters = 0;   // per CL instantiation
    if (!max_clusters) {
        max_clusters = sm_count() / CL;
        if (CL > 1) {
            cfg.gridDim = dim3(sm_count() / CL * CL);
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, ternary_gemm_tc_kernel<CL, NU>, &cfg) == cudaSuccess && n > 0) max_clusters = n;
            else cudaGetLastError();
        }
    }
    const int clusters = total < max_clusters ? total : max_clusters;
    cfg.gridDim = dim3(clusters * CL);
    return launch_check(cudaLaunchKernelEx(&cfg, ternary_gemm_tc_kernel<CL, NU>, tmap, p), "ternary_gemm_tc");
}

// Batched-decode form (the engine's phase kernels, 32 < M <= 128): S[m][n] += sum over this launch's K columns, added into the
// global fp32 accumulator `acc` ([M][acc_ld], zero on entry of the step phase) exactly like gemm_oneshot does — but on the
// tcgen05 int8 pipe (mma.sync int8 is a slow legacy path on sm_100: the M = 64 linears were IMMA-bound).  The (rows x K)
// plane is cut into 128-row x KBs-k-block tiles so that ~one tile lands on every SM; split-K partial sums meet in L2.
int launch_gemm_tc_red(const uint8_t* w_tile, int N, int K, const int8_t* q, int M, int ldq, int k_off, float* acc, int acc_ld,
                       cudaStream_t s, bool pdl) {
    gemm_tc_init_attrs();
    if (!g_encode) { set_error("gemm_tc: tcgen05 path unavailable (driver entry point / smem attribute)"); return WF_ERR_CUDA; }
    if (M < 1 || M > TC_BN || K % WF_QK || k_off % WF_QK || ldq % 16 || acc_ld % 4 || N % 8) { set_error("gemm_tc_red: unsupported shape"); return WF_ERR_UNSUPPORTED; }
    constexpr int NU = 2;
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)M};
    const cuuint64_t gstride[1] = {(cuuint64_t)ldq};
    const cuuint32_t box[2] = {TC_BK, TC_BN};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = g_encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void*)(q + k_off), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { set_error("gemm_tc_red: cuTensorMapEncodeTiled failed (%d)", (int)cr); return WF_ERR_CUDA; }
    TcParams p;
    memset(&p, 0, sizeof p);
    p.w = w_tile; p.w_layout = WF_W_TILE;
    p.M = M; p.N = N; p.K = K; p.KB = K / WF_QK;
    p.tiles_n = (N + TC_BM - 1) / TC_BM;
    p.tiles_m = 1;
    p.ep.M = M; p.ep.N = N; p.ep.out_mode = WF_OUT_I32_PLUS1; p.ep.nseg = 1; p.ep.ldo = acc_ld; p.ep.out = acc;
    p.red_acc = acc; p.red_ld = acc_ld;
    // largest divisor of KB that still gives every SM at most one tile
    const int sms = sm_count();
    int split = 1;
    for (int d = 1; d <= p.KB; ++d)
        if (p.KB % d == 0 && p.tiles_n * d <= sms) split = d;
    p.splitk = split; p.KBs = p.KB / split;
    { static const int ahead = getenv("WF_TC_L2_AHEAD") ? atoi(getenv("WF_TC_L2_AHEAD")) : TC_L2_AHEAD; p.l2_ahead = ahead; }
    const int total = p.tiles_n * split;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(total < sms ? total : sms);
    cfg.blockDim = dim3(tc_threads(NU));
    cfg.dynamicSmemBytes = TC_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, ternary_gemm_tc_kernel<1, NU>, tmap, p), "ternary_gemm_tc(decode)");
}

int launch_gemm_tc(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s) {
    gemm_tc_init_attrs();
    if (!g_encode) { set_error("gemm_tc: tcgen05 path unavailable (driver entry point / smem attribute)"); return WF_ERR_CUDA; }
    if (g_tc_cluster == 4) return launch_tc_cl<4, 2>(a, ep, s);
    if (g_tc_cluster == 2) return g_tc_nu == 2 ? launch_tc_cl<2, 2>(a, ep, s) : launch_tc_cl<2, 3>(a, ep, s);
    if (g_tc_nu == 1) return launch_tc_cl<1, 1>(a, ep, s);
    return g_tc_nu == 2 ? launch_tc_cl<1, 2>(a, ep, s) : launch_tc_cl<1, 3>(a, ep, s);
}

}  // namespace wf
