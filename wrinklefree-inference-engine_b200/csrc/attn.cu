// Paged KV cache + split-KV flash-decode attention (north-star item 5).
//
// Replaces the reference's KV utilities:
//   * KVCache.update / .get + attention_with_kv_cache  (src/wrinklefree_inference/kv_cache/kv_cache.py:230-369)
//   * the GQA decode attention of scripts/direct_inference_full.py:210-244 (RoPE, cache write, SDPA)
//   * sgl_kernel.kvcache.KVCacheManager scatter/gather  (tests/test_kv_cache_cpp.py:75-127)
//
// Pool layout: [page][layer][K|V][kv_head][slot][D] bf16 — the (page, layer, K|V, head) block is a
// contiguous page_size*D*2-byte run, so a 16-token block of one head is exactly 4 KB of K and 4 KB of V.
//
// Decode attention is HBM-bound (2*Hkv*D*2 bytes per token per layer, every step).  Design:
//   grid = (kv chunks, kv heads, batch); 4 warps per CTA, 2 CTAs per SM; every WARP is an independent
//   flash-decode worker over a strided set of 16-token blocks with its own 3-stage cp.async ring
//   (8 KB per stage, XOR-swizzled rows => conflict-free ldmatrix), so there is no block-level barrier in
//   the streaming loop and ~100 KB of loads stay in flight per SM.  The G = H/Hkv query heads of a KV
//   head form the (padded) M=16 rows of mma.m16n8k16 bf16: S = Q K^T (fp32), online softmax in fp32 with
//   exp2, O += P V.  Warps are merged through shared memory; chunks are merged by a second tiny kernel.
#include <cuda_fp16.h>

#include "fused_common.cuh"

namespace wf {

constexpr int AT_WARPS = 4;
constexpr int AT_EVEN_MAXB = 128;   // even work split (fused decode form): the per-sequence block counts are scanned by one CTA-wide pass
constexpr int AT_STAGES = 3;
constexpr int AT_BLK = 16;                              // tokens per stage
constexpr int AT_D = 128;                               // head_dim (all BASELINE shapes)
constexpr int AT_ROW = AT_D * 2;                        // bytes per token row
constexpr int AT_STAGE_BYTES = 2 * AT_BLK * AT_ROW;     // K + V
constexpr int AT_SMEM = AT_WARPS * AT_STAGES * AT_STAGE_BYTES;   // 96 KB

// ---- int8 KV option (the reference's KVCacheDtype.INT8, kv_cache/kv_cache.py:198-228): K and V are int8
// with one fp32 scale per (token, kv head).  Pool = int8 region [page][layer][K|V][head][slot][D] followed by the scale
// region [page][layer][K|V][head][slot].  The streaming loop moves half the bytes: int8 rows go through the same cp.async
// ring (6 stages of 4 KB + 128 B of scales), ldmatrix reads them as b16 pairs and ONE PRMT per f16x2 operand register turns
// two codes into f16: codes are stored offset-binary (u = code + 128) and 0x6400 | u is the f16 value 1024 + u = code + 1152,
// exactly.  The constant 1152 is not subtracted per element: S = sum q (code + 1152) - 1152 sum q and
// O = sum p (code + 1152) - 1152 sum p, i.e. one correction per score / per output row (fp32 accumulators: the offset costs
// ~3 of 24 bits).  The MMAs are f16 instead of bf16.  The K scale multiplies the score column, the V scale the probability
// (times 2^8 to stay clear of f16 subnormals; undone exactly at the end).
// The d index inside a 16-wide granule is permuted identically for Q and K (k = 4t.. instead of 2t.., 2t+8..).
constexpr int AT8_ROW = AT_D;
constexpr int AT8_STAGES = 4;                                    // 67.6 KB per CTA: three CTAs (12 warps) per SM
constexpr int AT8_SCALE_OFF = 2 * AT_BLK * AT8_ROW;                 // 16 K scales | 16 V scales
constexpr int AT8_STAGE_BYTES = AT8_SCALE_OFF + 2 * AT_BLK * 4;     // 4224
constexpr int AT8_SMEM = AT_WARPS * AT8_STAGES * AT8_STAGE_BYTES;   // 101,376
constexpr float AT8_PV_UP = 256.f, AT8_PV_DOWN = 1.f / 256.f;

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}
// two offset-binary bytes of y (selected by `sel`) -> f16x2 of (code + 1152)
__device__ __forceinline__ uint32_t u8pair_to_f16x2(uint32_t y, uint32_t sel) { return prmt(y, 0x64646464u, sel); }
constexpr float AT8_BIAS = 1152.f;
__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi) {
    __half2 t = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void mma_f16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// per-(token, head) symmetric int8 quantisation, the reference's arithmetic (kv_cache.py:198-218) at token granularity:
// scale = 127 / absmax (absmax < 1e-6 -> 1), code = clamp(round(x * scale), -127, 127), stored dequant scale = absmax / 127
__device__ __forceinline__ float kv8_inv(float& amax) {
    if (amax < 1e-6f) amax = 1.0f;
    return __fdiv_rn(127.0f, amax);
}
__device__ __forceinline__ int kv8_code(float x, float inv) { return (int)fminf(fmaxf(rintf(__fmul_rn(x, inv)), -127.f), 127.f); }
__device__ __forceinline__ uint8_t kv8_store(float x, float inv) { return (uint8_t)(kv8_code(x, inv) + 128); }   // offset binary
__device__ __forceinline__ float kv8_scale(float amax) { return __fdiv_rn(amax, 127.0f); }

// ex2.approx: one MUFU instead of exp2f()'s range-scaling sequence; exp2(-inf) = 0, 2 ulp — far inside the 16-bit rounding of P
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float ld_dsmem_f32(const float* local_ptr, int cta_rank) {   // the same shared-memory address in CTA `cta_rank` of this cluster
    uint32_t ra;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local_ptr)), "r"(cta_rank));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
    return v;
}

struct AttnParams {
    const __nv_bfloat16* q;      // [B][H][D]
    const __nv_bfloat16* pool;
    const int32_t* block_table;  // [B][max_blocks]
    const int32_t* seq_lens;     // [B]
    int B, H, Hkv, G;
    int layers, layer, page_size, max_blocks;
    int page_shift;              // log2(page_size) when it is a power of two, else -1
    int chunk_blocks;            // 16-token blocks per CTA
    int max_chunks;              // partial slots per (b, h)
    int min_cb;                  // > 0: a sequence's ACTUAL blocks are cut into max_chunks near-equal chunks of >= min_cb blocks (chunk_blocks, which
                                 // the host derives from the graph's length hint, is then only the upper bound); 0: fixed chunk_blocks
    float scale_log2e;           // softmax scale * log2(e)
    void* out;                   // [B][H][D] bf16 or f32
    int out_dtype;
    float* part_o;               // [B][H][max_chunks][D]
    float* part_ml;              // [B][H][max_chunks][2]
    int direct;                  // 1: single chunk per sequence everywhere -> write final output
    int cluster;                 // 1 (fused form, small grids): the chunks of a (sequence, kv head) pair are ONE thread-block cluster; their partials
                                 // meet in distributed shared memory (two cluster barriers) instead of global memory + fence + atomic + reload
    int two_short_first;         // two-class grid: dispatch the short CTAs before the whole pairs (WF_ATTN_TWO_ORDER=1; measured worse: 5.26 vs 5.10 ms
                                 // at bs64/4K, 8.43 vs 8.09 at 8K, profiles/r02ag_summary.txt)
    int two_pa, two_nb;          // two-class grid (fused form): the first two_pa pairs are ONE CTA each (no split, no merge), every later pair is
                                 // two_nb CTAs of chunk_blocks blocks; 1-D grid, long CTAs first (two_nb = 0: uniform grid)
    int chunk_major;             // legacy grid order: 1 = (kv head, sequence, chunk), 0 = (chunk, kv head, sequence)
    int even;                    // 1 (fused form only): the grid is ONE wave of persistent CTAs and the (sequence, kv head, block) space is cut
                                 // into gridDim.x equal contiguous ranges (see the kernel); chunk_blocks / max_chunks / direct are unused
    // ---- fused prologue / epilogue (decode engine): q, k, v come straight from the qkv linear's raw accumulator
    const float* qkv_acc;        // [B][qkv_ld] fp32 exact integers: sum (w+1)*a, rows q | k | v
    int qkv_ld;
    const int32_t* qkv_rowsum;   // [B]
    const float* qkv_ascale;     // [B]
    const float* qkv_wscale;     // device [3]: q, k, v weight scales
    const int32_t* positions;    // [B] position of the new token (== seq_len - 1)
    const __nv_bfloat16* cos_tab;
    const __nv_bfloat16* sin_tab;
    int max_pos;
    __nv_bfloat16* pool_w;       // pool, writable (append)
    int* merge_cnt;              // [B][Hkv] arrival counters (zero; self-cleaning)
    float* zero0; int zero0_n;   // accumulator range this phase clears
    // ---- tensor parallel: this rank's heads land in [B][out_ld] buffers at column out_col0, on every rank
    void* out_peer[TP_MAX];
    int n_out, out_ld, out_col0, signal_flag;
    TpComm tp;
    unsigned long long* trace;
    float* kv_scales;            // int8 KV: scale region of the pool (null = bf16 pool)
};

// final output element (b, local head h, d): local buffer, or the attn_full buffer of every rank (the all-gather)
__device__ __forceinline__ void attn_store(const AttnParams& p, int b, int h, int d, float v) {
    if (p.n_out <= 1) {
        const size_t o = (size_t)b * p.out_ld + p.out_col0 + (size_t)h * AT_D + d;
        if (p.out_dtype == WF_BF16) reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(v);
        else reinterpret_cast<float*>(p.out)[o] = v;
    } else {
        const size_t o = (size_t)b * p.out_ld + p.out_col0 + (size_t)h * AT_D + d;
        const __nv_bfloat16 bv = __float2bfloat16_rn(v);
        for (int r = 0; r < p.n_out; ++r) reinterpret_cast<__nv_bfloat16*>(p.out_peer[r])[o] = bv;
    }
}

__device__ __forceinline__ size_t kv_block_offset(int page, int layers, int layer, int kv, int Hkv, int head, int page_size) {
    return ((((size_t)page * layers + layer) * 2 + kv) * Hkv + head) * (size_t)page_size * AT_D;
}

__device__ __forceinline__ size_t kv_scale_offset(int page, int layers, int layer, int kv, int Hkv, int head, int page_size) {
    return ((((size_t)page * layers + layer) * 2 + kv) * Hkv + head) * (size_t)page_size;
}

// Blocks per chunk of a sequence with `nblocks` 16-token blocks.  The host sizes chunk_blocks from the length HINT of the captured graph
// (a power of two >= the longest sequence); a sequence well below the hint would fill only the first chunks (2049 tokens under a 4096
// hint: chunks of 128 + 1 blocks) — so every (sequence, kv head) pair cuts its own length evenly.  Same value in every CTA of the pair
// and in the merge kernel (it depends on seq_lens only).
__device__ __forceinline__ int chunk_blocks_of(int nblocks, int chunk_blocks, int max_chunks, int min_cb) {
    if (min_cb <= 0 || max_chunks <= 1) return chunk_blocks;
    int c = ((max(nblocks, 1) + max_chunks - 1) / max_chunks + 3) & ~3;
    c = max(c, min_cb);
    return min(c, chunk_blocks);
}

template <bool FUSED, bool I8>
__global__ void __launch_bounds__(AT_WARPS * 32, I8 ? 3 : 2) paged_decode_attn_kernel(const AttnParams p) {
    constexpr int ST = I8 ? AT8_STAGES : AT_STAGES;
    constexpr int SB = I8 ? AT8_STAGE_BYTES : AT_STAGE_BYTES;
    constexpr int ROW = I8 ? AT8_ROW : AT_ROW;
    // the even form (opt-in, measured slower) is not compiled into the int8 kernel: its range state pushed that instantiation over the
    // 170 registers three CTAs per SM allow (80 bytes of spills in an issue-bound loop; 84 vs 74 us per layer at batch 64 / 4K)
    constexpr bool EVEN_OK = FUSED && !I8;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int nctas = gridDim.x * gridDim.y * gridDim.z;
    const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;

    const bool tr = FUSED && !I8 && p.trace && bid == 0 && threadIdx.x == 0;
    if (tr) p.trace[0] = gtimer();
    if (FUSED) pdl_launch_dependents();

    // ---- work decomposition.
    // Legacy form (p.even == 0): CTA = (chunk, kv head, sequence), one segment.
    // Even form (fused decode step with enough work): the grid is one resident wave; the linearised space
    //   x = (sequence b, kv head, 16-token block)   — block fastest, W = Hkv * sum_b nblocks(b) points —
    // is cut into gridDim.x equal contiguous ranges [bid * W / n, (bid + 1) * W / n).  A CTA walks its range as up to a few
    // SEGMENTS (a segment = a run of blocks of one (b, kv head) pair).  Every CTA streams the same number of KV bytes (+-1
    // block), so there is no partial last wave (1280 CTAs on 296 slots = 4.32 waves cost 5) and no CTA launch/prologue per
    // 64 blocks.  A pair that lies inside one range is finished by its CTA; a pair cut by a range boundary leaves one partial
    // per CTA — only the first and the last segment of a range can be partial, so a CTA owns two partial slots — and the
    // last CTA to arrive for the pair merges them (merge_cnt, as in the legacy form).
    // Everything here depends on seq_lens only, which is stable during a decode step: it runs BEFORE griddepcontrol.wait.
    __shared__ int s_cum[AT_EVEN_MAXB + 1];
    __shared__ int s_wsum[AT_WARPS];
    long long W = 0, x_lin = 0, x_hi = 0;
    if (EVEN_OK && p.even) {
        int nbv = 0;
        if (tid < p.B) nbv = max(1, (p.seq_lens[tid] + AT_BLK - 1) / AT_BLK) * p.Hkv;
        int incl = nbv;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        int base = 0;
        for (int w = 0; w < warp; ++w) base += s_wsum[w];
        if (tid < p.B) s_cum[tid + 1] = base + incl;
        if (tid == 0) s_cum[0] = 0;
        __syncthreads();
        W = s_cum[p.B];
        x_lin = (long long)bid * W / nctas;
        x_hi = (long long)(bid + 1) * W / nctas;
    }
    const long long range_lo = x_lin;
    int seg_i = 0;          // segments done by this CTA
    int b_scan = 0;         // even form: sequence of the previous segment (ranges are walked in increasing x)
    bool waited = false;    // griddepcontrol.wait executed
    unsigned char* wsm = smem + warp * (ST * SB);

    for (;; ++seg_i) {
    // ---- this segment: blocks [blk0, blk1) of pair (b, kvh)
    int chunk, kvh, b, seq_len, nblocks, blk0, blk1;
    bool active, seg_direct;
    int nch = 1, my_c = 0, seg_c0 = 0, seg_c0_which = 0;   // partials of the pair: count, this CTA's index, first CTA, its slot
    if (!(EVEN_OK && p.even)) {
        if (seg_i > 0) break;
        // chunk index slowest in the dispatch order: the last (shortest) chunk of every sequence forms the final partial wave
        int cbk = p.chunk_blocks;
        bool whole = false;
        if (FUSED && p.two_nb > 0) {
            // two-class grid: as many whole pairs as fill the resident slots exactly stream unsplit; the leftover pairs — which would run
            // as a thin second wave at a fraction of the HBM bandwidth — are cut fine enough to fill one short wave of their own
            int pair;
            const int nshort = (p.B * p.Hkv - p.two_pa) * p.two_nb;
            const int wid = (!I8 && p.two_short_first) ? bid - nshort : bid;            // index among the whole pairs (when in range)
            if (wid >= 0 && wid < p.two_pa) { pair = wid; chunk = 0; whole = true; }
            else { const int j = (!I8 && p.two_short_first) ? bid : bid - p.two_pa; pair = p.two_pa + j / p.two_nb; chunk = j - (pair - p.two_pa) * p.two_nb; }
            b = pair / p.Hkv; kvh = pair - b * p.Hkv;
        } else if (!I8 && p.chunk_major) { kvh = blockIdx.x; b = blockIdx.y; chunk = blockIdx.z; }
        else { chunk = blockIdx.x; kvh = blockIdx.y; b = blockIdx.z; }
        seq_len = p.seq_lens[b];
        nblocks = (seq_len + AT_BLK - 1) / AT_BLK;
        cbk = chunk_blocks_of(nblocks, cbk, p.max_chunks, p.min_cb);
        if (whole) cbk = max(nblocks, 1);
        blk0 = chunk * cbk;
        active = !(blk0 >= nblocks && chunk != 0);
        blk1 = min(nblocks, blk0 + cbk);
        seg_direct = whole || p.direct != 0;
        nch = (nblocks + cbk - 1) / cbk;
        my_c = chunk;
    } else {
        if (x_lin >= x_hi) break;
        while (s_cum[b_scan + 1] <= x_lin) ++b_scan;
        b = b_scan;
        seq_len = p.seq_lens[b];
        nblocks = max(1, (seq_len + AT_BLK - 1) / AT_BLK);
        const int r = (int)(x_lin - s_cum[b]);
        kvh = r / nblocks;
        blk0 = r - kvh * nblocks;
        const long long left = x_hi - x_lin;
        blk1 = (left < (long long)(nblocks - blk0)) ? blk0 + (int)left : nblocks;
        chunk = 0;
        active = true;
        seg_direct = (blk0 == 0 && blk1 == nblocks);
        if (!seg_direct) {
            const long long pair_lo = (long long)s_cum[b] + (long long)kvh * nblocks, pair_hi = pair_lo + nblocks;
            seg_c0 = (int)(((pair_lo + 1) * nctas - 1) / W);            // CTA whose range holds the pair's first block
            const int c1 = (int)((pair_hi * nctas - 1) / W);             // ... and its last block
            nch = c1 - seg_c0 + 1;
            my_c = bid - seg_c0;
            seg_c0_which = ((long long)seg_c0 * W / nctas == pair_lo) ? 0 : 1;   // the pair is the first segment of seg_c0's range, or a later (= its last) one
        }
        x_lin += blk1 - blk0;
    }
    // slot of partial c (0 .. nch-1) of this pair, head hh
    auto pslot = [&](int c, int hh) -> size_t {
        if (!(EVEN_OK && p.even)) return ((size_t)b * p.H + (size_t)kvh * p.G + hh) * p.max_chunks + c;
        const int which = (c == 0) ? seg_c0_which : 0;               // every later CTA starts its range inside the pair
        return ((size_t)(seg_c0 + c) * 2 + which) * 8 + hh;
    };
    const int mstride = (EVEN_OK && p.even) ? nch : p.max_chunks;       // row stride of the merge scratch in shared memory
    const int blk_pos = (seq_len - 1) / AT_BLK;          // block of the token appended in this step (pos = seq_len - 1)
    const int my_n = active ? (blk1 - blk0 - warp + AT_WARPS - 1) / AT_WARPS : 0;   // blocks blk0 + warp + 4*i
    const int32_t* bt = p.block_table + (size_t)b * p.max_blocks;

    // page of this warp's block i, looked up THREE issues ahead (a 3-deep register queue): one issue = one block of compute (~0.3-0.5 us),
    // an L2 round trip under load is ~1 us — with a single entry ahead 13 % of the kernel's stall samples sat on this dependency
    // (profiles/r02_full_attn_fused_bs64.summary.txt, attn.cu kv_block_offset)
    auto page_of = [&](int i) -> int {
        const int tokf = (blk0 + warp + AT_WARPS * i) * AT_BLK;
        const int pi = p.page_shift >= 0 ? (tokf >> p.page_shift) : tokf / p.page_size;
        return bt[min(pi, p.max_blocks - 1)];
    };
    int pg_q0 = page_of(0), pg_q1 = page_of(1), pg_q2 = page_of(2);
    auto issue = [&](int i) {   // i-th block of this warp -> stage i % AT_STAGES
        if (i < my_n) {
            const int blk = blk0 + warp + AT_WARPS * i;
            const int tok0 = blk * AT_BLK;
            const int page = pg_q0;
            pg_q0 = pg_q1; pg_q1 = pg_q2; pg_q2 = page_of(i + 3);
            const int slot0 = p.page_shift >= 0 ? (tok0 & (p.page_size - 1)) : tok0 % p.page_size;
            const size_t ko = kv_block_offset(page, p.layers, p.layer, 0, p.Hkv, kvh, p.page_size) + (size_t)slot0 * AT_D;
            const size_t vo = kv_block_offset(page, p.layers, p.layer, 1, p.Hkv, kvh, p.page_size) + (size_t)slot0 * AT_D;
            unsigned char* ks = wsm + (i % ST) * SB;
            unsigned char* vs = ks + AT_BLK * ROW;
            const int valid = seq_len - tok0;   // tokens >= valid are zero-filled (pool slots may be uninitialised)
            if (!I8) {
                const unsigned char* kb = reinterpret_cast<const unsigned char*>(p.pool + ko);
                const unsigned char* vb = reinterpret_cast<const unsigned char*>(p.pool + vo);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int gi = lane + 32 * j;
                    const int r = gi >> 4, c = gi & 15;
                    const int dst = r * ROW + ((c ^ (r & 7)) << 4);
                    const int nb = (r < valid) ? 16 : 0;
                    cp_async16_zfill(ks + dst, kb + gi * 16, nb);
                    cp_async16_zfill(vs + dst, vb + gi * 16, nb);
                }
            } else {
                const unsigned char* kb = reinterpret_cast<const unsigned char*>(p.pool) + ko;
                const unsigned char* vb = reinterpret_cast<const unsigned char*>(p.pool) + vo;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int gi = lane + 32 * j;
                    const int r = gi >> 3, c = gi & 7;
                    const int dst = r * ROW + ((c ^ (r & 7)) << 4);
                    const int nb = (r < valid) ? 16 : 0;
                    cp_async16_zfill(ks + dst, kb + gi * 16, nb);
                    cp_async16_zfill(vs + dst, vb + gi * 16, nb);
                }
                // 16 K scales | 16 V scales, 16-byte granules through L2 (.cg): a scale appended by another CTA in this
                // launch must never be served from a stale L1 line
                if (lane < 8) {
                    const int kv = lane >> 2, q4 = (lane & 3) * 4;
                    const float* sc = p.kv_scales + kv_scale_offset(page, p.layers, p.layer, kv, p.Hkv, kvh, p.page_size) + slot0 + q4;
                    cp_async16_zfill(ks + AT8_SCALE_OFF + lane * 16, sc, 4 * max(0, min(4, valid - q4)));
                }
            }
        }
        cp_async_commit();
    };
    int pre = 0;   // stages issued before the dependency wait (first segment only)
    if (!waited) {
        while (pre < ST - 1 && (pre >= my_n || blk0 + warp + AT_WARPS * pre < blk_pos)) { issue(pre); ++pre; }
        pdl_wait_prior();
        waited = true;
        if (tr) p.trace[1] = gtimer();
        if (FUSED && p.zero0) {
            for (int i = (bid * AT_WARPS * 32 + tid) * 4; i < p.zero0_n; i += nctas * AT_WARPS * 32 * 4)
                *reinterpret_cast<float4*>(p.zero0 + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    if (!active && !p.cluster) break;   // (a cluster member without work still owes the cluster barriers: it contributes an empty partial)

    // ---- Q fragments (rows g < G valid; rows 8..15 are zero padding)
    uint32_t qa[8][2];
    if (!FUSED) {
        const __nv_bfloat16* qrow = p.q + ((size_t)b * p.H + (size_t)kvh * p.G + g) * AT_D;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            if (!I8) {
                qa[ks][0] = (g < p.G) ? *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + 2 * t) : 0u;
                qa[ks][1] = (g < p.G) ? *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + 8 + 2 * t) : 0u;
            } else {   // d = 16 ks + 4t .. 4t+3 (the order the int8 K bytes come out of ldmatrix), bf16 -> f16 is exact here
                uint2 w = make_uint2(0u, 0u);
                if (g < p.G) w = *reinterpret_cast<const uint2*>(qrow + ks * 16 + 4 * t);
                qa[ks][0] = pack_f16x2(bf16lo(w.x), bf16hi(w.x));
                qa[ks][1] = pack_f16x2(bf16lo(w.y), bf16hi(w.y));
            }
        }
    } else {
        // dequantise the qkv accumulator (direct_inference_full.py:204-206), round to bf16 as the linear's output
        // tensor, half-split RoPE with the reference's bf16 rounding points (:58-70), append k/v (:226-227).
        const int pos = p.positions[b];
        const int pc = min(pos, p.max_pos - 1);
        const float* arow = p.qkv_acc + (size_t)b * p.qkv_ld;
        const int rs = p.qkv_rowsum[b];
        const float sa = p.qkv_ascale[b];
        const __nv_bfloat16* ct = p.cos_tab + (size_t)pc * (AT_D / 2);
        const __nv_bfloat16* st = p.sin_tab + (size_t)pc * (AT_D / 2);
        {
            const float wq = p.qkv_wscale[0];
            const float* hq = arow + ((size_t)kvh * p.G + g) * AT_D;
            // The fragments depend on the lane only, not on the warp: each of the 4 warps dequantises + rotates a quarter
            // (i = 2*warp, 2*warp + 1) and the CTA shares them through 2 KB of shared memory - a quarter of the loads and of
            // the ALU work per warp on the prologue every CTA pays.  Loads are issued before the first use.
            __shared__ uint32_t qsh[16][32];
            float2 xa[2], xb[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int i = 2 * warp + u;
                const int d = I8 ? (i >> 1) * 16 + 4 * t + 2 * (i & 1) : (i >> 1) * 16 + (i & 1) * 8 + 2 * t;   // < 64; partner at d + 64
                xa[u] = make_float2(0.f, 0.f);
                xb[u] = make_float2(0.f, 0.f);
                if (g < p.G) {
                    asm volatile("ld.global.cg.v2.f32 {%0,%1}, [%2];" : "=f"(xa[u].x), "=f"(xa[u].y) : "l"(hq + d));
                    asm volatile("ld.global.cg.v2.f32 {%0,%1}, [%2];" : "=f"(xb[u].x), "=f"(xb[u].y) : "l"(hq + d + 64));
                }
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int i = 2 * warp + u;
                const int ks = i >> 1, half = i & 1;
                const int d = I8 ? ks * 16 + 4 * t + 2 * half : ks * 16 + half * 8 + 2 * t;
                const float a[2] = {xa[u].x, xa[u].y}, bq[2] = {xb[u].x, xb[u].y};
                float o1[2], o2[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    float x1 = 0.f, x2 = 0.f;
                    if (g < p.G) {
                        x1 = bf16_round(lin_value(a[e], rs, wq, sa));
                        x2 = bf16_round(lin_value(bq[e], rs, wq, sa));
                    }
                    const float c = __bfloat162float(ct[d + e]), sn = __bfloat162float(st[d + e]);
                    o1[e] = bf16_round(x1 * c) - bf16_round(x2 * sn);
                    o2[e] = bf16_round(x1 * sn) + bf16_round(x2 * c);
                }
                qsh[ks * 2 + half][lane] = I8 ? pack_f16x2(bf16_round(o1[0]), bf16_round(o1[1])) : pack_bf16x2(o1[0], o1[1]);
                qsh[(ks + 4) * 2 + half][lane] = I8 ? pack_f16x2(bf16_round(o2[0]), bf16_round(o2[1])) : pack_bf16x2(o2[0], o2[1]);
            }
            __syncthreads();
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) { qa[ks][0] = qsh[ks * 2][lane]; qa[ks][1] = qsh[ks * 2 + 1][lane]; }
        }
        // the CTA whose chunk holds the new token appends k (RoPE'd) and v before it streams that block
        const int blk_pos = pos / AT_BLK;
        if (blk_pos >= blk0 && blk_pos < blk1) {
            const int page = bt[pos / p.page_size];
            const int slot = pos % p.page_size;
            const int d = tid & 63;
            const int kv = tid >> 6;   // warps 0,1: k   warps 2,3: v
            float y1, y2;              // the bf16 values the bf16 cache would hold at d and d + 64
            if (kv == 0) {
                const float wk = p.qkv_wscale[1];
                const float* hk = arow + ((size_t)p.H + kvh) * AT_D;
                const float x1 = bf16_round(lin_value(ldcg_f(hk + d), rs, wk, sa));
                const float x2 = bf16_round(lin_value(ldcg_f(hk + d + 64), rs, wk, sa));
                const float c = __bfloat162float(ct[d]), s = __bfloat162float(st[d]);
                y1 = bf16_round(bf16_round(x1 * c) - bf16_round(x2 * s));
                y2 = bf16_round(bf16_round(x1 * s) + bf16_round(x2 * c));
            } else {
                const float wv = p.qkv_wscale[2];
                const float* hv = arow + ((size_t)p.H + p.Hkv + kvh) * AT_D;
                y1 = bf16_round(lin_value(ldcg_f(hv + d), rs, wv, sa));
                y2 = bf16_round(lin_value(ldcg_f(hv + d + 64), rs, wv, sa));
            }
            const size_t off = kv_block_offset(page, p.layers, p.layer, kv, p.Hkv, kvh, p.page_size) + (size_t)slot * AT_D;
            if (!I8) {
                __nv_bfloat16* dst = p.pool_w + off;
                dst[d] = __float2bfloat16_rn(y1);
                dst[d + 64] = __float2bfloat16_rn(y2);
            } else {
                __shared__ float s_amax[AT_WARPS];
                float am = fmaxf(fabsf(y1), fabsf(y2));
#pragma unroll
                for (int o = 16; o; o >>= 1) am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, o));
                if (lane == 0) s_amax[warp] = am;
                __syncthreads();
                float amax = fmaxf(s_amax[2 * kv], s_amax[2 * kv + 1]);
                const float inv = kv8_inv(amax);
                uint8_t* dst = reinterpret_cast<uint8_t*>(p.pool_w) + off;
                dst[d] = kv8_store(y1, inv);
                dst[d + 64] = kv8_store(y2, inv);
                if (d == 0) p.kv_scales[kv_scale_offset(page, p.layers, p.layer, kv, p.Hkv, kvh, p.page_size) + slot] = kv8_scale(amax);
            }
            __threadfence();
            __syncthreads();
        }
    }

    if (tr) p.trace[2] = gtimer();
    float o[16][4];
#pragma unroll
    for (int j = 0; j < 16; ++j) { o[j][0] = 0.f; o[j][1] = 0.f; o[j][2] = 0.f; o[j][3] = 0.f; }
    float m_run = -INFINITY, l_run = 0.f;   // row g (replicated over the 4 lanes of a group)
    float c_run = 0.f, q_bias = 0.f;        // int8 KV: running sum of the f16 P operands (per lane); 1152 * sum_d q[g][d]
    if (I8) {
        float qs = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&qa[ks][0]));
            const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&qa[ks][1]));
            qs += (a.x + a.y) + (c.x + c.y);
        }
        qs += __shfl_xor_sync(0xffffffffu, qs, 1);
        qs += __shfl_xor_sync(0xffffffffu, qs, 2);
        q_bias = AT8_BIAS * qs;
    }

    for (int s = pre; s < ST - 1; ++s) issue(s);

    for (int i = 0; i < my_n; ++i) {
        issue(i + ST - 1);
        cp_async_wait<ST - 1>();
        __syncwarp();
        const unsigned char* ks = wsm + (i % ST) * SB;
        const unsigned char* vs = ks + AT_BLK * ROW;
        const int tok0 = (blk0 + warp + AT_WARPS * i) * AT_BLK;

        // ---- S = Q K^T : 16 (heads) x 16 (tokens)
        float s[2][4];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) { s[nt][0] = 0.f; s[nt][1] = 0.f; s[nt][2] = 0.f; s[nt][3] = 0.f; }
        if (!I8) {
            const int r = (lane & 7) + ((lane >> 4) << 3);   // token row supplied by this lane
            const int gsel = (lane >> 3) & 1;                // granule parity supplied by this lane
#pragma unroll
            for (int ks_i = 0; ks_i < 8; ++ks_i) {
                const int c = 2 * ks_i + gsel;
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4(b0, b1, b2, b3, ks + r * ROW + ((c ^ (r & 7)) << 4));
                mma_bf16_16816(s[0], qa[ks_i][0], 0u, qa[ks_i][1], 0u, b0, b1);
                mma_bf16_16816(s[1], qa[ks_i][0], 0u, qa[ks_i][1], 0u, b2, b3);
            }
        } else {
            // matrices: (tokens 0-7, granule 2i) (tokens 8-15, 2i) (tokens 0-7, 2i+1) (tokens 8-15, 2i+1); a register is the
            // int8 d = 16c + 4t .. 4t+3 of token g: bytes (0,1) are the k slots 2t,2t+1 and bytes (2,3) the slots 2t+8,2t+9
            const int r = (lane & 7) + (((lane >> 3) & 1) << 3);
            const int gsel = lane >> 4;
#pragma unroll
            for (int i2 = 0; i2 < 4; ++i2) {
                const int c = 2 * i2 + gsel;
                uint32_t r0, r1, r2, r3;
                ldmatrix_x4(r0, r1, r2, r3, ks + r * ROW + ((c ^ (r & 7)) << 4));
                mma_f16_16816(s[0], qa[2 * i2][0], 0u, qa[2 * i2][1], 0u, u8pair_to_f16x2(r0, 0x4140u), u8pair_to_f16x2(r0, 0x4342u));
                mma_f16_16816(s[1], qa[2 * i2][0], 0u, qa[2 * i2][1], 0u, u8pair_to_f16x2(r1, 0x4140u), u8pair_to_f16x2(r1, 0x4342u));
                mma_f16_16816(s[0], qa[2 * i2 + 1][0], 0u, qa[2 * i2 + 1][1], 0u, u8pair_to_f16x2(r2, 0x4140u), u8pair_to_f16x2(r2, 0x4342u));
                mma_f16_16816(s[1], qa[2 * i2 + 1][0], 0u, qa[2 * i2 + 1][1], 0u, u8pair_to_f16x2(r3, 0x4140u), u8pair_to_f16x2(r3, 0x4342u));
            }
        }
        float2 ksc[2], vsc[2];   // int8 KV: scales of tokens 8nt + 2t, +1
        if (I8) {
            const float* scs = reinterpret_cast<const float*>(ks + AT8_SCALE_OFF);
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                ksc[nt] = *reinterpret_cast<const float2*>(scs + nt * 8 + 2 * t);
                vsc[nt] = *reinterpret_cast<const float2*>(scs + AT_BLK + nt * 8 + 2 * t);
            }
        }
        // ---- online softmax for row g over this block's 16 tokens (columns 8nt+2t, 8nt+2t+1)
        float mx = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int tok = tok0 + nt * 8 + 2 * t + e;
                float sv = s[nt][e];
                if (I8) sv = (sv - q_bias) * (e ? ksc[nt].y : ksc[nt].x);
                const float v = (tok < seq_len) ? sv * p.scale_log2e : -INFINITY;
                s[nt][e] = v;
                mx = fmaxf(mx, v);
            }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float m_new = fmaxf(m_run, mx);
        const float corr = fast_exp2(m_run - m_new);   // 0 on the first block (m_run = -inf)
        float psum = 0.f;
        uint32_t pa[4];
        {
            const float p00 = fast_exp2(s[0][0] - m_new), p01 = fast_exp2(s[0][1] - m_new);
            const float p10 = fast_exp2(s[1][0] - m_new), p11 = fast_exp2(s[1][1] - m_new);
            psum = (p00 + p01) + (p10 + p11);
            if (!I8) {
                pa[0] = pack_bf16x2(p00, p01);   // A frag: row g, k = 2t,2t+1
                pa[2] = pack_bf16x2(p10, p11);   // row g, k = 8+2t,..
            } else {                             // probability * V scale of the token (* 2^8, undone at the end)
                pa[0] = pack_f16x2(p00 * vsc[0].x * AT8_PV_UP, p01 * vsc[0].y * AT8_PV_UP);
                pa[2] = pack_f16x2(p10 * vsc[1].x * AT8_PV_UP, p11 * vsc[1].y * AT8_PV_UP);
                // sum of the ROUNDED operands (this lane's 4 tokens; the 4 lanes of a row are added up once, at the end)
                const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&pa[0]));
                const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&pa[2]));
                c_run = c_run * corr + ((f0.x + f0.y) + (f1.x + f1.y));
            }
            pa[1] = 0u;                          // rows 8..15 unused
            pa[3] = 0u;
        }
        psum += __shfl_xor_sync(0xffffffffu, psum, 1);
        psum += __shfl_xor_sync(0xffffffffu, psum, 2);
        l_run = l_run * corr + psum;
        m_run = m_new;
        if (__any_sync(0xffffffffu, corr != 1.0f)) {
#pragma unroll
            for (int j = 0; j < 16; ++j) { o[j][0] *= corr; o[j][1] *= corr; }
        }
        // ---- O += P V : V^T fragments via ldmatrix.trans
        if (!I8) {
            const int r = (lane & 7) + (((lane >> 3) & 1) << 3);   // token row
            const int gsel = lane >> 4;                            // d-granule parity
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
                const int c = 2 * jj + gsel;
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4_trans(b0, b1, b2, b3, vs + r * ROW + ((c ^ (r & 7)) << 4));
                mma_bf16_16816(o[2 * jj], pa[0], pa[1], pa[2], pa[3], b0, b1);
                mma_bf16_16816(o[2 * jj + 1], pa[0], pa[1], pa[2], pa[3], b2, b3);
            }
        } else {
            // transposed b16 units: a register holds V[tok 2t][d0], V[2t][d0+1], V[2t+1][d0], V[2t+1][d0+1] with d0 = 16c + 2g;
            // bytes (0,2) are the B fragment of column d0, bytes (1,3) of column d0 + 1: accumulator o[2c + par] element e
            // is d = 16c + 4t + 2e + par
            const int r = (lane & 7) + (((lane >> 3) & 1) << 3);
            const int gsel = lane >> 4;
#pragma unroll
            for (int i2 = 0; i2 < 4; ++i2) {
                const int c = 2 * i2 + gsel;
                uint32_t r0, r1, r2, r3;
                ldmatrix_x4_trans(r0, r1, r2, r3, vs + r * ROW + ((c ^ (r & 7)) << 4));
                mma_f16_16816(o[4 * i2 + 0], pa[0], pa[1], pa[2], pa[3], u8pair_to_f16x2(r0, 0x4240u), u8pair_to_f16x2(r1, 0x4240u));
                mma_f16_16816(o[4 * i2 + 1], pa[0], pa[1], pa[2], pa[3], u8pair_to_f16x2(r0, 0x4341u), u8pair_to_f16x2(r1, 0x4341u));
                mma_f16_16816(o[4 * i2 + 2], pa[0], pa[1], pa[2], pa[3], u8pair_to_f16x2(r2, 0x4240u), u8pair_to_f16x2(r3, 0x4240u));
                mma_f16_16816(o[4 * i2 + 3], pa[0], pa[1], pa[2], pa[3], u8pair_to_f16x2(r2, 0x4341u), u8pair_to_f16x2(r3, 0x4341u));
            }
        }
        __syncwarp();
    }
    cp_async_wait<0>();
    __syncthreads();
    if (tr) p.trace[3] = gtimer();

    // ---- merge the 4 warps through shared memory: [warp][G][D] fp32 + [warp][G][2]
    float* mo = reinterpret_cast<float*>(smem);                       // AT_WARPS * 8 * 128 floats = 16 KB
    float* mml = mo + AT_WARPS * 8 * AT_D;                            // AT_WARPS * 8 * 2
    if (g < p.G) {
        if (!I8) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                *reinterpret_cast<float2*>(mo + ((warp * 8 + g) * AT_D) + j * 8 + 2 * t) = make_float2(o[j][0], o[j][1]);
            }
        } else {
            float cs = c_run;
            cs += __shfl_xor_sync(0xfu << (lane & 28), cs, 1);
            cs += __shfl_xor_sync(0xfu << (lane & 28), cs, 2);
            const float ob = AT8_BIAS * cs;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                *reinterpret_cast<float4*>(mo + ((warp * 8 + g) * AT_D) + c * 16 + 4 * t) =
                    make_float4((o[2 * c][0] - ob) * AT8_PV_DOWN, (o[2 * c + 1][0] - ob) * AT8_PV_DOWN,
                                (o[2 * c][1] - ob) * AT8_PV_DOWN, (o[2 * c + 1][1] - ob) * AT8_PV_DOWN);
            }
        }
        if (t == 0) {
            mml[(warp * 8 + g) * 2 + 0] = m_run;
            mml[(warp * 8 + g) * 2 + 1] = l_run;
        }
    }
    __syncthreads();
    {
        const int d = tid;   // 128 threads == AT_D
        for (int hh = 0; hh < p.G; ++hh) {
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < AT_WARPS; ++w) M = fmaxf(M, mml[(w * 8 + hh) * 2]);
            float L = 0.f, O = 0.f;
#pragma unroll
            for (int w = 0; w < AT_WARPS; ++w) {
                const float mw = mml[(w * 8 + hh) * 2];
                const float f = (mw == -INFINITY) ? 0.f : exp2f(mw - M);
                L += mml[(w * 8 + hh) * 2 + 1] * f;
                O += mo[(w * 8 + hh) * AT_D + d] * f;
            }
            const int h = kvh * p.G + hh;
            if (seg_direct) {
                attn_store(p, b, h, d, (L > 0.f) ? O / L : 0.f);
            } else if (FUSED && p.cluster) {
                float* cl_o = mml + AT_WARPS * 8 * 2;          // [8][AT_D] + [8][2], behind the 4-warp merge scratch
                cl_o[hh * AT_D + d] = O;
                if (d == 0) { cl_o[8 * AT_D + hh * 2] = M; cl_o[8 * AT_D + hh * 2 + 1] = L; }
            } else {
                const size_t slot = pslot(my_c, hh);
                p.part_o[slot * AT_D + d] = O;
                if (d == 0) { p.part_ml[slot * 2] = M; p.part_ml[slot * 2 + 1] = L; }
            }
        }
    }
    if (tr) p.trace[4] = gtimer();
    if (FUSED && p.cluster && !seg_direct) {
        // cluster form: every chunk's (M, L, O) sits in its CTA's shared memory and is read over DSMEM.  ~1-2 us instead of the
        // ~4 us chain of the global form below (partial store, fence, atomic, reload) — at batch 1 that chain was a third of the kernel
        const float* cl_o = mml + AT_WARPS * 8 * 2;
        const int ncl = gridDim.x;                                  // cluster = all chunks of the pair (inactive ones hold L = 0)
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        {   // the pair's heads are dealt round-robin to the cluster's CTAs: each merges its heads from every CTA's partial
            const int d = tid;
            for (int hh = chunk; hh < p.G; hh += ncl) {
                float mc[8], lc[8], oc[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    mc[c] = -INFINITY; lc[c] = 0.f; oc[c] = 0.f;
                    if (c < ncl) {
                        mc[c] = ld_dsmem_f32(cl_o + 8 * AT_D + hh * 2, c);
                        lc[c] = ld_dsmem_f32(cl_o + 8 * AT_D + hh * 2 + 1, c);
                        oc[c] = ld_dsmem_f32(cl_o + hh * AT_D + d, c);
                    }
                }
                float M = -INFINITY;
#pragma unroll
                for (int c = 0; c < 8; ++c) M = fmaxf(M, mc[c]);
                float L = 0.f, O = 0.f;
#pragma unroll
                for (int c = 0; c < 8; ++c) {                       // same order and arithmetic as the global form
                    const float f = (mc[c] == -INFINITY) ? 0.f : exp2f(mc[c] - M);
                    L += lc[c] * f;
                    O += oc[c] * f;
                }
                attn_store(p, b, kvh * p.G + hh, d, (L > 0.f) ? O / L : 0.f);
            }
        }
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");   // nobody leaves while a peer may still read its shared memory
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    if (FUSED && !seg_direct && !p.cluster) {
        // the last chunk to arrive for this (sequence, kv head) merges all of them — no separate merge launch
        __shared__ int s_last;
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            int* c = p.merge_cnt + (size_t)b * p.Hkv + kvh;
            const int prev = atomicAdd(c, 1);
            s_last = (prev == nch - 1);
            if (s_last) *c = 0;
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            // (m, l) of every (head, chunk) pair: one parallel round of loads into smem, factors computed once
            float* fml = reinterpret_cast<float*>(smem);            // [G][nch][2] -> factor, unused
            float* fL = fml + 2 * 8 * mstride;                      // [G] total l
            for (int i = tid; i < p.G * nch; i += AT_WARPS * 32) {
                const int hh = i / nch, c = i - hh * nch;
                const size_t slot = pslot(c, hh);
                float2 ml;
                asm volatile("ld.global.cg.v2.f32 {%0,%1}, [%2];" : "=f"(ml.x), "=f"(ml.y) : "l"(p.part_ml + slot * 2));
                fml[(hh * mstride + c) * 2] = ml.x;
                fml[(hh * mstride + c) * 2 + 1] = ml.y;
            }
            __syncthreads();
            if (tid < p.G) {
                const int hh = tid;
                float M = -INFINITY;
                for (int c = 0; c < nch; ++c) M = fmaxf(M, fml[(hh * mstride + c) * 2]);
                float L = 0.f;
                for (int c = 0; c < nch; ++c) {
                    const float mc = fml[(hh * mstride + c) * 2];
                    const float f = (mc == -INFINITY) ? 0.f : exp2f(mc - M);
                    L += fml[(hh * mstride + c) * 2 + 1] * f;
                    fml[(hh * mstride + c) * 2] = f;
                }
                fL[hh] = L;
            }
            __syncthreads();
            const int d = tid;
            for (int hh = 0; hh < p.G; ++hh) {
                const int h = kvh * p.G + hh;
                float O = 0.f;
                for (int c0 = 0; c0 < nch; c0 += 8) {   // 8 independent loads in flight
                    float v[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) v[u] = (c0 + u < nch) ? ldcg_f(p.part_o + pslot(c0 + u, hh) * AT_D + d) : 0.f;
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (c0 + u < nch) O += v[u] * fml[(hh * mstride + c0 + u) * 2];
                }
                const float L = fL[hh];
                attn_store(p, b, h, d, (L > 0.f) ? O / L : 0.f);
            }
        }
    }
    if (tr) p.trace[5] = gtimer();
    __syncthreads();   // the next segment's K/V ring (and its q fragments) reuse the shared memory the merge just read
    }   // segment loop
    if (!waited) {     // a CTA without work (legacy form: chunk past the sequence; even form: empty range) still owes the dependency
        pdl_wait_prior();   // wait and its share of the accumulator clearing
        if (FUSED && p.zero0) {
            for (int i = (bid * AT_WARPS * 32 + tid) * 4; i < p.zero0_n; i += nctas * AT_WARPS * 32 * 4)
                *reinterpret_cast<float4*>(p.zero0 + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    if (FUSED && p.signal_flag >= 0) tp_signal_grid_done(p.tp, p.signal_flag, nctas);
    if (!FUSED) pdl_launch_dependents();
}

struct MergeParams {
    const float* part_o;
    const float* part_ml;
    const int32_t* seq_lens;
    int H, max_chunks, chunk_blocks, min_cb;
    void* out;
    int out_dtype;
};

// grid = (H, B), block = 128 (one thread per d)
__global__ void __launch_bounds__(AT_D) attn_merge_kernel(const MergeParams p) {
    const int h = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
    pdl_wait_prior();
    const int seq_len = p.seq_lens[b];
    const int nblocks = (seq_len + AT_BLK - 1) / AT_BLK;
    const int cbk = chunk_blocks_of(nblocks, p.chunk_blocks, p.max_chunks, p.min_cb);
    const int nch = (nblocks + cbk - 1) / cbk;
    const size_t base = ((size_t)b * p.H + h) * p.max_chunks;
    float M = -INFINITY;
    for (int c = 0; c < nch; ++c) M = fmaxf(M, p.part_ml[(base + c) * 2]);
    float L = 0.f, O = 0.f;
    for (int c = 0; c < nch; ++c) {
        const float mc = p.part_ml[(base + c) * 2];
        const float f = (mc == -INFINITY) ? 0.f : exp2f(mc - M);
        L += p.part_ml[(base + c) * 2 + 1] * f;
        O += p.part_o[(base + c) * AT_D + d] * f;
    }
    const float v = (L > 0.f) ? O / L : 0.f;
    if (p.out_dtype == WF_BF16) reinterpret_cast<__nv_bfloat16*>(p.out)[((size_t)b * p.H + h) * AT_D + d] = __float2bfloat16_rn(v);
    else reinterpret_cast<float*>(p.out)[((size_t)b * p.H + h) * AT_D + d] = v;
    pdl_launch_dependents();
}

// ---------------------------------------------------------------------------------------------
// RoPE + KV append.  grid = (H + 2*Hkv, B), block = D/2 threads.
// Rounding follows the reference's bf16 tensor ops (direct_inference_full.py:58-70):
//   out1 = bf16( bf16(x1*cos) - bf16(x2*sin) ),  out2 = bf16( bf16(x1*sin) + bf16(x2*cos) )
struct RopeParams {
    const __nv_bfloat16* qkv;   // [B][(H+2Hkv)*D]
    int B, H, Hkv, D;
    __nv_bfloat16* pool;
    int layers, layer, page_size, max_blocks;
    const int32_t* positions;
    const int32_t* block_table;
    const __nv_bfloat16* cos_tab;   // [max_pos][D/2]
    const __nv_bfloat16* sin_tab;
    int max_pos;
    __nv_bfloat16* q_out;       // [B][H][D]
    float* kv_scales;           // int8 pool: scale region (null = bf16 pool); D == 128
};

__global__ void rope_kv_append_kernel(const RopeParams p) {
    const int head = blockIdx.x, b = blockIdx.y, i = threadIdx.x;
    const int half = p.D / 2;
    pdl_wait_prior();
    const int pos = p.positions[b];
    const __nv_bfloat16* src = p.qkv + (size_t)b * (p.H + 2 * p.Hkv) * p.D + (size_t)head * p.D;
    const float x1 = __bfloat162float(src[i]), x2 = __bfloat162float(src[i + half]);
    __nv_bfloat16* dst;
    size_t blk = 0;      // (page, layer, K|V, head) block index of a K/V head
    int slot = 0, kv = -1;
    if (head < p.H) {
        dst = p.q_out + ((size_t)b * p.H + head) * p.D;
    } else {
        kv = (head < p.H + p.Hkv) ? 0 : 1;
        const int kh = head - p.H - kv * p.Hkv;
        const int page = p.block_table[(size_t)b * p.max_blocks + pos / p.page_size];
        slot = pos % p.page_size;
        blk = (((size_t)page * p.layers + p.layer) * 2 + kv) * p.Hkv + kh;
        dst = p.pool + blk * (size_t)p.page_size * p.D + (size_t)slot * p.D;
    }
    float o1 = x1, o2 = x2;   // V: plain copy
    if (kv != 1) {
        const int pc = min(pos, p.max_pos - 1);
        const float c = __bfloat162float(p.cos_tab[(size_t)pc * half + i]);
        const float s = __bfloat162float(p.sin_tab[(size_t)pc * half + i]);
        o1 = bf16_round(bf16_round(x1 * c) - bf16_round(x2 * s));
        o2 = bf16_round(bf16_round(x1 * s) + bf16_round(x2 * c));
    }
    if (kv >= 0 && p.kv_scales) {   // int8 pool: per-(token, head) absmax over the 2 warps of this block (D == 128)
        __shared__ float s_amax[2];
        float am = fmaxf(fabsf(o1), fabsf(o2));
#pragma unroll
        for (int o = 16; o; o >>= 1) am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, o));
        if ((i & 31) == 0) s_amax[i >> 5] = am;
        __syncthreads();
        float amax = fmaxf(s_amax[0], s_amax[1]);
        const float inv = kv8_inv(amax);
        uint8_t* d8 = reinterpret_cast<uint8_t*>(p.pool) + blk * (size_t)p.page_size * p.D + (size_t)slot * p.D;
        d8[i] = kv8_store(o1, inv);
        d8[i + half] = kv8_store(o2, inv);
        if (i == 0) p.kv_scales[blk * p.page_size + slot] = kv8_scale(amax);
    } else {
        dst[i] = __float2bfloat16_rn(o1);
        dst[i + half] = __float2bfloat16_rn(o2);
    }
    pdl_launch_dependents();
}

// int8 pool <-> bf16 rows: one CTA per token, one warp per (K|V, head), 4 values per lane (D == 128)
template <bool SCATTER>
__global__ void kv_scatter_gather_i8_kernel(uint8_t* pool, float* scales, __nv_bfloat16* k, __nv_bfloat16* v, const int32_t* pages,
                                            const int32_t* slots, int layers, int layer, int Hkv, int page_size, int max_pages) {
    const int tkn = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int page = pages[tkn], slot = slots[tkn];
    if (page < 0 || page >= max_pages || slot < 0 || slot >= page_size) return;
    for (int idx = warp; idx < 2 * Hkv; idx += nw) {
        const int kv = idx / Hkv, hh = idx % Hkv;
        const size_t blk = (((size_t)page * layers + layer) * 2 + kv) * Hkv + hh;
        __nv_bfloat16* row = (kv ? v : k) + ((size_t)tkn * Hkv + hh) * AT_D + 4 * lane;
        uint8_t* q = pool + blk * (size_t)page_size * AT_D + (size_t)slot * AT_D + 4 * lane;
        if (SCATTER) {
            const uint2 w = *reinterpret_cast<const uint2*>(row);
            const float x[4] = {bf16lo(w.x), bf16hi(w.x), bf16lo(w.y), bf16hi(w.y)};
            float am = fmaxf(fmaxf(fabsf(x[0]), fabsf(x[1])), fmaxf(fabsf(x[2]), fabsf(x[3])));
#pragma unroll
            for (int o = 16; o; o >>= 1) am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, o));
            const float inv = kv8_inv(am);
            uchar4 c;
            c.x = kv8_store(x[0], inv); c.y = kv8_store(x[1], inv);
            c.z = kv8_store(x[2], inv); c.w = kv8_store(x[3], inv);
            *reinterpret_cast<uchar4*>(q) = c;
            if (lane == 0) scales[blk * page_size + slot] = kv8_scale(am);
        } else {   // dequantise like the reference's get(): bf16(float(code) * scale)  (kv_cache.py:220-228)
            const uchar4 c = *reinterpret_cast<const uchar4*>(q);
            const float sc = scales[blk * page_size + slot];
            uint2 w;
            w.x = pack_bf16x2((float)((int)c.x - 128) * sc, (float)((int)c.y - 128) * sc);
            w.y = pack_bf16x2((float)((int)c.z - 128) * sc, (float)((int)c.w - 128) * sc);
            *reinterpret_cast<uint2*>(row) = w;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// KVCacheManager scatter / gather: one CTA per token, rows [Hkv*D] <-> (page, slot, layer)
template <typename T, bool SCATTER>
__global__ void kv_scatter_gather_kernel(T* pool, T* k, T* v, const int32_t* pages, const int32_t* slots, int layers, int layer,
                                         int Hkv, int page_size, int D, int max_pages) {
    const int tkn = blockIdx.x;
    const int page = pages[tkn], slot = slots[tkn];
    if (page < 0 || page >= max_pages || slot < 0 || slot >= page_size) return;
    const int HD = Hkv * D;
    for (int e = threadIdx.x; e < HD; e += blockDim.x) {
        const int hh = e / D, d = e % D;
        const size_t ok = ((((size_t)page * layers + layer) * 2 + 0) * Hkv + hh) * (size_t)page_size * D + (size_t)slot * D + d;
        const size_t ov = ok + (size_t)Hkv * page_size * D;
        if (SCATTER) { pool[ok] = k[(size_t)tkn * HD + e]; pool[ov] = v[(size_t)tkn * HD + e]; }
        else { k[(size_t)tkn * HD + e] = pool[ok]; v[(size_t)tkn * HD + e] = pool[ov]; }
    }
}

static int elt_size(int dtype) { return dtype == WF_F32 ? 4 : dtype == WF_I8 ? 1 : 2; }

int launch_rope_kv_append(const RopeParams& p, cudaStream_t s, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.H + 2 * p.Hkv, p.B);
    cfg.blockDim = dim3(p.D / 2);
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, rope_kv_append_kernel, p), "rope_kv_append");
}

// picks the chunking, launches attention (+ merge when sequences span several chunks)
int launch_paged_decode_attn(AttnParams p, int max_seq_hint, void* workspace, size_t workspace_bytes, cudaStream_t s, bool pdl) {
    static DevOnce attr;
    if (attr.need()) {
        if (cuda_check(cudaFuncSetAttribute(paged_decode_attn_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM), "attn smem")) return WF_ERR_CUDA;
        if (cuda_check(cudaFuncSetAttribute(paged_decode_attn_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM), "attn smem")) return WF_ERR_CUDA;
        if (cuda_check(cudaFuncSetAttribute(paged_decode_attn_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT8_SMEM), "attn smem")) return WF_ERR_CUDA;
        if (cuda_check(cudaFuncSetAttribute(paged_decode_attn_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT8_SMEM), "attn smem")) return WF_ERR_CUDA;
    }
    const int max_blocks16 = (max_seq_hint + AT_BLK - 1) / AT_BLK;
    // Chunking of a sequence's KV stream over CTAs (large grids; small ones take the cluster form below).  Cost model, calibrated on
    // forced chunk counts (profiles/r02ab_summary.txt, r02ac_summary.txt), in units of one 16-token block of a CTA:
    //   * J = pairs * n CTAs of length L = chunk + C0 (prologue, ring fill, 4-warp merge) run in waves of `slots` resident CTAs;
    //   * the KV stream is bandwidth-bound long before every slot is busy: R ~ 3/8 of the slots saturate HBM, so a wave of m >= R
    //     CTAs takes m * L / R, a wave of fewer takes L (each CTA then runs at its own latency-bound rate);
    //   * a split pair pays 1.4 blocks per partial for the serial merge by its last CTA.
    // Consequences the old rule (>= 4 CTAs per slot, chunks >= 512 tokens) missed: with pairs between R and `slots` ONE chunk per pair
    // is best (8B bs32/2K 4.74 -> 4.21 ms, 2B bs32/2K 2.88 -> 2.55, 2B bs48/4K 4.54 -> 4.41, 70B bs16/8K 24.7 -> 22.9), while
    // pairs just above `slots` (2B bs64: 320 on 296) need the split to dissolve the 24-CTA second wave (n = 1: 5.81 ms, n = 4: 5.28).
    // Among 3..5 chunks the model's differences there are ~3 % and the measured ones ~1 %, so the old rule's count keeps the tie.
    const long long pairs = (long long)p.B * p.Hkv;
    static const int want_mult = getenv("WF_ATTN_WANT") ? atoi(getenv("WF_ATTN_WANT")) : 0;
    static const int cm_c0 = getenv("WF_ATTN_C0") ? atoi(getenv("WF_ATTN_C0")) : 10;
    int chunk_blocks, nchunks;
    int old_cb, old_n;
    {
        const long long want = (long long)(want_mult > 0 ? want_mult : 4) * 2 * sm_count();
        int chunks = (int)((want + pairs - 1) / pairs);
        if (chunks < 1) chunks = 1;
        old_cb = (max_blocks16 + chunks - 1) / chunks;
        if (old_cb < 32) old_cb = 32;
        old_cb = (old_cb + 3) & ~3;
        old_n = (max_blocks16 + old_cb - 1) / old_cb;
        old_cb = (((max_blocks16 + old_n - 1) / old_n) + 3) & ~3;
        old_n = (max_blocks16 + old_cb - 1) / old_cb;
    }
    if (want_mult > 0) {
        chunk_blocks = old_cb; nchunks = old_n;
    } else {
        const long long slots = (long long)(p.kv_scales ? 3 : 2) * sm_count();
        // the int8 form is issue-bound, not HBM-bound: it gains from every resident CTA (R = slots).  Forced counts, profiles/r02ak_summary.txt:
        // 2B bs32/2K one chunk 2.81 ms, two 2.45; bs48/4K one 4.09, three 3.77; bs64/4K and bs64/2K one chunk (320 CTAs on 444 slots) best
        const long long R = p.kv_scales ? slots : slots * 3 / 8;
        auto model = [&](int nn, int cb) -> long long {
            const long long J = pairs * nn, L = cb + (p.kv_scales && !getenv("WF_ATTN_C0") ? 30 : cm_c0);   // int8: n = 4 at bs64/4K predicted 14 % ahead of n = 1 with C0 = 10, measured 3 % behind
            const long long full = J / slots, rem = J % slots;
            auto wave = [&](long long m) -> long long { return m >= R ? m * L * 10 / R : (m > 0 ? L * 10 : 0); };
            return full * wave(slots) + wave(rem) + (nn > 1 ? 14 * nn : 0);
        };
        long long best = -1;
        chunk_blocks = (max_blocks16 + 3) & ~3; nchunks = 1;
        for (int n = 1; n <= 64; ++n) {
            const int cb = (((max_blocks16 + n - 1) / n) + 3) & ~3;
            if (n > 1 && cb < 16) break;                                  // never below 256 tokens per CTA
            const int nn = (max_blocks16 + cb - 1) / cb;
            if (nn != n) continue;                                        // the rounding collapsed this count onto a smaller one
            const long long cost = model(nn, cb);
            if (best < 0 || cost < best) { best = cost; chunk_blocks = cb; nchunks = nn; }
        }
        if (!p.kv_scales && model(old_n, old_cb) * 100 <= best * 104) { chunk_blocks = old_cb; nchunks = old_n; }   // (a bf16 calibration)
    }
    {   // experiments: WF_ATTN_NCHUNKS forces the chunk count of the large-grid form
        static const int force_n = getenv("WF_ATTN_NCHUNKS") ? atoi(getenv("WF_ATTN_NCHUNKS")) : 0;
        if (force_n > 0) {
            chunk_blocks = (((max_blocks16 + force_n - 1) / force_n) + 3) & ~3;
            nchunks = (max_blocks16 + chunk_blocks - 1) / chunk_blocks;
        }
    }
    // Small grids (every CTA resident at once, latency-bound): the chunks of a pair form a thread-block cluster of <= 8 CTAs and
    // merge through DSMEM.  A split is cheap there (~1 us), so chunks are short (~12 blocks = 3 per warp) instead of >= 32.
    const bool fused_q = p.qkv_acc != nullptr;
    const char* cl_s = getenv("WF_ATTN_CLUSTER");
    p.cluster = 0;
    if (fused_q && (!cl_s || atoi(cl_s) != 0)) {
        // measured (profiles/r02k_summary.txt, 2B): batch 1 / 512: 1.375 -> 1.278 ms, batch 1 / 4K: 1.499 -> 1.395, batch 16 / 2K: 2.40 ->
        // 2.19; but batch 8 / 2K with 7-CTA clusters 1.87 -> 2.01 and batch 1 / 16K with 8 x 129-block chunks 1.84 -> 1.90: clusters
        // of <= 4 once the grid exceeds one CTA per SM, and only while a chunk stays short (more chunks than a cluster holds = the
        // global form, which can use every SM)
        const long long pr = pairs > 0 ? pairs : 1;
        long long nch_max = 2LL * sm_count() / pr;
        if (nch_max > 8) nch_max = 8;
        if (pr * nch_max > sm_count() && nch_max > 4) nch_max = 4;
        if (nch_max >= 2 && max_blocks16 > 12) {
            int nch = (max_blocks16 + 11) / 12;
            if (nch > nch_max) nch = (int)nch_max;
            const int cb = (((max_blocks16 + nch - 1) / nch) + 3) & ~3;
            if (cb <= 48) { chunk_blocks = cb; p.cluster = 1; }
        }
    }
    nchunks = (max_blocks16 + chunk_blocks - 1) / chunk_blocks;
    if (fused_q && cl_s && atoi(cl_s) == 2) p.cluster = 1;   // WF_ATTN_CLUSTER=2: clusters on large grids too — measured slower (5.40 vs 5.31 ms at batch 64 / 4K, profiles/r02y_summary.txt)
    if (p.cluster && (nchunks < 2 || nchunks > 8)) p.cluster = 0;
    p.chunk_blocks = chunk_blocks;
    p.max_chunks = nchunks;
    p.page_shift = -1;
    if ((p.page_size & (p.page_size - 1)) == 0) { p.page_shift = 0; while ((1 << p.page_shift) < p.page_size) ++p.page_shift; }
    p.direct = (nchunks <= 1);
    // Even work split (see the kernel): the fused decode form with at least 512 tokens of KV per CTA of one resident wave.
    const bool fused_form = p.qkv_acc != nullptr && p.merge_cnt != nullptr;
    const bool i8_pool = p.kv_scales != nullptr;
    const char* even_s = getenv("WF_ATTN_EVEN");   // read per launch (launches are graph-captured): tests switch it inside one process
    const int even_env = even_s ? atoi(even_s) : 0;   // opt-in: measured 5.65 vs 5.46 ms per bs64/4K step (profiles/r02i_*): the per-segment
                                                       // prologue / merge of a CTA costs what the missing partial wave saves
    { const char* cm = getenv("WF_ATTN_CHUNK_MAJOR"); p.chunk_major = cm ? atoi(cm) : 0; }   // measured: chunk-major dispatch 5.71 vs 5.47 ms (bs64/4K)
    int even_ctas = 0;
    p.even = 0;
    if (fused_form && even_env && !i8_pool && p.B <= AT_EVEN_MAXB) {
        long long nc = pairs * max_blocks16 / 32;
        const long long slots = (long long)(i8_pool ? 3 : 2) * sm_count();
        if (nc > slots) nc = slots;
        const size_t need = (size_t)nc * 2 * 8 * (AT_D + 2) * sizeof(float);
        if (nc >= sm_count() && workspace && workspace_bytes >= need) {
            p.even = 1; p.direct = 0; even_ctas = (int)nc;
            p.part_o = (float*)workspace;
            p.part_ml = p.part_o + (size_t)nc * 2 * 8 * AT_D;
        }
    }
    if (p.even) p.cluster = 0;
    // Two-class grid (see the kernel): pairs = q * slots + r with a thin remainder r < R.
    p.two_pa = 0; p.two_nb = 0;
    { static const int sf = getenv("WF_ATTN_TWO_ORDER") ? atoi(getenv("WF_ATTN_TWO_ORDER")) : 0; p.two_short_first = sf; }
    if (i8_pool) { p.chunk_major = 0; p.two_short_first = 0; }   // experiment orders: not compiled into the int8 kernel (register budget)
    {
        const char* two_s = getenv("WF_ATTN_TWO");   // read per launch: the parity test switches it inside one process
        const int two_env = two_s ? atoi(two_s) : 1;
        const long long slots = (long long)(i8_pool ? 3 : 2) * sm_count();
        const long long q = pairs / slots, r = pairs % slots;
        if (two_env && fused_form && !p.even && !p.cluster && q >= 1 && r > 0 && r < slots * 3 / 8 && max_blocks16 >= 64) {
            static const int two_fill = getenv("WF_ATTN_TWO_FILL") ? atoi(getenv("WF_ATTN_TWO_FILL")) : 4;   // eighths of the slots the short wave fills: measured 8: 5.22, 6: 5.14, 4: 5.07, 3: 5.09 ms (bs64/4K; uniform grid 5.29)
            int nb2 = (int)((slots * two_fill / 8 + r - 1) / r);
            if (nb2 > max_blocks16 / 16) nb2 = max_blocks16 / 16;        // >= 256 tokens per CTA
            if (nb2 > 16) nb2 = 16;
            if (nb2 >= 2) {
                chunk_blocks = (((max_blocks16 + nb2 - 1) / nb2) + 3) & ~3;
                nchunks = (max_blocks16 + chunk_blocks - 1) / chunk_blocks;
                p.chunk_blocks = chunk_blocks; p.max_chunks = nchunks; p.direct = 0;
                p.two_pa = (int)(q * slots); p.two_nb = nchunks;
            }
        }
    }
    {   // per-sequence chunking (WF_ATTN_ADAPT=0: fixed chunks of the hint, the round-2 parity twin); read per launch like WF_ATTN_TWO
        const char* ad = getenv("WF_ATTN_ADAPT");
        p.min_cb = (ad && atoi(ad) == 0) ? 0 : (p.cluster ? 4 : 16);
    }
    if (!p.direct && !p.even && !p.cluster) {
        const size_t need = (size_t)p.B * p.H * nchunks * (AT_D + 2) * sizeof(float);
        if (!workspace || workspace_bytes < need) { set_error("paged_decode_attn: workspace too small (%zu < %zu)", workspace_bytes, need); return WF_ERR_INVALID; }
        p.part_o = (float*)workspace;
        p.part_ml = p.part_o + (size_t)p.B * p.H * nchunks * AT_D;
    }
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = nchunks < 1 ? 1 : nchunks; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    {
        cudaLaunchConfig_t cfg = {};
        const int nchg = nchunks < 1 ? 1 : nchunks;
        cfg.gridDim = p.even ? dim3(even_ctas) : p.two_nb > 0 ? dim3((unsigned)(p.two_pa + (pairs - p.two_pa) * p.two_nb))
                             : (p.chunk_major ? dim3(p.Hkv, p.B, nchg) : dim3(nchg, p.Hkv, p.B));
        cfg.blockDim = dim3(AT_WARPS * 32);
        const bool i8 = p.kv_scales != nullptr;
        cfg.dynamicSmemBytes = i8 ? AT8_SMEM : AT_SMEM;
        cfg.stream = s;
        cudaLaunchAttribute la[2];
        int na = 0;
        if (pdl) la[na++] = at[0];
        if (p.cluster) la[na++] = at[1];
        cfg.attrs = la;
        cfg.numAttrs = na;
        const bool fused = p.qkv_acc != nullptr;
        cudaError_t le;
        if (i8) le = fused ? cudaLaunchKernelEx(&cfg, paged_decode_attn_kernel<true, true>, p) : cudaLaunchKernelEx(&cfg, paged_decode_attn_kernel<false, true>, p);
        else le = fused ? cudaLaunchKernelEx(&cfg, paged_decode_attn_kernel<true, false>, p) : cudaLaunchKernelEx(&cfg, paged_decode_attn_kernel<false, false>, p);
        if (launch_check(le, "paged_decode_attn")) return WF_ERR_CUDA;
        if (fused) return WF_OK;
    }
    if (!p.direct) {
        MergeParams mp;
        mp.part_o = p.part_o; mp.part_ml = p.part_ml; mp.seq_lens = p.seq_lens;
        mp.H = p.H; mp.max_chunks = p.max_chunks; mp.chunk_blocks = p.chunk_blocks; mp.min_cb = p.min_cb;
        mp.out = p.out; mp.out_dtype = p.out_dtype;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(p.H, p.B);
        cfg.blockDim = dim3(AT_D);
        cfg.stream = s;
        cfg.attrs = at;
        cfg.numAttrs = pdl ? 1 : 0;
        if (launch_check(cudaLaunchKernelEx(&cfg, attn_merge_kernel, mp), "attn_merge")) return WF_ERR_CUDA;
    }
    return WF_OK;
}


int rope_append(const void* qkv, int B, int H, int Hkv, int D, void* pool, int layers, int layer, int page_size,
                const int32_t* positions, const int32_t* block_table, int max_blocks, const void* cos_tab, const void* sin_tab,
                int max_pos, void* q_out, cudaStream_t s, bool pdl, float* kv_scales) {
    RopeParams p;
    p.kv_scales = kv_scales;
    p.qkv = (const __nv_bfloat16*)qkv; p.B = B; p.H = H; p.Hkv = Hkv; p.D = D;
    p.pool = (__nv_bfloat16*)pool; p.layers = layers; p.layer = layer; p.page_size = page_size; p.max_blocks = max_blocks;
    p.positions = positions; p.block_table = block_table;
    p.cos_tab = (const __nv_bfloat16*)cos_tab; p.sin_tab = (const __nv_bfloat16*)sin_tab; p.max_pos = max_pos;
    p.q_out = (__nv_bfloat16*)q_out;
    return launch_rope_kv_append(p, s, pdl);
}

int attn_decode(const void* q, const void* pool, int layers, int layer, int Hkv, int page_size, const int32_t* block_table,
                int max_blocks, const int32_t* seq_lens, int B, int H, float scale, void* out, int out_dtype, int max_seq_hint,
                void* ws, size_t ws_bytes, cudaStream_t s, bool pdl, float* kv_scales) {
    AttnParams p = {};
    p.kv_scales = kv_scales;
    p.q = (const __nv_bfloat16*)q; p.pool = (const __nv_bfloat16*)pool;
    p.block_table = block_table; p.seq_lens = seq_lens;
    p.B = B; p.H = H; p.Hkv = Hkv; p.G = H / Hkv;
    p.layers = layers; p.layer = layer; p.page_size = page_size; p.max_blocks = max_blocks;
    p.scale_log2e = scale * 1.4426950408889634f;
    p.out = out; p.out_dtype = out_dtype;
    p.n_out = 1; p.out_ld = H * AT_D; p.out_col0 = 0; p.signal_flag = -1;
    return launch_paged_decode_attn(p, max_seq_hint, ws, ws_bytes, s, pdl);
}

// decode-engine form: q/k/v are dequantised from the qkv linear's raw accumulator inside the kernel (RoPE, KV append,
// attention and the chunk merge are one launch)
int attn_decode_fused(const FusedAttnArgs& a, cudaStream_t s, bool pdl) {
    AttnParams p = {};
    p.pool = (const __nv_bfloat16*)a.pool; p.pool_w = (__nv_bfloat16*)a.pool;
    p.block_table = a.block_table; p.seq_lens = a.seq_lens;
    p.B = a.B; p.H = a.H; p.Hkv = a.Hkv; p.G = a.H / a.Hkv;
    p.layers = a.layers; p.layer = a.layer; p.page_size = a.page_size; p.max_blocks = a.max_blocks;
    p.scale_log2e = a.scale * 1.4426950408889634f;
    p.out = a.out; p.out_dtype = WF_BF16;
    p.qkv_acc = a.qkv_acc; p.qkv_ld = a.qkv_ld; p.qkv_rowsum = a.qkv_rowsum; p.qkv_ascale = a.qkv_ascale; p.qkv_wscale = a.qkv_wscale;
    p.positions = a.positions; p.cos_tab = (const __nv_bfloat16*)a.cos_tab; p.sin_tab = (const __nv_bfloat16*)a.sin_tab; p.max_pos = a.max_pos;
    p.merge_cnt = a.merge_cnt; p.zero0 = a.zero0; p.zero0_n = a.zero0_n;
    p.n_out = a.n_out < 1 ? 1 : a.n_out; p.out_ld = a.out_ld > 0 ? a.out_ld : a.H * AT_D; p.out_col0 = a.out_col0;
    for (int r = 0; r < TP_MAX; ++r) p.out_peer[r] = a.out_peer[r];
    p.signal_flag = a.n_out > 1 ? a.signal_flag : -1; p.tp = a.tp; p.trace = a.trace;
    p.kv_scales = a.kv_scales;
    return launch_paged_decode_attn(p, a.max_seq_hint, a.ws, a.ws_bytes, s, pdl);
}

}  // namespace wf

using namespace wf;

extern "C" size_t wf_kv_pool_bytes(int max_pages, int layers, int kv_heads, int page_size, int head_dim, int dtype) {
    const size_t elems = (size_t)max_pages * layers * 2 * kv_heads * page_size * head_dim;
    if (dtype == WF_I8) return elems + elems / head_dim * sizeof(float);   // int8 region + one fp32 scale per (token, head)
    return elems * elt_size(dtype);
}

// scale region of an int8 pool
static float* kv_scales_of(const wf_kv_layout* kv) {
    if (kv->dtype != WF_I8) return nullptr;
    return reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(kv->pool) +
                                    (size_t)kv->max_pages * kv->layers * 2 * kv->kv_heads * kv->page_size * kv->head_dim);
}

static int check_kv(const wf_kv_layout* kv, const char* who) {
    if (!kv || !kv->pool) { set_error("%s: null kv layout/pool", who); return WF_ERR_INVALID; }
    if (kv->max_pages <= 0 || kv->layers <= 0 || kv->kv_heads <= 0 || kv->page_size <= 0 || kv->head_dim <= 0) { set_error("%s: bad kv layout", who); return WF_ERR_INVALID; }
    return WF_OK;
}

template <bool SC>
static int scatter_gather(const wf_kv_layout* kv, void* k, void* v, const int32_t* pages, const int32_t* slots, int T, int layer, void* stream) {
    int rc = check_kv(kv, SC ? "wf_kv_scatter" : "wf_kv_gather");
    if (rc) return rc;
    if (!k || !v || !pages || !slots) { set_error("kv scatter/gather: null pointer"); return WF_ERR_INVALID; }
    if (layer < 0 || layer >= kv->layers) { set_error("kv scatter/gather: layer %d out of range", layer); return WF_ERR_INVALID; }
    if (T <= 0) return WF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (kv->dtype == WF_F32)
        kv_scatter_gather_kernel<float, SC><<<T, 256, 0, s>>>((float*)kv->pool, (float*)k, (float*)v, pages, slots, kv->layers, layer, kv->kv_heads, kv->page_size, kv->head_dim, kv->max_pages);
    else if (kv->dtype == WF_BF16)
        kv_scatter_gather_kernel<__nv_bfloat16, SC><<<T, 256, 0, s>>>((__nv_bfloat16*)kv->pool, (__nv_bfloat16*)k, (__nv_bfloat16*)v, pages, slots, kv->layers, layer, kv->kv_heads, kv->page_size, kv->head_dim, kv->max_pages);
    else if (kv->dtype == WF_I8) {   // k / v rows are bf16; quantised per (token, head) on the way in, dequantised on the way out
        if (kv->head_dim != AT_D) { set_error("kv scatter/gather: int8 pools need head_dim 128"); return WF_ERR_UNSUPPORTED; }
        kv_scatter_gather_i8_kernel<SC><<<T, 128, 0, s>>>((uint8_t*)kv->pool, kv_scales_of(kv), (__nv_bfloat16*)k, (__nv_bfloat16*)v, pages, slots, kv->layers, layer, kv->kv_heads, kv->page_size, kv->max_pages);
    } else { set_error("kv scatter/gather: dtype must be f32, bf16 or int8"); return WF_ERR_INVALID; }
    return launch_check(cudaGetLastError(), "kv scatter/gather");
}

extern "C" int wf_kv_scatter(const wf_kv_layout* kv, const void* k, const void* v, const int32_t* pages, const int32_t* slots, int T, int layer, void* stream) {
    return scatter_gather<true>(kv, (void*)k, (void*)v, pages, slots, T, layer, stream);
}
extern "C" int wf_kv_gather(const wf_kv_layout* kv, void* k_out, void* v_out, const int32_t* pages, const int32_t* slots, int T, int layer, void* stream) {
    return scatter_gather<false>(kv, k_out, v_out, pages, slots, T, layer, stream);
}

extern "C" int wf_rope_kv_append(const void* qkv, int B, int H, const wf_kv_layout* kv, int layer, const int32_t* positions,
                                 const int32_t* block_table, int max_blocks, const void* cos_tab, const void* sin_tab, int max_pos,
                                 void* q_out, void* stream) {
    int rc = check_kv(kv, "wf_rope_kv_append");
    if (rc) return rc;
    if (!qkv || !positions || !block_table || !cos_tab || !sin_tab || !q_out) { set_error("wf_rope_kv_append: null pointer"); return WF_ERR_INVALID; }
    if (kv->dtype != WF_BF16 && kv->dtype != WF_I8) { set_error("wf_rope_kv_append: pool must be bf16 or int8"); return WF_ERR_UNSUPPORTED; }
    if (kv->head_dim % 2 || kv->head_dim > 2048) { set_error("wf_rope_kv_append: bad head_dim"); return WF_ERR_UNSUPPORTED; }
    if (kv->dtype == WF_I8 && kv->head_dim != AT_D) { set_error("wf_rope_kv_append: int8 pools need head_dim 128"); return WF_ERR_UNSUPPORTED; }
    if (B <= 0) return WF_OK;
    return rope_append(qkv, B, H, kv->kv_heads, kv->head_dim, kv->pool, kv->layers, layer, kv->page_size, positions, block_table,
                       max_blocks, cos_tab, sin_tab, max_pos, q_out, (cudaStream_t)stream, false, kv_scales_of(kv));
}

extern "C" size_t wf_attn_workspace_bytes(int B, int H, int head_dim, int max_seq) {
    const size_t chunks = (size_t)(max_seq + 63) / 64 + 1;
    const size_t legacy = (size_t)B * H * chunks * (head_dim + 2) * sizeof(float);
    // even work split: two partial slots of 8 heads per CTA of one resident wave (<= 3 CTAs on each of <= 160 SMs)
    const size_t even = (size_t)3 * 160 * 2 * 8 * (head_dim + 2) * sizeof(float);
    return legacy > even ? legacy : even;
}

extern "C" int wf_paged_decode_attn(const void* q, const wf_kv_layout* kv, int layer, const int32_t* block_table, int max_blocks,
                                    const int32_t* seq_lens, int B, int H, float scale, void* out, int out_dtype, void* workspace,
                                    size_t workspace_bytes, void* stream) {
    int rc = check_kv(kv, "wf_paged_decode_attn");
    if (rc) return rc;
    if (!q || !block_table || !seq_lens || !out) { set_error("wf_paged_decode_attn: null pointer"); return WF_ERR_INVALID; }
    if (kv->dtype != WF_BF16 && kv->dtype != WF_I8) { set_error("wf_paged_decode_attn: pool must be bf16 or int8"); return WF_ERR_UNSUPPORTED; }
    if (kv->head_dim != AT_D) { set_error("wf_paged_decode_attn: head_dim must be 128 (got %d)", kv->head_dim); return WF_ERR_UNSUPPORTED; }
    if (kv->page_size % AT_BLK) { set_error("wf_paged_decode_attn: page_size must be a multiple of 16 (got %d)", kv->page_size); return WF_ERR_UNSUPPORTED; }
    if (H % kv->kv_heads || H / kv->kv_heads > 8) { set_error("wf_paged_decode_attn: H/Hkv must be an integer <= 8"); return WF_ERR_UNSUPPORTED; }
    if (layer < 0 || layer >= kv->layers) { set_error("wf_paged_decode_attn: layer out of range"); return WF_ERR_INVALID; }
    if (out_dtype != WF_BF16 && out_dtype != WF_F32) { set_error("wf_paged_decode_attn: out dtype"); return WF_ERR_INVALID; }
    if (B <= 0) return WF_OK;
    return attn_decode(q, kv->pool, kv->layers, layer, kv->kv_heads, kv->page_size, block_table, max_blocks, seq_lens, B, H, scale, out,
                       out_dtype, max_blocks * kv->page_size, workspace, workspace_bytes, (cudaStream_t)stream, false, kv_scales_of(kv));
}

namespace wf {
void attn_init_attrs() {
    cudaFuncSetAttribute(paged_decode_attn_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
    cudaFuncSetAttribute(paged_decode_attn_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
    cudaFuncSetAttribute(paged_decode_attn_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT8_SMEM);
    cudaFuncSetAttribute(paged_decode_attn_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT8_SMEM);
}
}
