// Streaming batched ternary GEMM for the decode step, 1..32 activation rows (the engine uses it up to 16 rows, 24 where K < 8192, and for batch 1 at the 8B / 70B widths) (bitnet_gemm at decode batch sizes; the
// engine's linears on the 8B / 70B matrices, DESIGN 7 gap 3).
//
//   acc[m, n] += sum_k code[n, k] * a[m, k]        exact integers in fp32 (|S| < 2^24), RED into the step's split-K accumulators
//   (reference: one ggml_vec_dot_i2_i8_s per output element, extern/BitNet/src/ggml-bitnet-mad.cpp:94-186)
//
// The one-shot GEMM (gemm_oneshot.cu) and the tcgen05 split-K form (gemm_tc.cu) are sized for the 2B matrices: on 30-230 MB matrices they
// move the packed weights at 9-15 % of the HBM peak (one-shot: a RED round per 4 k-blocks; tcgen05: the 2-bit -> u8 unpack costs 7 ALU
// instructions per packed word, worth it only when 128 tokens share it).  The GEMV kernel (gemv_stream.cu) streams the same matrices at
// 54-86 % but keeps all M activation rows of the whole K in shared memory (M <= 8, K * M <= ~150 KB).  This kernel is its batched form:
//   * K is cut into nwin windows of KW <= 32 k-blocks (4096 k): only the M x (KW * 128) activation window is resident (32 rows x 4096 =
//     128 KB), and in b200_tile_v1 the KW tiles of one 16-row block inside a window are ONE contiguous run of KW * 512 <= 16 KB bytes:
//     a unit = (window, row-block) = one cp.async.bulk = one ring stage;
//   * persistent grid; units are dealt in window-major order, a contiguous range per CTA (a CTA changes window at most a few times); the
//     producer lane never waits on the previous kernel (programmatic dependent launch: the weight stream starts under the row glue);
//   * 8 consumer warps, warp j takes k-blocks j, j + 8, ... of the unit.  A lane's 16-byte LDS is its four mma A-fragment words; the 2-bit
//     fields are masked IN PLACE (one AND per fragment, four k-groups -> four accumulators, shifted back at the flush: exact), activations
//     come through ldmatrix from a skewed copy, M/8 n-tiles of mma.sync.m16n8k32.u8.s8 share every weight fragment;
//   * per unit: warps add their partial sums into a double-buffered shared tile laid out [n-tile][fragment register][lane] (every lane its own
//     bank), one consumer barrier, 4 * M lanes (a duty that rotates over the warps) issue red.global.add.v4.f32 into the accumulator (or, tensor parallel, into the owner rank's / every rank's accumulator, as the one-shot GEMM).
#include "fused_common.cuh"

namespace wf {

constexpr int GS_KW_MAX = 32;                  // k-blocks per window: a stage is at most 16 KB
constexpr int GS_STAGE_BYTES = GS_KW_MAX * WF_TILE_BYTES;
constexpr int GS_MAX_STAGES = 8;

struct GsParams {
    const uint8_t* w;
    const int8_t* q;
    int M, N, ldq, k_off;
    int KB, RB, KW, nwin;
    int U;                // units = nwin * RB
    int act_stride;       // bytes per activation row of a window in shared memory (KW * 128 + 16: rows land 16 B apart mod 128)
    int stages;
    TpDest dest;
    int acc_ld;
    TpComm tp;
};

// NTL n-tiles of 8 activation rows per pass, NP passes over the same weight fragments (M <= 8 * NTL * NP), CW consumer warps.
// <1,1,8> / <2,1,8>: up to 8 / 16 rows, two CTAs per SM.  <2,2,16>: 17..32 rows — the 128 KB activation window leaves room for one CTA
// per SM, so that CTA brings 16 consumer warps, and two passes of two n-tiles keep the accumulators at 32 registers (four n-tiles at once:
// 168 registers, 8 warps per SM, measured slower than the one-shot GEMM).
template <int NTL, int NP, int CW>
__global__ void __launch_bounds__((CW + 1) * 32, (CW == 8 && NTL <= 2) ? 2 : 1) gemm_stream_kernel(const GsParams p) {
    constexpr int GS_CW = CW, GS_NT = (CW + 1) * 32, GS_CT = CW * 32, TPW = GS_KW_MAX / CW, NTT = NTL * NP;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;
    unsigned char* acts = ring + p.stages * GS_STAGE_BYTES;
    int* sacc = reinterpret_cast<int*>(acts + ((p.M * p.act_stride + 127) & ~127));   // [2][8 * NTT][16]
    __shared__ __align__(8) uint64_t full_bar[GS_MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[GS_MAX_STAGES];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int u0 = (int)((long long)blockIdx.x * p.U / gridDim.x), u1 = (int)((long long)(blockIdx.x + 1) * p.U / gridDim.x);
    const int S = p.stages;
    const uint32_t unit_bytes = (uint32_t)p.KW * WF_TILE_BYTES;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], GS_CW); }
        fence_mbar_init();
    }
    pdl_launch_dependents();   // dependents still wait for this grid to finish; this only lets their weight streams start
    for (int i = tid; i < 2 * 8 * NTT * 16; i += GS_NT) sacc[i] = 0;
    __syncthreads();

    if (warp == GS_CW) {
        // ------------------------------------------------ producer (one lane): independent of the previous kernel
        if (lane == 0) {
            int win = u0 / p.RB, rb = u0 - win * p.RB;
            int s = 0, ph = 1;
            for (int u = u0; u < u1; ++u) {
                mbar_wait(&empty_bar[s], ph);
                mbar_expect_tx(&full_bar[s], unit_bytes);
                bulk_g2s(ring + s * GS_STAGE_BYTES, p.w + ((size_t)rb * p.KB + (size_t)win * p.KW) * WF_TILE_BYTES, unit_bytes, &full_bar[s]);
                if (++s == S) { s = 0; ph ^= 1; }
                if (++rb == p.RB) { rb = 0; ++win; }
            }
        }
    } else {
        // ------------------------------------------------ consumers
        pdl_wait_prior();      // the int8 activations are written by the row glue in front of this kernel
        int acc[NTL][16];
#pragma unroll
        for (int i = 0; i < NTL; ++i)
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[i][j] = 0;
        const unsigned char* abase[NTT];
#pragma unroll
        for (int i = 0; i < NTT; ++i) abase[i] = acts + min(8 * i + (lane & 7), p.M - 1) * p.act_stride + (lane >> 3) * 16;
        const int g = lane >> 2, t = lane & 3;
        int win = u0 / p.RB, rb = u0 - win * p.RB;
        int cur_win = -1;
        int s = 0, ph = 0, par = 0;
        for (int u = u0; u < u1; ++u) {
            if (win != cur_win) {
                // a new K window: everybody is done with the old activations (the barrier at the end of the previous unit), stage the new ones
                const int gran = p.KW * 8;                       // 16-byte granules per row
                const int8_t* src = p.q + p.k_off + (size_t)win * p.KW * WF_QK;
                for (int i = tid; i < p.M * gran; i += GS_CT) {
                    const int m = i / gran, c = i - m * gran;
                    cp_async16(acts + m * p.act_stride + c * 16, src + (size_t)m * p.ldq + c * 16);
                }
                cp_async_commit();
                cp_async_wait<0>();
                asm volatile("bar.sync 1, %0;" ::"n"(GS_CT) : "memory");
                cur_win = win;
            }
            mbar_wait(&full_bar[s], ph);
            int* dst = sacc + par * (8 * NTT * 16);
            {
                const unsigned char* wp = ring + s * GS_STAGE_BYTES + lane * 16;
                uint4 W[TPW];
#pragma unroll
                for (int j = 0; j < TPW; ++j) {
                    const int kb = warp + GS_CW * j;
                    if (kb < p.KW) W[j] = *reinterpret_cast<const uint4*>(wp + kb * WF_TILE_BYTES);
                }
                __syncwarp();
                if (NP == 1 && lane == 0) mbar_arrive(&empty_bar[s]);   // the stage's bytes are in registers
#pragma unroll
                for (int pp = 0; pp < NP; ++pp) {
#pragma unroll
                    for (int j = 0; j < TPW; ++j) {
                        const int kb = warp + GS_CW * j;
                        if (kb < p.KW) {
#pragma unroll
                            for (int i = 0; i < NTL; ++i) {
                                uint32_t b[8];
                                const unsigned char* ap = abase[pp * NTL + i] + kb * 128;
                                ldmatrix_x4(b[0], b[1], b[2], b[3], ap);
                                ldmatrix_x4(b[4], b[5], b[6], b[7], ap + 64);
#pragma unroll
                                for (int kg = 0; kg < 4; ++kg) {
                                    const uint32_t mask = 0x03030303u << (6 - 2 * kg);   // k-group kg sits in bits (7-2kg .. 6-2kg)
                                    int(&d)[4] = *reinterpret_cast<int(*)[4]>(&acc[i][kg * 4]);
                                    mma_u8s8_16832(d, W[j].x & mask, W[j].y & mask, W[j].z & mask, W[j].w & mask, b[2 * kg], b[2 * kg + 1]);
                                }
                            }
                        }
                    }
                    // ---- the warp's partial sums of this pass -> shared [m][16] (weight row g / g + 8, activation rows 8i + 2t, + 1)
#pragma unroll
                    for (int i = 0; i < NTL; ++i) {
                        int v[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            v[j] = (acc[i][0 + j] >> 6) + (acc[i][4 + j] >> 4) + (acc[i][8 + j] >> 2) + acc[i][12 + j];
#pragma unroll
                            for (int kg = 0; kg < 4; ++kg) acc[i][kg * 4 + j] = 0;
                        }
                        // shared tile [n-tile][j][lane]: every lane adds into its own bank (the eight warps hold the same (row, token) in the same lane)
                        int* dl = dst + ((pp * NTL + i) * 4) * 32 + lane;
                        const int m = 8 * (pp * NTL + i) + 2 * t;
                        if (m < p.M) {
                            if (v[0]) atomicAdd(dl, v[0]);
                            if (v[2]) atomicAdd(dl + 64, v[2]);
                        }
                        if (m + 1 < p.M) {
                            if (v[1]) atomicAdd(dl + 32, v[1]);
                            if (v[3]) atomicAdd(dl + 96, v[3]);
                        }
                    }
                }
                if (NP > 1) { __syncwarp(); if (lane == 0) mbar_arrive(&empty_bar[s]); }
            }
            if (++s == S) { s = 0; ph ^= 1; }
            asm volatile("bar.sync 1, %0;" ::"n"(GS_CT) : "memory");
            // ---- exact fp32 RED of the unit's [M][16] tile (and the tile is left zero for the unit after next)
            // (the duty rotates over the warps from unit to unit: the lanes that issue the REDs fall behind the others by that much)
            const int rt = (tid - ((u - u0) & 3) * (GS_CT / 4)) & (GS_CT - 1);
            if (rt < p.M * 4) {
                const int m = rt >> 2, c = rt & 3;                     // activation row m, weight rows 4c .. 4c + 3 of the block
                // element (row r, token m) sits at [n-tile m / 8][j = 2 * (r / 8) + (m & 1)][lane = 4 * (r & 7) + (m & 7) / 2]
                int* sp = dst + ((m >> 3) * 4 + 2 * (c >> 1) + (m & 1)) * 32 + 4 * (4 * (c & 1)) + ((m & 7) >> 1);
                const int vx = sp[0], vy = sp[4], vz = sp[8], vw = sp[12];
                const int n = rb * 16 + 4 * c;
                if ((vx | vy | vz | vw) != 0) {
                    sp[0] = 0; sp[4] = 0; sp[8] = 0; sp[12] = 0;
                    if (n < p.N) {
                        const float fx = (float)vx, fy = (float)vy, fz = (float)vz, fw = (float)vw;
                        if (p.dest.own_cols > 0) red_add_v4_f32(p.dest.acc[n / p.dest.own_cols] + (size_t)m * p.acc_ld + n, fx, fy, fz, fw);
                        else
                            for (int d = 0; d < p.dest.n; ++d) red_add_v4_f32(p.dest.acc[d] + (size_t)m * p.acc_ld + n, fx, fy, fz, fw);
                    }
                }
            }
            par ^= 1;
            if (++rb == p.RB) { rb = 0; ++win; }
        }
    }
    if (p.dest.signal_flag >= 0) tp_signal_grid_done(p.tp, p.dest.signal_flag, gridDim.x);
}

// ---------------------------------------------------------------------------------------------- host
static size_t g_gs_attr_smem = 0;
void gemm_stream_init_attrs() {
    bool ok = true;
    ok &= cudaFuncSetAttribute(gemm_stream_kernel<1, 1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) == cudaSuccess;
    ok &= cudaFuncSetAttribute(gemm_stream_kernel<2, 1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) == cudaSuccess;
    ok &= cudaFuncSetAttribute(gemm_stream_kernel<2, 1, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) == cudaSuccess;
    ok &= cudaFuncSetAttribute(gemm_stream_kernel<2, 2, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) == cudaSuccess;
    if (ok) g_gs_attr_smem = 220 * 1024;
}

// window split: the fewest windows of equal length <= GS_KW_MAX k-blocks
static int gs_windows(int KB) {
    int nwin = (KB + GS_KW_MAX - 1) / GS_KW_MAX;
    while (KB % nwin) ++nwin;
    return nwin;
}

bool gemm_stream_fits(int M, int N, int K) {
    if (M < 1 || M > 32 || K % WF_QK || N % 4) return false;
    const int KB = K / WF_QK, KW = KB / gs_windows(KB);
    if (KW < 8) return false;                                    // (a prime KB: windows of a few k-blocks would starve the 8 warps)
    const size_t acts = ((size_t)M * (KW * 128 + 16) + 127) & ~(size_t)127;
    return 3 * (size_t)GS_STAGE_BYTES + acts + 2 * 32 * 16 * 4 <= 220 * 1024;
}

int launch_gemm_stream(const OneshotArgs& a, cudaStream_t s, bool pdl) {
    if (!gemm_stream_fits(a.M, a.N, a.K)) { set_error("gemm_stream: M=%d N=%d K=%d not supported", a.M, a.N, a.K); return WF_ERR_UNSUPPORTED; }
    if (a.ldq % 16 || a.k_off % WF_QK || a.acc_ld % 4) { set_error("gemm_stream: ldq %% 16, k_off %% 128, acc_ld %% 4 must be 0"); return WF_ERR_INVALID; }
    GsParams p;
    p.w = a.w; p.q = a.q; p.M = a.M; p.N = a.N; p.ldq = a.ldq; p.k_off = a.k_off;
    p.KB = a.K / WF_QK;
    p.RB = (a.N + 15) / 16;
    p.nwin = gs_windows(p.KB);
    p.KW = p.KB / p.nwin;
    p.U = p.nwin * p.RB;
    p.act_stride = p.KW * 128 + 16;
    p.dest = a.dest; p.acc_ld = a.acc_ld; p.tp = a.tp;
    const int ntl = a.M <= 8 ? 1 : (a.M <= 16 ? 2 : 4);
    const size_t acts = ((size_t)a.M * p.act_stride + 127) & ~(size_t)127;
    const size_t fixed = acts + (size_t)2 * 8 * ntl * 16 * 4;
    // two CTAs per SM (two independent TMA streams) while both fit with >= 3 stages each; one CTA with a deeper ring otherwise
    int cps = 1;
    p.stages = GS_MAX_STAGES;
    if (ntl <= 2 && 2 * (3 * (size_t)GS_STAGE_BYTES + fixed + 1024) <= 226 * 1024) {
        cps = 2;
        while (p.stages > 3 && 2 * (p.stages * (size_t)GS_STAGE_BYTES + fixed + 1024) > 226 * 1024) --p.stages;
    } else {
        while (p.stages > 3 && p.stages * (size_t)GS_STAGE_BYTES + fixed > 220 * 1024) --p.stages;
    }
    const size_t smem = p.stages * (size_t)GS_STAGE_BYTES + fixed;
    if (smem > g_gs_attr_smem) {
        gemm_stream_init_attrs();
        if (smem > g_gs_attr_smem) { set_error("gemm_stream: cannot opt in to %zu bytes of shared memory", smem); return WF_ERR_CUDA; }
    }
    int grid = sm_count() * cps;
    if (grid > p.U) grid = p.U;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    // 9..16 rows whose activation window leaves room for only one CTA per SM (K windows of >= 3.5K at 16 rows: the 8B / 70B widths): that CTA
    // brings 16 consumer warps instead of 8
    static const int wide_env = getenv("WF_GS_WIDE") ? atoi(getenv("WF_GS_WIDE")) : 0;   // measured slower: 8B bs16 3.19 vs 3.05 ms, 70B bs16 20.3 vs 19.9 (profiles/r02au_summary.txt)
    const bool wide = ntl == 2 && cps == 1 && wide_env;
    cfg.blockDim = dim3((ntl == 4 || wide) ? 17 * 32 : 9 * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e;
    if (ntl == 1) e = cudaLaunchKernelEx(&cfg, gemm_stream_kernel<1, 1, 8>, p);
    else if (ntl == 2 && wide) e = cudaLaunchKernelEx(&cfg, gemm_stream_kernel<2, 1, 16>, p);
    else if (ntl == 2) e = cudaLaunchKernelEx(&cfg, gemm_stream_kernel<2, 1, 8>, p);
    else e = cudaLaunchKernelEx(&cfg, gemm_stream_kernel<2, 2, 16>, p);
    return launch_check(e, "gemm_stream");
}

}  // namespace wf
