// Batched-decode ternary GEMM (8 < M <= 64): "one-shot" split tiles, exact fp32 RED accumulation.
//
//   S[m,n] += sum_{k in slice} code[n,k] * a[m,k]     (reference: one ggml_vec_dot_i2_i8_s per (row, token),
//   extern/BitNet/src/ggml-bitnet-mad.cpp:94-186, under bitnet_gemm's OpenMP loop)
//
// At decode batch sizes every 2B-8B matrix is smaller than what the chip keeps in flight (1.6-29 MB vs
// ~6.5 MB of outstanding loads), so a K-loop pipeline only adds latency.  Instead the (rows x K) plane is cut
// into (64*RBW rows) x (KC k-blocks) units, one CTA each, 3-4 CTAs per SM: a CTA issues ALL its loads at once
// (weights before griddepcontrol.wait — they do not depend on the previous kernel — activations after),
// computes with mma.sync.m16n8k32.u8.s8 and adds its partial sums into the global fp32 accumulator with
// red.global.add.v4.f32 (integers < 2^24: exact, order-independent).  The consumer of the accumulator
// (glue_rows_kernel / attention prologue) applies scales, residual, norm and re-quantisation.
#include "fused_common.cuh"
#include <type_traits>

namespace wf {

constexpr int GO_T = 256;
constexpr int GO_MT = 64;   // batch columns per CTA

struct OneshotParams {
    const uint8_t* w;
    const int8_t* q;
    int M, N, K, KB, RB, KC;
    int ldq, k_off;
    TpDest dest;
    int acc_ld;
    TpComm tp;
};

template <int RBW, bool FULLM>   // FULLM: M is a multiple of 32, every active warp owns four live 8-token tiles
__global__ void __launch_bounds__(GO_T) gemm_oneshot_kernel(const OneshotParams p) {
    constexpr int RBS = 4 * RBW;        // row-blocks per CTA
    constexpr int RT = 16 * RBS;        // rows per CTA
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wr = warp & 3, wb = warp >> 2;
    const int rb0 = blockIdx.x * RBS;
    const int kb0 = blockIdx.y * p.KC;
    const int kc = min(p.KC, p.KB - kb0);
    const int a_stride = p.KC * 128 + 16;
    unsigned char* ws = smem;                              // [RBS][KC][512]
    unsigned char* as = smem + RBS * p.KC * WF_TILE_BYTES;  // [64][a_stride]

    pdl_launch_dependents();
    // ---- weights: RBS contiguous strips of kc*512 bytes
    {
        const int gran_per_strip = kc * 32;
        for (int g = tid; g < RBS * gran_per_strip; g += GO_T) {
            const int rbl = g / gran_per_strip, gi = g - rbl * gran_per_strip;
            const int rb = rb0 + rbl;
            const bool ok = rb < p.RB;
            const unsigned char* src = p.w + ((size_t)(ok ? rb : 0) * p.KB + kb0) * WF_TILE_BYTES + (size_t)gi * 16;
            cp_async16_zfill(ws + (size_t)rbl * p.KC * WF_TILE_BYTES + (size_t)gi * 16, src, ok ? 16 : 0);
        }
        cp_async_commit();
    }
    pdl_wait_prior();
    // ---- activations: 64 rows x kc*128 bytes
    {
        const int gran_per_row = kc * 8;
        for (int g = tid; g < GO_MT * gran_per_row; g += GO_T) {
            const int m = g / gran_per_row, gi = g - m * gran_per_row;
            const bool ok = m < p.M;
            const int8_t* src = p.q + (size_t)(ok ? m : 0) * p.ldq + p.k_off + (size_t)kb0 * 128 + (size_t)gi * 16;
            cp_async16_zfill(as + (size_t)m * a_stride + (size_t)gi * 16, src, ok ? 16 : 0);
        }
        cp_async_commit();
    }
    int acc[RBW][4][4];
#pragma unroll
    for (int i = 0; i < RBW; ++i)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][nt][j] = 0;
    cp_async_wait<0>();
    __syncthreads();

    const bool warp_active = (wb * 32) < p.M;
    const int nt_cnt = min(4, (p.M - wb * 32 + 7) >> 3);   // 8-token column tiles this warp really has (M = 16: two, not four)
    // FULL = all four 8-token column tiles of this warp are live (compile-time trip counts: the predicated form costs
    // ~7 us per launch at M = 64); otherwise only nt_cnt of them (M = 16: two — half the IMMA work)
    auto main_loop = [&](auto full_tag) {
        constexpr bool FULL = decltype(full_tag)::value;
        for (int kbl = 0; kbl < kc; ++kbl) {
            uint32_t b[4][8];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                if (FULL || nt < nt_cnt) {
                    const unsigned char* ap = as + (size_t)((wb * 4 + nt) * 8 + (lane & 7)) * a_stride + kbl * 128 + (lane >> 3) * 16;
                    ldmatrix_x4(b[nt][0], b[nt][1], b[nt][2], b[nt][3], ap);
                    ldmatrix_x4(b[nt][4], b[nt][5], b[nt][6], b[nt][7], ap + 64);
                }
            }
#pragma unroll
            for (int i = 0; i < RBW; ++i) {
                const uint4 Wv = *reinterpret_cast<const uint4*>(ws + ((size_t)(wr * RBW + i) * p.KC + kbl) * WF_TILE_BYTES + lane * 16);
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const int sh = 6 - 2 * g;   // field of k-group g (ggml-bitnet-mad.cpp:76-81)
                    const uint32_t a0 = (Wv.x >> sh) & 0x03030303u, a1 = (Wv.y >> sh) & 0x03030303u;
                    const uint32_t a2 = (Wv.z >> sh) & 0x03030303u, a3 = (Wv.w >> sh) & 0x03030303u;
#pragma unroll
                    for (int nt = 0; nt < 4; ++nt)
                        if (FULL || nt < nt_cnt) mma_u8s8_16832(acc[i][nt], a0, a1, a2, a3, b[nt][2 * g], b[nt][2 * g + 1]);
                }
            }
        }
    };
    if (warp_active) main_loop(std::integral_constant<bool, FULLM>{});
    __syncthreads();   // everyone is done with the load buffers

    // ---- accumulators -> St[m][RT] (fp32, stride RT+4) -> vectorised exact RED
    float* St = reinterpret_cast<float*>(smem);
    constexpr int SS = RT + 4;
    if (warp_active) {
        const int g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int i = 0; i < RBW; ++i)
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                const int ml = (wb * 4 + nt) * 8 + 2 * t;
                const int nl = (wr * RBW + i) * 16 + g;
                St[ml * SS + nl] = (float)acc[i][nt][0];
                St[(ml + 1) * SS + nl] = (float)acc[i][nt][1];
                St[ml * SS + nl + 8] = (float)acc[i][nt][2];
                St[(ml + 1) * SS + nl + 8] = (float)acc[i][nt][3];
            }
    }
    __syncthreads();
    const int mcount = min(GO_MT, p.M);
    const int row0 = rb0 * 16;
    for (int idx = tid; idx < mcount * (RT / 4); idx += GO_T) {
        const int ml = idx / (RT / 4), c = idx - ml * (RT / 4);
        const int n = row0 + 4 * c;
        if (n >= p.N) continue;
        const float4 v = *reinterpret_cast<const float4*>(St + ml * SS + 4 * c);
        if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) {
            // row-parallel linears under TP: the RED into every rank's accumulator IS the all-reduce
            // all-reduce = the RED into every rank's accumulator; reduce-scatter (own_cols > 0) = the RED into the owner's only:
            // 1/T of the NVLink packets, the consumer pulls the columns from their owners
            if (p.dest.own_cols > 0) red_add_v4_f32(p.dest.acc[n / p.dest.own_cols] + (size_t)ml * p.acc_ld + n, v.x, v.y, v.z, v.w);
            else
                for (int d = 0; d < p.dest.n; ++d) red_add_v4_f32(p.dest.acc[d] + (size_t)ml * p.acc_ld + n, v.x, v.y, v.z, v.w);
        }
    }
    if (p.dest.signal_flag >= 0) tp_signal_grid_done(p.tp, p.dest.signal_flag, gridDim.x * gridDim.y);
}

template <int RBW, bool FULLM>
static int launch_rbw(const OneshotParams& p, cudaStream_t s, bool pdl) {
    constexpr int RBS = 4 * RBW;
    const size_t smem_load = (size_t)RBS * p.KC * WF_TILE_BYTES + (size_t)GO_MT * (p.KC * 128 + 16);
    const size_t smem_epi = (size_t)GO_MT * (16 * RBS + 4) * 4;
    const size_t smem = smem_load > smem_epi ? smem_load : smem_epi;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((p.RB + RBS - 1) / RBS, (p.KB + p.KC - 1) / p.KC);
    cfg.blockDim = dim3(GO_T);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, gemm_oneshot_kernel<RBW, FULLM>, p), "gemm_oneshot");
}

static int g_kc_max = 0;

int launch_gemm_oneshot(const OneshotArgs& a, cudaStream_t s, bool pdl) {
    if (a.M < 1 || a.M > GO_MT) { set_error("gemm_oneshot: M=%d outside 1..64", a.M); return WF_ERR_UNSUPPORTED; }
    if (a.K % WF_QK || a.N % 4 || a.acc_ld % 4) { set_error("gemm_oneshot: K %% 128, N %% 4, acc_ld %% 4 must be 0"); return WF_ERR_INVALID; }
    OneshotParams p;
    p.w = a.w; p.q = a.q; p.M = a.M; p.N = a.N; p.K = a.K;
    p.KB = a.K / WF_QK;
    p.RB = (a.N + 15) / 16;
    p.ldq = a.ldq; p.k_off = a.k_off;
    p.dest = a.dest; p.acc_ld = a.acc_ld; p.tp = a.tp;
    if (a.ldq % 16 || a.k_off % WF_QK) { set_error("gemm_oneshot: ldq %% 16 and k_off %% 128 must be 0"); return WF_ERR_INVALID; }
    if (!g_kc_max) {
        const char* e = getenv("WF_ONESHOT_KC");
        g_kc_max = e ? atoi(e) : 4;
        if (g_kc_max < 1 || g_kc_max > 12) g_kc_max = 4;
    }
    // wide matrices take 128-row units (halves the L2->SM activation traffic), the rest 64-row units
    const bool wide = p.RB >= 8 * 64;
    const int kcmax = wide ? (g_kc_max > 5 ? g_kc_max - 1 : g_kc_max) : g_kc_max;
    const int splits = (p.KB + kcmax - 1) / kcmax;
    p.KC = (p.KB + splits - 1) / splits;
    if (p.M % 32 == 0) return wide ? launch_rbw<2, true>(p, s, pdl) : launch_rbw<1, true>(p, s, pdl);
    return wide ? launch_rbw<2, false>(p, s, pdl) : launch_rbw<1, false>(p, s, pdl);
}

void linear_small_init_attrs();
void fused_init_attrs() {
    linear_small_init_attrs();
    layer_tail_init_attrs();
    layer_tail_batch_init_attrs();
    gemm_stream_init_attrs();
    cudaFuncSetAttribute(gemm_oneshot_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(gemm_oneshot_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(gemm_oneshot_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(gemm_oneshot_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
}

}  // namespace wf
