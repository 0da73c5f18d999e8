// Ternary GEMM for small / medium batch (M up to a few hundred): the general kernel behind
// bitnet_gemm and the batched-decode linears.
//
//   S[m,n] = sum_k code[n,k] * a[m,k],  code = w+1 in {0,1,2} (2-bit, I2_S or b200_tile_v1),  a int8
//
// Computes exactly what ggml_vec_dot_i2_i8_s (extern/BitNet/src/ggml-bitnet-mad.cpp:94-186) computes per
// (row, activation vector) — in exact int32, without that kernel's int16 partial sums.
//
// Design (bandwidth-bound, weights are read once from HBM):
//   * CTA = 8 warps; CTA tile = (WR*16 weight rows) x (WB*NT*8 batch columns); K is walked in 256-wide
//     chunks through a 3-stage cp.async (LDGSTS, 16 B, L2-only) ring;
//   * packed weights stay packed in shared memory; each lane pulls its four A-fragment source words with
//     ONE 16-byte LDS (tile layout) and unpacks 2-bit -> u8 in registers with shift+mask;
//   * int8 activations are fetched as mma B fragments with ldmatrix.x4 from a padded (conflict-free) tile;
//   * mma.sync.m16n8k32.u8.s8 -> int32 accumulators in registers (this is an HBM-bound op: the tensor pipe
//     runs at a few percent, tcgen05 buys nothing here; the prefill-size contraction lives in gemm_tc.cu);
//   * optional split-K over gridDim.z: exact int32 atomics into a zeroed workspace, the last CTA of a tile
//     applies the epilogue and re-zeroes the workspace (integer sums are order-independent => deterministic).
#include "linear_common.cuh"

namespace wf {

constexpr int GK = 256;        // k per pipeline stage (2 I2_S blocks)
constexpr int GSTAGES = 3;
constexpr int ACT_STRIDE = GK + 16;  // bytes per activation row in smem (16-B skew => conflict-free ldmatrix)

struct GemmParams {
    const uint8_t* w;
    const int8_t* q;
    int M, N, K, ldq;
    int chunks_total;      // ceil(K / 256)
    int chunks_per_split;
    int* ws_acc;           // [M][N] int32 (split-K)
    int* ws_cnt;           // [row tiles * batch tiles]
    EpiParams ep;
};

template <int LAYOUT, int WR, int WB, int NT>
__global__ void __launch_bounds__(256) ternary_gemm_mma_kernel(const GemmParams p) {
    constexpr int ROWS = WR * 16;
    constexpr int MT = WB * NT * 8;
    constexpr int W_STAGE = WR * 2 * WF_TILE_BYTES;
    constexpr int A_STAGE = MT * ACT_STRIDE;
    constexpr int STAGE = W_STAGE + A_STAGE;
    extern __shared__ __align__(128) unsigned char smem[];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wr = warp % WR, wb = warp / WR;
    const int row0 = blockIdx.x * ROWS;
    const int m0 = blockIdx.y * MT;
    const int KB = p.K / WF_QK;
    const int c_begin = blockIdx.z * p.chunks_per_split;
    const int c_end = min(p.chunks_total, c_begin + p.chunks_per_split);
    const int nchunks = c_end - c_begin;

    auto load_stage = [&](int chunk, int stage) {
        unsigned char* ws = smem + stage * STAGE;
        unsigned char* as = ws + W_STAGE;
        const int kb0 = chunk * 2;
        // weights: WR row-blocks x 2 k-blocks x 512 B = WR*64 16-byte granules
        for (int g = tid; g < WR * 64; g += 256) {
            const int rbl = g >> 6, r = g & 63;       // r: granule inside the 1024-B (2 k-block) strip
            const int kbl = r >> 5, gi = r & 31;      // k-block within chunk, granule within tile
            const int kb = kb0 + kbl;
            const int rb = blockIdx.x * WR + rbl;
            const void* src;
            int ok;
            if (LAYOUT == WF_W_TILE) {
                ok = (kb < KB) && (rb * 16 < p.N);
                src = p.w + ((size_t)rb * KB + kb) * WF_TILE_BYTES + gi * 16;
            } else {
                const int row = rb * 16 + (gi >> 1);
                ok = (kb < KB) && (row < p.N);
                src = p.w + (size_t)row * (p.K / 4) + kb * 32 + (gi & 1) * 16;
            }
            cp_async16_zfill(ws + (rbl * 2 + kbl) * WF_TILE_BYTES + gi * 16, ok ? src : p.w, ok ? 16 : 0);
        }
        // activations: MT rows x 256 B = MT*16 granules
        for (int g = tid; g < MT * 16; g += 256) {
            const int r = g >> 4, c = g & 15;
            const int m = m0 + r;
            const int k = chunk * GK + c * 16;
            const int ok = (m < p.M) && (k < p.K);
            cp_async16_zfill(as + r * ACT_STRIDE + c * 16, ok ? (const void*)(p.q + (size_t)m * p.ldq + k) : (const void*)p.q,
                             ok ? 16 : 0);
        }
    };

    int acc[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0;

    pdl_wait_prior();
#pragma unroll
    for (int s = 0; s < GSTAGES - 1; ++s) {
        if (s < nchunks) load_stage(c_begin + s, s);
        cp_async_commit();
    }

    for (int c = 0; c < nchunks; ++c) {
        cp_async_wait<GSTAGES - 2>();
        __syncthreads();
        {   // prefetch chunk c + STAGES-1 into the slot consumed at iteration c-1
            const int nc = c + GSTAGES - 1;
            if (nc < nchunks) load_stage(c_begin + nc, nc % GSTAGES);
            cp_async_commit();
        }
        const unsigned char* ws = smem + (c % GSTAGES) * STAGE;
        const unsigned char* as = ws + W_STAGE;
#pragma unroll
        for (int kbl = 0; kbl < 2; ++kbl) {
            const unsigned char* tile = ws + (wr * 2 + kbl) * WF_TILE_BYTES;
            uint32_t W0, W1, W2, W3;
            if (LAYOUT == WF_W_TILE) {
                const uint4 t = *reinterpret_cast<const uint4*>(tile + lane * 16);
                W0 = t.x; W1 = t.y; W2 = t.z; W3 = t.w;
            } else {
                const int g = lane >> 2, t4 = (lane & 3) * 4;
                W0 = *reinterpret_cast<const uint32_t*>(tile + g * 32 + t4);
                W1 = *reinterpret_cast<const uint32_t*>(tile + (g + 8) * 32 + t4);
                W2 = *reinterpret_cast<const uint32_t*>(tile + g * 32 + 16 + t4);
                W3 = *reinterpret_cast<const uint32_t*>(tile + (g + 8) * 32 + 16 + t4);
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const unsigned char* arow =
                    as + ((wb * NT + nt) * 8 + (lane & 7)) * ACT_STRIDE + kbl * 128 + (lane >> 3) * 16;
                uint32_t b[8];
                ldmatrix_x4(b[0], b[1], b[2], b[3], arow);        // k-steps 0,1
                ldmatrix_x4(b[4], b[5], b[6], b[7], arow + 64);   // k-steps 2,3
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const int sh = 6 - 2 * g;   // field of k-group g (ggml-bitnet-mad.cpp:76-81)
                    mma_u8s8_16832(acc[nt], (W0 >> sh) & 0x03030303u, (W1 >> sh) & 0x03030303u,
                                   (W2 >> sh) & 0x03030303u, (W3 >> sh) & 0x03030303u, b[2 * g], b[2 * g + 1]);
                }
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();

    // ---- accumulators -> shared int32 tile St[m][n] (row stride ROWS+1) ----
    int* St = reinterpret_cast<int*>(smem);
    constexpr int SS = ROWS + 1;
    {
        const int g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int mloc = (wb * NT + nt) * 8 + t * 2;
            const int nloc = wr * 16 + g;
            St[mloc * SS + nloc] = acc[nt][0];
            St[(mloc + 1) * SS + nloc] = acc[nt][1];
            St[mloc * SS + nloc + 8] = acc[nt][2];
            St[(mloc + 1) * SS + nloc + 8] = acc[nt][3];
        }
    }
    __syncthreads();

    const int mcount = min(MT, p.M - m0);
    const int ncount = min(ROWS, p.N - row0);
    bool finalize = true;
    if (gridDim.z > 1) {
        for (int i = tid; i < mcount * ROWS; i += 256) {
            const int ml = i / ROWS, nl = i % ROWS;
            if (nl < ncount) {
                const int v = St[ml * SS + nl];
                if (v != 0) atomicAdd(p.ws_acc + (size_t)(m0 + ml) * p.N + row0 + nl, v);
            }
        }
        __threadfence();
        __syncthreads();
        __shared__ int s_last;
        if (tid == 0) {
            const int slot = blockIdx.y * gridDim.x + blockIdx.x;
            const int prev = atomicAdd(p.ws_cnt + slot, 1);
            s_last = (prev == (int)gridDim.z - 1);
            if (s_last) p.ws_cnt[slot] = 0;   // self-cleaning for the next launch
        }
        __syncthreads();
        finalize = s_last != 0;
        if (finalize) {
            __threadfence();
            for (int i = tid; i < mcount * ROWS; i += 256) {
                const int ml = i / ROWS, nl = i % ROWS;
                if (nl < ncount) {
                    int* a = p.ws_acc + (size_t)(m0 + ml) * p.N + row0 + nl;
                    St[ml * SS + nl] = ldcg_i32(a);
                    *a = 0;
                }
            }
            __syncthreads();
        }
    }
    if (finalize) {
        if (p.ep.out_mode == WF_OUT_LINEAR && (p.ep.epi_flags & WF_EPI_RELU2_MUL)) {
            // pairs (gate row nl, up row nl+16) inside 32-row groups
            for (int i = tid; i < mcount * (ROWS / 2); i += 256) {
                const int ml = i / (ROWS / 2), j = i % (ROWS / 2);
                const int nl = (j >> 4) * 32 + (j & 15);
                if (row0 + nl + 16 < p.N)
                    epilogue_store_relu2(p.ep, m0 + ml, row0 + nl, St[ml * SS + nl], St[ml * SS + nl + 16]);
            }
        } else {
            for (int i = tid; i < mcount * ROWS; i += 256) {
                const int ml = i / ROWS, nl = i % ROWS;
                if (nl < ncount) epilogue_store(p.ep, m0 + ml, row0 + nl, St[ml * SS + nl]);
            }
        }
    }
    pdl_launch_dependents();
}

template <int LAYOUT, int WR, int WB, int NT>
static int launch_cfg(const GemmParams& p, int splits, cudaStream_t s) {
    constexpr int ROWS = WR * 16, MT = WB * NT * 8;
    constexpr int STAGE = WR * 2 * WF_TILE_BYTES + MT * ACT_STRIDE;
    constexpr int SM_PIPE = STAGE * GSTAGES;
    constexpr int SM_EPI = MT * (ROWS + 1) * 4;
    constexpr int SMEM = SM_PIPE > SM_EPI ? SM_PIPE : SM_EPI;
    auto kern = ternary_gemm_mma_kernel<LAYOUT, WR, WB, NT>;
    static DevOnce attr;
    if (attr.need()) {
        if (cuda_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM), "smem attr")) return WF_ERR_CUDA;
    }
    dim3 grid((p.N + ROWS - 1) / ROWS, (p.M + MT - 1) / MT, splits);
    kern<<<grid, 256, SMEM, s>>>(p);
    return launch_check(cudaGetLastError(), "ternary_gemm_mma");
}

template <int LAYOUT, int WR, int WB, int NT>
static void set_attr_cfg() {
    constexpr int ROWS = WR * 16, MT = WB * NT * 8;
    constexpr int STAGE = WR * 2 * WF_TILE_BYTES + MT * ACT_STRIDE;
    constexpr int SM_PIPE = STAGE * GSTAGES;
    constexpr int SM_EPI = MT * (ROWS + 1) * 4;
    constexpr int SMEM = SM_PIPE > SM_EPI ? SM_PIPE : SM_EPI;
    cudaFuncSetAttribute(ternary_gemm_mma_kernel<LAYOUT, WR, WB, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
}

// called once at engine creation so that no function attribute is touched during stream capture
void gemm_mma_init_attrs() {
    set_attr_cfg<WF_W_TILE, 8, 1, 1>(); set_attr_cfg<WF_W_TILE, 8, 1, 2>(); set_attr_cfg<WF_W_TILE, 8, 1, 4>(); set_attr_cfg<WF_W_TILE, 4, 2, 4>();
    set_attr_cfg<WF_W_I2S, 8, 1, 1>(); set_attr_cfg<WF_W_I2S, 8, 1, 2>(); set_attr_cfg<WF_W_I2S, 8, 1, 4>(); set_attr_cfg<WF_W_I2S, 4, 2, 4>();
}

size_t gemm_mma_workspace_bytes(int M, int N, int K) {
    (void)K;
    // int32 [M][N] accumulators + one counter per (16-row block, 8-column block) upper bound
    return (size_t)M * N * 4 + (size_t)((N + 15) / 16) * ((M + 7) / 8) * 4 + 256;
}

template <int LAYOUT>
static int dispatch(const wf_linear_args* a, GemmParams& p, cudaStream_t s) {
    const int M = a->M;
    const int chunks = p.chunks_total;
    // pick tile config
    int rows, mt;
    if (M <= 8) { rows = 128; mt = 8; }
    else if (M <= 16) { rows = 128; mt = 16; }
    else if (M <= 32) { rows = 128; mt = 32; }
    else { rows = 64; mt = 64; }
    const int ctas = ((a->N + rows - 1) / rows) * ((M + mt - 1) / mt);
    // split-K only while the int32 partial-sum traffic (M*rows*4 per split) stays below the weight bytes
    int splits = 1;
    if (a->workspace && a->workspace_bytes >= gemm_mma_workspace_bytes(M, a->N, a->K)) {
        const int target = 2 * sm_count();
        while (splits < chunks && ctas * splits < target && (mt * 4 * 2) * (splits + 1) <= (a->K / 4)) ++splits;
    }
    p.chunks_per_split = (chunks + splits - 1) / splits;
    splits = (chunks + p.chunks_per_split - 1) / p.chunks_per_split;
    p.ws_acc = (int*)a->workspace;
    p.ws_cnt = a->workspace ? (int*)((char*)a->workspace + (size_t)M * a->N * 4) : nullptr;
    if (M <= 8) return launch_cfg<LAYOUT, 8, 1, 1>(p, splits, s);
    if (M <= 16) return launch_cfg<LAYOUT, 8, 1, 2>(p, splits, s);
    if (M <= 32) return launch_cfg<LAYOUT, 8, 1, 4>(p, splits, s);
    return launch_cfg<LAYOUT, 4, 2, 4>(p, splits, s);
}

int launch_gemm_mma(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s) {
    GemmParams p;
    p.w = a->w;
    p.q = a->q;
    p.M = a->M; p.N = a->N; p.K = a->K; p.ldq = a->ldq;
    p.chunks_total = (a->K + GK - 1) / GK;
    p.chunks_per_split = p.chunks_total;
    p.ws_acc = nullptr; p.ws_cnt = nullptr;
    p.ep = ep;
    if (a->w_layout == WF_W_TILE) return dispatch<WF_W_TILE>(a, p, s);
    return dispatch<WF_W_I2S>(a, p, s);
}

}  // namespace wf
