// Shared epilogue of the ternary linear kernels.
//
// Every kernel produces S = sum_k (w+1)*a as exact int32 (the reference kernel's accumulator,
// extern/BitNet/src/ggml-bitnet-mad.cpp:136-186) and hands it to `wf::epilogue_store`, which applies
// one of the output modes of include/wf_bitnet.h.
#pragma once
#include "common.cuh"
#include "wf_bitnet.h"

namespace wf {

struct EpiParams {
    int M, N;
    const int32_t* rowsum;
    const float* act_scale;
    int act_scale_per_token;
    float w_scale_host[4];  // unused on device; kept for debugging
    const float* w_scale;
    int nseg;
    int seg_end[4];
    float host_scale;
    int out_mode, epi_flags, out_dtype;
    void* out;
    int ldo;
    const __nv_bfloat16* residual;
};

__device__ __forceinline__ float seg_scale(const EpiParams& p, int n) {
    int s = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
        if (i + 1 < p.nseg && n >= p.seg_end[i]) s = i + 1;
    return p.w_scale[s];
}

// value of the (dequantised) linear output for accumulator S (= sum (w+1) a) at (m, n)
__device__ __forceinline__ float linear_value(const EpiParams& p, int m, float wscale, int S) {
    // the reference's own rounding sequence (scripts/direct_inference_full.py:202-206), so that the fp32 value — and the bf16 it is
    // rounded to — is bit-identical to the CPU path, not merely within 1e-3:
    //   out = bitnet_gemv(...)            -> fl32(S * s_w)            (float)sumi * scale, one rounding
    //   out - scale * act_sum             -> fl32(out - fl32(s_w * Sigma a))   (the python-double product is exact, the scalar is cast to fp32)
    //   (...) * act_scale                 -> fl32(... * s_a)
    const float sa = p.act_scale[p.act_scale_per_token ? m : 0];
    const float y = __fmul_rn((float)S, wscale);
    const float c = __fmul_rn(wscale, (float)p.rowsum[m]);
    return __fmul_rn(__fsub_rn(y, c), sa);
}

// Non-fused modes and LINEAR without RELU2_MUL: one element.
__device__ __forceinline__ void epilogue_store(const EpiParams& p, int m, int n, int S) {
    switch (p.out_mode) {
        case WF_OUT_I32_PLUS1:
            reinterpret_cast<int32_t*>(p.out)[(size_t)m * p.ldo + n] = S;
            break;
        case WF_OUT_I32:
            reinterpret_cast<int32_t*>(p.out)[(size_t)m * p.ldo + n] = S - p.rowsum[m];
            break;
        case WF_OUT_COMPAT_F32:
            // (float)sumi (mad.cpp:186) times the op's scale argument: one fp32 rounding
            reinterpret_cast<float*>(p.out)[(size_t)m * p.ldo + n] = __fmul_rn((float)S, p.host_scale);
            break;
        default: {  // WF_OUT_LINEAR
            float v = linear_value(p, m, seg_scale(p, n), S);
            if (p.epi_flags & WF_EPI_RESIDUAL) {
                const float r = __bfloat162float(p.residual[(size_t)m * p.ldo + n]);
                v = r + bf16_round(v);  // hidden + proj, both bf16 (direct_inference_full.py:276,283)
            }
            if (p.out_dtype == WF_BF16)
                reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)m * p.ldo + n] = __float2bfloat16_rn(v);
            else
                reinterpret_cast<float*>(p.out)[(size_t)m * p.ldo + n] = v;
        }
    }
}

// RELU2_MUL: rows are interleaved gate/up in 16-row blocks; pair (Sg at row n_g, Su at row n_g+16)
// -> out[m, j] with j = (n_g / 32) * 16 + (n_g % 16).  bf16 rounding after every torch op:
//   torch.relu(gate).square() * up   (direct_inference_full.py:259)
__device__ __forceinline__ void epilogue_store_relu2(const EpiParams& p, int m, int n_g, int Sg, int Su) {
    const float g = bf16_round(linear_value(p, m, p.w_scale[0], Sg));
    const float u = bf16_round(linear_value(p, m, p.w_scale[p.nseg > 1 ? 1 : 0], Su));
    const float r = fmaxf(g, 0.f);
    const float h = bf16_round(bf16_round(r * r) * u);
    const int j = (n_g >> 5) * 16 + (n_g & 15);
    if (p.out_dtype == WF_BF16)
        reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)m * p.ldo + j] = __float2bfloat16_rn(h);
    else
        reinterpret_cast<float*>(p.out)[(size_t)m * p.ldo + j] = h;
}

int fill_epi(const wf_linear_args* a, EpiParams* p);  // validates + copies (host)

// launchers implemented in the kernel files
int launch_gemm_mma(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s);
int launch_gemv_stream(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s, bool pdl);
bool gemv_stream_fits(int M, int K);
int launch_gemm_tc(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s);
size_t gemm_mma_workspace_bytes(int M, int N, int K);

}  // namespace wf
