// Activation quantisation (north-star item 2) and the RMSNorm that precedes it.
//
// Reference arithmetic: quantize_activations_i8, tests/test_bitnet_core.py:139-146 (copy of the
// missing sgl_kernel/quantization/bitnet.py); rms_norm, scripts/direct_inference_full.py:181-184.
// HBM/L2-bound elementwise + reduction work: one CTA per token row, 16-byte vector loads, the row is
// read twice (absmax pass, quantise pass) — the second read hits L2/L1.
#include "common.cuh"
#include "wf_bitnet.h"
#include <cuda_fp16.h>

namespace wf {

constexpr int QT = 256;  // threads per CTA

template <typename T> struct Vec;  // 16-byte vector of T
template <> struct Vec<float> {
    static constexpr int N = 4;
    __device__ static void load(const float* p, float (&v)[8]) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};
template <> struct Vec<__nv_bfloat16> {
    static constexpr int N = 8;
    __device__ static void load(const __nv_bfloat16* p, float (&v)[8]) {
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        v[0] = bf16lo(t.x); v[1] = bf16hi(t.x); v[2] = bf16lo(t.y); v[3] = bf16hi(t.y);
        v[4] = bf16lo(t.z); v[5] = bf16hi(t.z); v[6] = bf16lo(t.w); v[7] = bf16hi(t.w);
    }
};
template <> struct Vec<__half> {
    static constexpr int N = 8;
    __device__ static void load(const __half* p, float (&v)[8]) {
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        const __half2* h = reinterpret_cast<const __half2*>(&t);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
};

template <typename T>
__device__ __forceinline__ float row_absmax(const T* row, int K) {
    constexpr int VN = Vec<T>::N;
    float amax = 0.f;
    float v[8];
    for (int k = threadIdx.x * VN; k < K; k += QT * VN) {
        Vec<T>::load(row + k, v);
#pragma unroll
        for (int i = 0; i < VN; ++i) amax = fmaxf(amax, fabsf(v[i]));
    }
    return amax;
}

// quantise one row with a known scale; returns this thread's partial sum of q
template <typename T>
__device__ __forceinline__ int row_quant(const T* row, int K, float scale, int8_t* q) {
    constexpr int VN = Vec<T>::N;
    int psum = 0;
    float v[8];
    const QDiv qd = make_qdiv(scale);
    for (int k = threadIdx.x * VN; k < K; k += QT * VN) {
        Vec<T>::load(row + k, v);
        uint32_t w[2] = {0, 0};
#pragma unroll
        for (int i = 0; i < VN; ++i) {
            const int qi = quant_i8(v[i], qd);
            psum += qi;
            w[i >> 2] |= (uint32_t)(qi & 0xff) << (8 * (i & 3));
        }
        if (VN == 4) *reinterpret_cast<uint32_t*>(q + k) = w[0];
        else *reinterpret_cast<uint2*>(q + k) = make_uint2(w[0], w[1]);
    }
    return psum;
}

// per-token: grid = M
template <typename T>
__global__ void __launch_bounds__(QT) act_quant_rows_kernel(const T* __restrict__ x, int K, int ldx, int8_t* __restrict__ q,
                                                            float* __restrict__ scale_out, int32_t* __restrict__ rowsum) {
    __shared__ float fs[33];
    __shared__ int is[33];
    const int m = blockIdx.x;
    const T* row = x + (size_t)m * ldx;
    const float amax = block_max<QT>(row_absmax(row, K), fs);
    const float scale = act_scale_from_absmax(amax);
    const int total = block_sum_i<QT>(row_quant(row, K, scale, q + (size_t)m * K), is);
    if (threadIdx.x == 0) {
        scale_out[m] = scale;
        if (rowsum) rowsum[m] = total;
    }
}

// per-tensor pass 1: grid-stride absmax -> atomicMax on the float bit pattern (values are >= 0)
template <typename T>
__global__ void __launch_bounds__(QT) act_absmax_kernel(const T* __restrict__ x, int M, int K, int ldx,
                                                        unsigned int* __restrict__ amax_bits) {
    float amax = 0.f;
    for (int m = blockIdx.x; m < M; m += gridDim.x) amax = fmaxf(amax, row_absmax(x + (size_t)m * ldx, K));
    amax = warp_max(amax);
    if ((threadIdx.x & 31) == 0 && amax > 0.f) atomicMax(amax_bits, __float_as_uint(amax));
}
// per-tensor pass 2: grid = M
template <typename T>
__global__ void __launch_bounds__(QT) act_quant_tensor_kernel(const T* __restrict__ x, int K, int ldx,
                                                              const unsigned int* __restrict__ amax_bits,
                                                              int8_t* __restrict__ q, float* __restrict__ scale_out,
                                                              int32_t* __restrict__ rowsum) {
    __shared__ int is[33];
    const int m = blockIdx.x;
    const float scale = act_scale_from_absmax(__uint_as_float(*amax_bits));
    const int total = block_sum_i<QT>(row_quant(x + (size_t)m * ldx, K, scale, q + (size_t)m * K), is);
    if (threadIdx.x == 0) {
        if (m == 0) scale_out[0] = scale;
        if (rowsum) rowsum[m] = total;
    }
}

// RMSNorm (+ optional store) + per-token quant.  Rounding points follow the reference on bf16 tensors:
//   variance = mean(float(x)^2)                         (fp32)
//   r        = bf16( rsqrt(variance + eps) )            (`.to(x.dtype)`)
//   y        = bf16( bf16(x * r) * gamma )              (two bf16 multiplies)
//   q        = quantize_activations_i8(float(y))        (per token)
// The normalised row is kept in shared memory as bf16 (K*2 bytes, dynamic).
__global__ void __launch_bounds__(QT) rmsnorm_quant_kernel(const __nv_bfloat16* __restrict__ x,
                                                           const __nv_bfloat16* __restrict__ gamma, float eps, int K,
                                                           __nv_bfloat16* __restrict__ y_out, int8_t* __restrict__ q,
                                                           float* __restrict__ scale_out, int32_t* __restrict__ rowsum) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __nv_bfloat16* ys = reinterpret_cast<__nv_bfloat16*>(smem_raw);
    __shared__ float fs[33];
    __shared__ int is[33];
    const int m = blockIdx.x;
    const __nv_bfloat16* row = x + (size_t)m * K;
    pdl_wait_prior();
    float ss = 0.f;
    float v[8];
    for (int k = threadIdx.x * 8; k < K; k += QT * 8) {
        Vec<__nv_bfloat16>::load(row + k, v);
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += v[i] * v[i];
    }
    ss = block_sum<QT>(ss, fs);
    const float var = ss / (float)K;
    const float r = bf16_round(1.0f / sqrtf(var + eps));
    float amax = 0.f;
    for (int k = threadIdx.x * 8; k < K; k += QT * 8) {
        float g[8];
        Vec<__nv_bfloat16>::load(row + k, v);
        Vec<__nv_bfloat16>::load(gamma + k, g);
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
            const float y0 = bf16_round(bf16_round(v[i] * r) * g[i]);
            const float y1 = bf16_round(bf16_round(v[i + 1] * r) * g[i + 1]);
            amax = fmaxf(amax, fmaxf(fabsf(y0), fabsf(y1)));
            w[i >> 1] = pack_bf16x2(y0, y1);
        }
        *reinterpret_cast<uint4*>(ys + k) = make_uint4(w[0], w[1], w[2], w[3]);
        if (y_out) *reinterpret_cast<uint4*>(y_out + (size_t)m * K + k) = make_uint4(w[0], w[1], w[2], w[3]);
    }
    amax = block_max<QT>(amax, fs);  // contains __syncthreads: ys is complete
    const float scale = act_scale_from_absmax(amax);
    const int total = block_sum_i<QT>(row_quant<__nv_bfloat16>(ys, K, scale, q + (size_t)m * K), is);
    if (threadIdx.x == 0) {
        scale_out[m] = scale;
        if (rowsum) rowsum[m] = total;
    }
    pdl_launch_dependents();
}

}  // namespace wf

using namespace wf;

template <typename T>
static int launch_act_quant(const T* x, int M, int K, int ldx, int per_token, int8_t* q, float* scale_out, int32_t* rowsum,
                            void* workspace, cudaStream_t s) {
    if (per_token) {
        act_quant_rows_kernel<T><<<M, QT, 0, s>>>(x, K, ldx, q, scale_out, rowsum);
    } else {
        if (cuda_check(cudaMemsetAsync(workspace, 0, 4, s), "memset absmax")) return WF_ERR_CUDA;
        int g = M < sm_count() * 4 ? M : sm_count() * 4;
        act_absmax_kernel<T><<<g, QT, 0, s>>>(x, M, K, ldx, (unsigned int*)workspace);
        act_quant_tensor_kernel<T><<<M, QT, 0, s>>>(x, K, ldx, (const unsigned int*)workspace, q, scale_out, rowsum);
    }
    return launch_check(cudaGetLastError(), "wf_act_quant_i8");
}

extern "C" int wf_act_quant_i8(const void* x, int x_dtype, int M, int K, int ldx, int per_token, int8_t* q, float* scale_out,
                               int32_t* rowsum_out, void* workspace, void* stream) {
    if (!x || !q || !scale_out) { set_error("wf_act_quant_i8: null pointer"); return WF_ERR_INVALID; }
    if (M < 0 || K <= 0) { set_error("wf_act_quant_i8: bad shape M=%d K=%d", M, K); return WF_ERR_INVALID; }
    if (M == 0) return WF_OK;
    if (K % 8 != 0 || ldx % 8 != 0) { set_error("wf_act_quant_i8: K (%d) and ldx (%d) must be multiples of 8", K, ldx); return WF_ERR_UNSUPPORTED; }
    if (!per_token && !workspace) { set_error("wf_act_quant_i8: per-tensor mode needs a 4-byte workspace"); return WF_ERR_INVALID; }
    cudaStream_t s = (cudaStream_t)stream;
    switch (x_dtype) {
        case WF_F32: return launch_act_quant((const float*)x, M, K, ldx, per_token, q, scale_out, rowsum_out, workspace, s);
        case WF_BF16: return launch_act_quant((const __nv_bfloat16*)x, M, K, ldx, per_token, q, scale_out, rowsum_out, workspace, s);
        case WF_F16: return launch_act_quant((const __half*)x, M, K, ldx, per_token, q, scale_out, rowsum_out, workspace, s);
    }
    set_error("wf_act_quant_i8: bad dtype %d", x_dtype);
    return WF_ERR_INVALID;
}

namespace wf {
int launch_rmsnorm_quant(const void* x, const void* gamma, float eps, int M, int K, void* y_out, int8_t* q, float* scale_out,
                         int32_t* rowsum_out, cudaStream_t s, bool pdl) {
    const size_t smem = (size_t)K * 2;
    static DevOnce attr_set;
    if (attr_set.need()) cudaFuncSetAttribute(rmsnorm_quant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(M);
    cfg.blockDim = dim3(QT);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return launch_check(cudaLaunchKernelEx(&cfg, rmsnorm_quant_kernel, (const __nv_bfloat16*)x, (const __nv_bfloat16*)gamma, eps, K,
                                         (__nv_bfloat16*)y_out, q, scale_out, rowsum_out),
                      "wf_rmsnorm_quant");
}
}  // namespace wf

extern "C" int wf_rmsnorm_quant(const void* x, const void* gamma, float eps, int M, int K, void* y_out, int8_t* q,
                                float* scale_out, int32_t* rowsum_out, void* stream) {
    if (!x || !gamma || !q || !scale_out) { set_error("wf_rmsnorm_quant: null pointer"); return WF_ERR_INVALID; }
    if (M <= 0 || K <= 0 || K % 8 != 0) { set_error("wf_rmsnorm_quant: bad shape M=%d K=%d (K %% 8 == 0)", M, K); return WF_ERR_INVALID; }
    if ((size_t)K * 2 > 128 * 1024) { set_error("wf_rmsnorm_quant: K=%d exceeds the 64K-element row limit", K); return WF_ERR_UNSUPPORTED; }
    return launch_rmsnorm_quant(x, gamma, eps, M, K, y_out, q, scale_out, rowsum_out, (cudaStream_t)stream, false);
}

namespace wf {
void quant_init_attrs() { cudaFuncSetAttribute(rmsnorm_quant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024); }
}
