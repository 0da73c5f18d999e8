// Weight-format kernels: ternary float -> I2_S, I2_S -> float, HF checkpoint -> I2_S,
// I2_S <-> b200_tile_v1 (the GPU tile layout), 16-row interleave for gate|up fusion.
//
// Reference semantics: src/wrinklefree_inference/cache/bitnet_converter.py:26-77
// (_pack_ternary_weights), extern/BitNet/src/ggml-bitnet-mad.cpp:46-92 (quantize_i2_s),
// tests/test_bitnet_core.py:21-53 (_unpack_ternary_weights, HF layout), :109-136
// (_unpack_i2_to_ternary).  All pure byte/integer work, HBM-bound; one thread per output
// u32 (coalesced stores), grid-stride.
#include "common.cuh"
#include "wf_bitnet.h"
#include <cuda_fp16.h>

namespace wf {

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

__device__ __forceinline__ uint32_t ternary_code(float w) {
    float t = rintf(w);  // torch.round: ties to even
    t = fminf(fmaxf(t, -1.f), 1.f);
    return (uint32_t)((int)t + 1);
}

// one thread -> 4 consecutive packed bytes (one u32) of one row
template <typename T>
__global__ void pack_ternary_kernel(const T* __restrict__ w, int N, int K, int bb, int nb, uint8_t* __restrict__ out,
                                    unsigned int* __restrict__ amax_bits) {
    const int K4 = K / 4;
    const int words_per_row = (K4 + 3) / 4;
    const long long total = (long long)N * words_per_row;
    float amax = 0.f;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(idx / words_per_row);
        const int j0 = (int)(idx % words_per_row) * 4;
        const T* row = w + (size_t)n * K;
        uint32_t word = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int j = j0 + e;  // packed byte index in the row
            if (j >= K4) break;
            uint32_t byte = 0;
            if (j < nb * bb) {
                const int b = j / bb, k = j % bb;
                const int base = b * bb * 4;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const float v = to_f32(row[base + k + g * bb]);
                    byte |= ternary_code(v) << (6 - 2 * g);
                }
            }
            word |= byte << (8 * e);
        }
        // abs-max over ALL K elements of the row segment this thread covers: the 16 floats at k-index
        // positions are scattered, so sweep the 16 source elements we touched plus (for tails) nothing else.
        if ((K4 & 3) == 0 && j0 + 4 <= K4) {
            *reinterpret_cast<uint32_t*>(out + (size_t)n * K4 + j0) = word;
        } else {
            for (int e = 0; j0 + e < K4; ++e) out[(size_t)n * K4 + j0 + e] = (uint8_t)(word >> (8 * e));
        }
    }
    // separate coalesced pass for max|w| (covers tails that packing skips, like weight.abs().max())
    const long long nel = (long long)N * K;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nel; i += (long long)gridDim.x * blockDim.x)
        amax = fmaxf(amax, fabsf(to_f32(w[i])));
    amax = warp_max(amax);
    if ((threadIdx.x & 31) == 0 && amax > 0.f) atomicMax(amax_bits, __float_as_uint(amax));
}

__global__ void finalize_weight_scale_kernel(float* scale) {
    float s = *scale;
    if (s < 1e-6f) *scale = 1.0f;
}

__global__ void unpack_i2s_kernel(const uint8_t* __restrict__ p, int N, int K, int bb, int nb, float* __restrict__ w) {
    const long long total = (long long)N * K;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / K), k = (int)(i % K);
        float v = 0.f;
        if (k < nb * bb * 4) {
            const int b = k / (bb * 4), r = k % (bb * 4);
            const int g = r / bb, j = r % bb;
            const uint8_t byte = p[(size_t)n * (K / 4) + b * bb + j];
            v = (float)((byte >> (6 - 2 * g)) & 3) - 1.0f;
        }
        w[i] = v;
    }
}

// HF u8 [N4, K] (bits 2i+1..2i of row r = weight row i*N4 + r) -> I2_S [4*N4, K/4].
// One thread per output u32 (4 packed bytes = 16 weights of one output row).
__global__ void hf_to_i2s_kernel(const uint8_t* __restrict__ hf, int N4, int K, int bb, int nb, uint8_t* __restrict__ out) {
    const int K4 = K / 4;
    const int words_per_row = K4 / 4;
    const long long total = (long long)4 * N4 * words_per_row;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(idx / words_per_row);
        const int j0 = (int)(idx % words_per_row) * 4;
        const int i = n / N4, r = n % N4;
        const uint8_t* src = hf + (size_t)r * K;
        uint32_t word = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int j = j0 + e;
            uint32_t byte = 0;
            if (j < nb * bb) {
                const int b = j / bb, k = j % bb;
#pragma unroll
                for (int g = 0; g < 4; ++g) byte |= ((uint32_t)(src[b * bb * 4 + k + g * bb] >> (2 * i)) & 3u) << (6 - 2 * g);
            }
            word |= byte << (8 * e);
        }
        *reinterpret_cast<uint32_t*>(out + (size_t)n * K4 + j0) = word;
    }
}

// b200_tile_v1: tile (rb, kb) at byte ((rb * KB + kb) * 512); inside, u32 index (4*l + i) for lane l,
// word i holds I2_S bytes [j, j+4) of row (rb*16 + (l>>2) + 8*(i&1)), j = 4*(l&3) + 16*(i>>1), block kb.
__global__ void i2s_to_tile_kernel(const uint8_t* __restrict__ i2s, int N, int K, uint32_t* __restrict__ tile) {
    const int KB = K / WF_QK;
    const int RB = (N + 15) / 16;
    const long long total = (long long)RB * KB * 128;  // u32 words
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int wi = (int)(idx & 127);
        const long long t = idx >> 7;
        const int kb = (int)(t % KB), rb = (int)(t / KB);
        const int l = wi >> 2, i = wi & 3;
        const int row = rb * 16 + (l >> 2) + 8 * (i & 1);
        const int j = 4 * (l & 3) + 16 * (i >> 1);
        uint32_t v = 0x55555555u;  // code 1 == weight 0 for padded rows
        if (row < N) v = *reinterpret_cast<const uint32_t*>(i2s + (size_t)row * (K / 4) + kb * 32 + j);
        tile[idx] = v;
    }
}

__global__ void tile_to_i2s_kernel(const uint32_t* __restrict__ tile, int N, int K, uint8_t* __restrict__ i2s) {
    const int KB = K / WF_QK;
    const int RB = (N + 15) / 16;
    const long long total = (long long)RB * KB * 128;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int wi = (int)(idx & 127);
        const long long t = idx >> 7;
        const int kb = (int)(t % KB), rb = (int)(t / KB);
        const int l = wi >> 2, i = wi & 3;
        const int row = rb * 16 + (l >> 2) + 8 * (i & 1);
        const int j = 4 * (l & 3) + 16 * (i >> 1);
        if (row < N) *reinterpret_cast<uint32_t*>(i2s + (size_t)row * (K / 4) + kb * 32 + j) = tile[idx];
    }
}

__global__ void interleave_rows16_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int N, int K4v,
                                         uint4* __restrict__ out) {
    const long long total = (long long)2 * N * K4v;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(idx % K4v);
        const int orow = (int)(idx / K4v);
        const int blk = orow / 32, r = orow % 32;
        const int srow = blk * 16 + (r & 15);
        out[idx] = (r < 16 ? a : b)[(size_t)srow * K4v + c];
    }
}

static inline int grid_for(long long total, int threads) {
    long long g = (total + threads - 1) / threads;
    const long long cap = (long long)sm_count() * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

static void block_geom(int K, int* bb, int* nb) {
    *bb = 32;
    *nb = K / 128;
    if (*nb == 0) { *bb = K / 4; *nb = 1; }
}

}  // namespace wf

using namespace wf;

extern "C" int wf_pack_ternary(const void* w, int w_dtype, int N, int K, uint8_t* i2s, float* scale_out, void* stream) {
    if (!w || !i2s || !scale_out || N <= 0 || K <= 0) { set_error("wf_pack_ternary: null pointer or empty shape"); return WF_ERR_INVALID; }
    if (K % 4 != 0) { set_error("in_features (%d) must be divisible by 4", K); return WF_ERR_INVALID; }
    cudaStream_t s = (cudaStream_t)stream;
    int bb, nb;
    block_geom(K, &bb, &nb);
    if (cuda_check(cudaMemsetAsync(scale_out, 0, sizeof(float), s), "memset scale")) return WF_ERR_CUDA;
    const long long total = (long long)N * K;
    const int g = grid_for(total, 256);
    if (w_dtype == WF_F32)
        pack_ternary_kernel<float><<<g, 256, 0, s>>>((const float*)w, N, K, bb, nb, i2s, (unsigned int*)scale_out);
    else if (w_dtype == WF_BF16)
        pack_ternary_kernel<__nv_bfloat16><<<g, 256, 0, s>>>((const __nv_bfloat16*)w, N, K, bb, nb, i2s, (unsigned int*)scale_out);
    else if (w_dtype == WF_F16)
        pack_ternary_kernel<__half><<<g, 256, 0, s>>>((const __half*)w, N, K, bb, nb, i2s, (unsigned int*)scale_out);
    else { set_error("wf_pack_ternary: bad dtype %d", w_dtype); return WF_ERR_INVALID; }
    finalize_weight_scale_kernel<<<1, 1, 0, s>>>(scale_out);
    return launch_check(cudaGetLastError(), "wf_pack_ternary");
}

extern "C" int wf_unpack_i2s(const uint8_t* i2s, int N, int K, float* w, void* stream) {
    if (!i2s || !w || N <= 0 || K <= 0 || K % 4) { set_error("wf_unpack_i2s: bad arguments"); return WF_ERR_INVALID; }
    int bb, nb;
    block_geom(K, &bb, &nb);
    unpack_i2s_kernel<<<grid_for((long long)N * K, 256), 256, 0, (cudaStream_t)stream>>>(i2s, N, K, bb, nb, w);
    return launch_check(cudaGetLastError(), "wf_unpack_i2s");
}

extern "C" int wf_hf_to_i2s(const uint8_t* hf, int N4, int K, uint8_t* i2s, void* stream) {
    if (!hf || !i2s || N4 <= 0 || K <= 0) { set_error("wf_hf_to_i2s: bad arguments"); return WF_ERR_INVALID; }
    if (K % 16 != 0) { set_error("wf_hf_to_i2s: K (%d) must be a multiple of 16", K); return WF_ERR_INVALID; }
    int bb, nb;
    block_geom(K, &bb, &nb);
    hf_to_i2s_kernel<<<grid_for((long long)N4 * K / 4, 256), 256, 0, (cudaStream_t)stream>>>(hf, N4, K, bb, nb, i2s);
    return launch_check(cudaGetLastError(), "wf_hf_to_i2s");
}

extern "C" size_t wf_tile_bytes(int N, int K) {
    if (N <= 0 || K <= 0 || K % WF_QK) return 0;
    return (size_t)((N + 15) / 16) * (K / WF_QK) * WF_TILE_BYTES;
}

extern "C" int wf_repack_i2s_to_tile(const uint8_t* i2s, int N, int K, uint8_t* tile, void* stream) {
    if (!i2s || !tile || N <= 0) { set_error("wf_repack_i2s_to_tile: bad arguments"); return WF_ERR_INVALID; }
    if (K <= 0 || K % WF_QK) { set_error("in_features (%d) must be multiple of %d", K, WF_QK); return WF_ERR_INVALID; }
    const long long words = (long long)wf_tile_bytes(N, K) / 4;
    i2s_to_tile_kernel<<<grid_for(words, 256), 256, 0, (cudaStream_t)stream>>>(i2s, N, K, (uint32_t*)tile);
    return launch_check(cudaGetLastError(), "wf_repack_i2s_to_tile");
}

extern "C" int wf_repack_tile_to_i2s(const uint8_t* tile, int N, int K, uint8_t* i2s, void* stream) {
    if (!i2s || !tile || N <= 0) { set_error("wf_repack_tile_to_i2s: bad arguments"); return WF_ERR_INVALID; }
    if (K <= 0 || K % WF_QK) { set_error("in_features (%d) must be multiple of %d", K, WF_QK); return WF_ERR_INVALID; }
    const long long words = (long long)wf_tile_bytes(N, K) / 4;
    tile_to_i2s_kernel<<<grid_for(words, 256), 256, 0, (cudaStream_t)stream>>>((const uint32_t*)tile, N, K, i2s);
    return launch_check(cudaGetLastError(), "wf_repack_tile_to_i2s");
}

extern "C" int wf_interleave_rows16(const uint8_t* a, const uint8_t* b, int N, int K4, uint8_t* out, void* stream) {
    if (!a || !b || !out || N <= 0 || N % 16 || K4 <= 0 || K4 % 16) { set_error("wf_interleave_rows16: N %% 16 and K/4 %% 16 must be 0"); return WF_ERR_INVALID; }
    const int K4v = K4 / 16;
    interleave_rows16_kernel<<<grid_for((long long)2 * N * K4v, 256), 256, 0, (cudaStream_t)stream>>>(
        (const uint4*)a, (const uint4*)b, N, K4v, (uint4*)out);
    return launch_check(cudaGetLastError(), "wf_interleave_rows16");
}
