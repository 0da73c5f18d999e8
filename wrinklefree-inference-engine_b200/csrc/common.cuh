// Shared device helpers for the sm_100a BitNet kernels (inline PTX wrappers, bf16 rounding
// helpers that mirror the reference's torch-CPU rounding points, small reductions).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define WF_OK 0
#define WF_ERR_INVALID (-1)
#define WF_ERR_UNSUPPORTED (-2)
#define WF_ERR_CUDA (-3)

#define WF_QK 128          // BITNET_BLOCK_SIZE / QK_I2_S (reference: sglang_backend/bitnet_quantization.py:45-46)
#define WF_TILE_ROWS 16    // rows per weight tile
#define WF_TILE_BYTES 512  // 16 rows x 128 k x 2 bit

namespace wf {

// ---------------------------------------------------------------------------------------------
// error plumbing (host)
void set_error(const char* fmt, ...);
int cuda_check(cudaError_t e, const char* what);
int launch_check(cudaError_t e, const char* what);   // cuda_check + counts one kernel launch
long long launch_count();
int sm_count();
// "once per device" guard: cudaFuncSetAttribute and the SM count are per-device state, and the op-level C ABI may be called
// from several GPUs of one process (one bit per device ordinal)
struct DevOnce {
    unsigned long long done = 0;
    bool need() {
        int d = 0;
        if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d > 63) return true;
        if ((done >> d) & 1ull) return false;
        done |= 1ull << d;
        return true;
    }
};

// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// streaming 16-byte global load, no L1 allocation (weights / KV are read once)
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const void* p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ int ldcg_i32(const int* p) {
    int r;
    asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

// cp.async (LDGSTS) 16 B, L2-only caching
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
// zero-fill variant: copies src_bytes (0 or 16) and zero-fills the rest
__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------------------------------------
// mbarrier + 1-D bulk async copy (TMA engine, no tensor map needed)
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) { }
}
// global -> shared::cta bulk copy; bytes % 16 == 0, both addresses 16-B aligned
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// L2 prefetch of a byte range (no smem destination)
__device__ __forceinline__ void bulk_prefetch_l2(const void* gsrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}

// Programmatic dependent launch (PDL)
__device__ __forceinline__ void pdl_wait_prior() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// legacy-path tensor ops used by the bandwidth-bound small-batch kernels
// D(16x8,s32) += A(16x32,u8,row) * B(32x8,s8,col)
__device__ __forceinline__ void mma_u8s8_16832(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* smem_row) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(smem_u32(smem_row)));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* smem_row) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(smem_u32(smem_row)));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t& r0, uint32_t& r1, const void* smem_row) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(smem_u32(smem_row)));
}

// ---------------------------------------------------------------------------------------------
// bf16 helpers.  The reference keeps activations in bf16 between ops on the CPU
// (scripts/direct_inference_full.py:76), so every torch op rounds to bf16; these
// reproduce those rounding points.
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ float bf16_to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}

// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide reductions through a small smem scratch (>= 33 floats); all threads get the result
template <int NT>
__device__ __forceinline__ float block_sum(float v, float* scratch) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) scratch[w] = v;
    __syncthreads();
    if (w == 0) {
        float t = (l < NT / 32) ? scratch[l] : 0.f;
        t = warp_sum(t);
        if (l == 0) scratch[32] = t;
    }
    __syncthreads();
    return scratch[32];
}
template <int NT>
__device__ __forceinline__ float block_max(float v, float* scratch) {
    v = warp_max(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) scratch[w] = v;
    __syncthreads();
    if (w == 0) {
        float t = (l < NT / 32) ? scratch[l] : 0.f;   // callers reduce non-negative values
        t = warp_max(t);
        if (l == 0) scratch[32] = t;
    }
    __syncthreads();
    return scratch[32];
}
template <int NT>
__device__ __forceinline__ int block_sum_i(int v, int* scratch) {
    v = warp_sum_i(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) scratch[w] = v;
    __syncthreads();
    if (w == 0) {
        int t = (l < NT / 32) ? scratch[l] : 0;
        t = warp_sum_i(t);
        if (l == 0) scratch[32] = t;
    }
    __syncthreads();
    return scratch[32];
}

// The reference's activation quantiser (tests/test_bitnet_core.py:139-146):
//   scale = fl32( double(max) / 127.0 )  (1.0 substituted when max < 1e-6);  q = clamp(rint(x / scale), -128, 127)
// true IEEE division, ties-to-even; a reciprocal multiply would differ by one LSB on some inputs.
// Evaluated WITHOUT fp64 (the B200 fp64 pipe would put ~1 us on the decode critical path per normalisation):
//  * 127 = 2^7 - 1, so the binary expansion of max/127 is periodic with period 7: it can never sit within 2^-53 of an
//    fp32 rounding boundary unless it is exact, hence fl32(fl64(max/127)) == fl32(max/127) == __fdiv_rn(max, 127.f);
//  * (double)max < 1e-6  <=>  max <= fl32(1e-6) = 0x358637BD (the largest float below 1e-6).
__device__ __forceinline__ float act_scale_from_absmax(float amax) {
    if (amax <= 1e-6f) amax = 1.0f;
    return __fdiv_rn(amax, 127.0f);
}
// q = clamp(rint(x / scale), -128, 127) with x / scale CORRECTLY ROUNDED (true IEEE division, as the reference's
// torch x / scale) for a divisor that is shared by a whole row: Markstein's correction.  With y = RN(1/scale) and
// q0 = RN(x*y), q = RN(q0 + RN(x - scale*q0)*y) is the correctly rounded quotient unless scale's significand is all
// ones (then: plain __fdiv_rn).  3 instructions per element instead of a ~40-instruction division subroutine — the
// quantisation pass sits on the decode critical path of every linear.
struct QDiv {
    float scale, rcp;
    int slow;
};
__device__ __forceinline__ QDiv make_qdiv(float scale) {
    QDiv d;
    d.scale = scale;
    d.rcp = __frcp_rn(scale);
    d.slow = ((__float_as_uint(scale) & 0x007fffffu) == 0x007fffffu) ? 1 : 0;
    return d;
}
__device__ __forceinline__ int quant_i8(float x, const QDiv& d) {
    float q;
    if (d.slow) {
        q = __fdiv_rn(x, d.scale);
    } else {
        const float q0 = __fmul_rn(x, d.rcp);
        const float e = __fmaf_rn(-d.scale, q0, x);
        q = __fmaf_rn(e, d.rcp, q0);
    }
    float r = rintf(q);
    r = fminf(fmaxf(r, -128.f), 127.f);
    return (int)r;
}

}  // namespace wf
