// Fused decode path: conventions shared by the phase kernels (linear_small.cu, gemm_oneshot.cu,
// glue_rows.cu, attn.cu).
//
// Every ternary linear of the decode step leaves its RAW accumulator S[m][n] = sum_k (w+1)*a in a global
// fp32 buffer.  All values are integers with |S| <= 2*128*28672 < 2^24, so fp32 holds them exactly and
// fp32 atomic adds (RED, vectorised .v4 where possible) are exact and order-independent — split-K needs no
// "last CTA" tail.  The CONSUMER of a linear applies the reference's epilogue
//   out = (S - s_w-free Sigma a) * s_w * s_a      (scripts/direct_inference_full.py:204-206)
// together with whatever follows it in the reference step (residual add, relu^2*up, RMSNorm / sub-norm,
// per-token int8 quantisation: direct_inference_full.py:181-184,247,259,262,276,283) — "glue".  The bf16
// rounding points are the reference's (every torch op on a bf16 tensor rounds).
#pragma once
#include "common.cuh"
#include "tp_comm.cuh"
#include "wf_bitnet.h"

namespace wf {

enum GlueMode {
    GL_EMBED = 0,   // x = embed[token]                                  (direct_inference_full.py:290)
    GL_RESID = 1,   // x = bf16(hid_in + bf16(lin(acc)))                 (:276, :283)
    GL_PLAIN = 2,   // x = src (bf16)                                    (attention output, :244)
    GL_RELU2 = 3,   // x = bf16(bf16(relu(g)^2) * u), g/u = bf16(lin(acc gate/up))   (:259)
};

struct GlueParams {
    int mode, M, K;
    const __nv_bfloat16* embed;     // GL_EMBED: [V][K]
    const int32_t* tokens;          //           [M]
    const __nv_bfloat16* hid_in;    // GL_RESID: [M][K]
    __nv_bfloat16* hid_out;         // GL_EMBED / GL_RESID: [M][K] (may be null)
    const __nv_bfloat16* src;       // GL_PLAIN: [M][K]
    const float* acc;               // GL_RESID: [M][acc_ld]; GL_RELU2: gate|up interleaved in 16-row blocks
    int acc_ld;
    // tensor parallel, batched path: the accumulator is reduce-SCATTERED — column k lives on rank k / acc_own_cols and is
    // pulled from that rank's buffer over NVLink (0 = the local buffer holds every column)
    const float* acc_peer[TP_MAX];
    int acc_own_cols;
    const int32_t* acc_rowsum;      // [M] Sigma a of the linear that produced acc
    const float* acc_ascale;        // [M] its per-token activation scale
    const float* acc_wscale;        // device: [1] (RESID) or [2] gate, up (RELU2)
    const __nv_bfloat16* gamma;     // [K] norm weight
    float eps;
    float* zero0; int zero0_n;      // fp32 ranges this phase must clear (accumulators consumed one phase ago)
    float* zero1; int zero1_n;
    // outputs of the standalone glue kernel (large batch); the small-batch linear keeps them in smem
    int8_t* q_out;                  // [M][K]
    float* scale_out;               // [M]
    int32_t* rowsum_out;            // [M]
    __nv_bfloat16* y_out;           // [M][K] normalised activations (final norm -> LM head), optional
    // tensor parallel: the source of this phase was produced by an exchange — wait for every rank's flag first
    int wait_flag;                  // flag id, -1 = none
    TpComm tp;
    unsigned long long* trace;      // timeline taps (WF_TRACE=1 debugging): 8 globaltimer stamps per launch, or null
};
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// where a linear's partial sums go: the local accumulator, or (row-parallel o / down under TP) the same accumulator
// on every rank — the all-reduce is the RED itself
struct TpDest {
    int n;                          // destinations (1 = local only)
    int own_cols;                   // > 0: reduce-scatter — column n is added on rank n / own_cols only (acc[] indexed by rank)
    float* acc[TP_MAX];
    int signal_flag;                // flag id published when the whole grid has finished its REDs, -1 = none
};

__device__ __forceinline__ float4 ldcg_f4(const float* p) {
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ldcg_u4(const void* p) {
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float ldcg_f(const float* p) {
    float r;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
// exact: |S| < 2^24.  The reference's rounding sequence (scripts/direct_inference_full.py:202-206; see linear_value in
// linear_common.cuh): fl32( fl32( fl32(S*s_w) - fl32(s_w*Sigma a) ) * s_a ) — bit-identical to the CPU path's fp32 value.
__device__ __forceinline__ float lin_value(float S, int rowsum, float ws, float sa) {
    return __fmul_rn(__fsub_rn(__fmul_rn(S, ws), __fmul_rn(ws, (float)rowsum)), sa);
}
__device__ __forceinline__ void red_add_f32(float* p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add_v4_f32(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// named barrier over the first NT threads of the CTA (warps 0..NT/32-1)
template <int ID, int NT>
__device__ __forceinline__ void group_bar() {
    asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(NT) : "memory");
}

// block reductions over a thread GROUP of NT threads (ids 0..NT-1), synchronised with named barrier ID
template <int ID, int NT>
__device__ __forceinline__ float group_sum(float v, float* scratch) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    group_bar<ID, NT>();
    if (l == 0) scratch[w] = v;
    group_bar<ID, NT>();
    float t = (l < NT / 32) ? scratch[l] : 0.f;
    return warp_sum(t);
}
template <int ID, int NT>
__device__ __forceinline__ float group_max(float v, float* scratch) {
    v = warp_max(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    group_bar<ID, NT>();
    if (l == 0) scratch[w] = v;
    group_bar<ID, NT>();
    float t = (l < NT / 32) ? scratch[l] : 0.f;   // non-negative values
    return warp_max(t);
}
template <int ID, int NT>
__device__ __forceinline__ int group_sum_i(int v, int* scratch) {
    v = warp_sum_i(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    group_bar<ID, NT>();
    if (l == 0) scratch[w] = v;
    group_bar<ID, NT>();
    int t = (l < NT / 32) ? scratch[l] : 0;
    return warp_sum_i(t);
}

// Single-barrier reductions over the NT-thread group: every warp drops its partial into its OWN scratch slot, one
// barrier, every thread folds the NT/32 partials itself.  The caller gives each reduction of a row a distinct scratch
// range, so no second "scratch free" barrier is needed (the next reuse is separated by the following barriers).
template <int ID, int NT>
__device__ __forceinline__ float group_sum1(float v, float* scratch) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    group_bar<ID, NT>();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < NT / 32; ++w) t += scratch[w];
    return t;
}
template <int ID, int NT>
__device__ __forceinline__ float group_max1(float v, float* scratch) {
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    group_bar<ID, NT>();
    float t = 0.f;   // non-negative values
#pragma unroll
    for (int w = 0; w < NT / 32; ++w) t = fmaxf(t, scratch[w]);
    return t;
}
template <int ID, int NT>
__device__ __forceinline__ int group_sum_i1(int v, int* scratch) {
    v = warp_sum_i(v);
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    group_bar<ID, NT>();
    int t = 0;
#pragma unroll
    for (int w = 0; w < NT / 32; ++w) t += scratch[w];
    return t;
}

// Norm weights do not depend on the previous kernel: pull them towards the SM before the dependency wait.
template <int NT>
__device__ __forceinline__ void glue_prefetch_gamma(const GlueParams& p) {
    const char* g = reinterpret_cast<const char*>(p.gamma);
    for (int o = threadIdx.x * 128; o < p.K * 2; o += NT * 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(g + o));
}

// where the int8 codes of a row go: plain [K]; the dp4a staging order of gemv_stream.cu; or the chunk-major image the batched
// tail kernel copies straight into its shared-memory ring — [K/512][64 rows][512 + 16 B skew] (layer_tail_batch.cu), qdst =
// base + row * QROW_BYTES
enum { QMAP_PLAIN = 0, QMAP_PERM = 1, QMAP_CHUNKED = 2 };
constexpr int QROW_BYTES = 4 * 128 + 16;
constexpr int QCHUNK_BYTES = 64 * QROW_BYTES;

// One activation row through the glue: source -> (residual / relu^2*up) -> RMSNorm -> per-token int8.
// Executed by the NT-thread group (thread ids 0..NT-1).  ys: K bf16 of scratch (shared); qdst: K int8
// (shared or global, 8-byte aligned); fs/is: >= 32 floats / ints of shared scratch.
// Returns the activation scale and Sigma q through the references (identical in every thread).
// Flag-in-data hand-off (HandOff, optional; the batch-1 tail kernel): the source of the row (the previous phase's accumulator or bf16
// output) is being written by OTHER CTAs of the same persistent grid.  Instead of a grid barrier followed by a load, the consumer polls
// the words themselves: the owner of a row stores plain data over a sentinel (0xffffffff: a NaN no integer-valued accumulator and no
// bf16 pair produced by the epilogue can equal), so one L2 hop replaces release fence + atomic arrival + acquire poll + load.
constexpr uint32_t WF_HANDOFF_SENTINEL = 0xffffffffu;
struct HandOff {
    bool poll;                        // the phase's source (acc for RESID, src for PLAIN) is a hand-off buffer
    bool use_prev;                    // RESID: activation scale / rowsum of the producing linear come from the caller (computed locally
    float prev_scale; int prev_rowsum;   //       by every CTA one phase ago) instead of global memory written by CTA 0
    const __nv_bfloat16* hid_sm_in;   // RESID: residual row in shared memory (kept by this CTA two phases ago) instead of global hid_in
    __nv_bfloat16* hid_sm_out;        // RESID: keep the new residual row here
    int* err;
    int tag;                          // error code reported when a poll times out (identifies the phase)
};
// The poll must be a STRONG load (ld.relaxed.gpu): a weak ld.global.cg may be served again and again from a copy of the line in the
// near L2 partition and never observe the owner's store (measured: every hand-off timed out with .cg loads).
__device__ __forceinline__ uint4 ld_relaxed_gpu_u4(const void* p) {
    uint4 r;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ uint4 handoff_poll_u4(const void* p, int* err, int tag = 3) {
    uint4 v = ld_relaxed_gpu_u4(p);
    if (v.x == WF_HANDOFF_SENTINEL || v.y == WF_HANDOFF_SENTINEL || v.z == WF_HANDOFF_SENTINEL || v.w == WF_HANDOFF_SENTINEL) {
        if (err && ld_relaxed_gpu(err)) return v;            // after one timeout nothing waits any more: the step is lost anyway
        const unsigned long long t0 = spin_now();
        do {
            v = ld_relaxed_gpu_u4(p);
            if (spin_now() - t0 > WF_SPIN_BUDGET_NS) { if (err) atomicExch(err, tag); break; }
        } while (v.x == WF_HANDOFF_SENTINEL || v.y == WF_HANDOFF_SENTINEL || v.z == WF_HANDOFF_SENTINEL || v.w == WF_HANDOFF_SENTINEL);
    }
    return v;
}
__device__ __forceinline__ float4 handoff_poll_f4(const float* p, int* err, int tag = 3) {
    const uint4 v = handoff_poll_u4(p, err, tag);
    return make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w));
}

// Cross-CTA part of the three row reductions of glue_row.  NoXR: the row belongs to one CTA.  ClusterXR: the row is cut into contiguous
// slices over the CTAs of a thread-block cluster (long rows at small batch: one 512-thread CTA per 28672-element row is 10 us); a CTA
// leaves its partial in its own shared memory, one cluster barrier, everybody sums the partials in rank order over DSMEM (the same
// total in every CTA).  The fp32 sum of squares is then cut differently from the one-CTA form — like between the one-CTA forms of
// different block sizes, the bf16 rounding of 1/rms absorbs that; absmax and the int32 row sum are exact.
struct NoXR {
    __device__ __forceinline__ float sum(float v) const { return v; }
    __device__ __forceinline__ float max(float v) const { return v; }
    __device__ __forceinline__ int sumi(int v) const { return v; }
};
struct ClusterXR {
    float* slot;     // this CTA's shared float[4] (one word per reduction: no reuse, so no second barrier)
    int n;           // CTAs in the cluster
    __device__ __forceinline__ void sync() const {
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    __device__ __forceinline__ float peer(const float* q, int r) const {
        uint32_t ra; float v;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(q)), "r"(r));
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
        return v;
    }
    __device__ __forceinline__ float sum(float v) const {
        if (threadIdx.x == 0) slot[0] = v;
        sync();
        float t = 0.f;
        for (int r = 0; r < n; ++r) t += peer(slot, r);
        return t;
    }
    __device__ __forceinline__ float max(float v) const {
        if (threadIdx.x == 0) slot[1] = v;
        sync();
        float t = 0.f;
        for (int r = 0; r < n; ++r) t = fmaxf(t, peer(slot + 1, r));
        return t;
    }
    __device__ __forceinline__ int sumi(int v) const {
        if (threadIdx.x == 0) slot[2] = __int_as_float(v);
        sync();
        int t = 0;
        for (int r = 0; r < n; ++r) t += __float_as_int(peer(slot + 2, r));
        return t;
    }
};
template <int ID, int NT, class XR = NoXR>
__device__ __forceinline__ void glue_row(const GlueParams& p, int m, __nv_bfloat16* ys, int8_t* qdst, float* fs, int* is,
                                         bool write_side, float& scale_ret, int& rowsum_ret, unsigned long long* tr = nullptr,
                                         int qmap = QMAP_PLAIN, int kb = 0, int ke = -1, XR xr = XR()) {
    // [kb, ke): the slice of the row this CTA works on (the whole row by default); ys holds the slice
    const int K = p.K;
    const int KE = ke < 0 ? K : ke;
    __nv_bfloat16* ysb = ys - kb;
    const int tid = threadIdx.x;
    float ss = 0.f;
    // ---- pass 1: x (bf16) -> ys, sum of squares.  All global loads of GU iterations are issued before the first use:
    //      this pass is a chain of L2 round trips (~0.7 us each) on the layer's critical path.
    constexpr int GU = 2;
    if (p.mode == GL_EMBED || p.mode == GL_PLAIN) {
        const __nv_bfloat16* row = (p.mode == GL_EMBED) ? p.embed + (size_t)p.tokens[m] * K : p.src + (size_t)m * K;
        for (int k0 = kb + tid * 8; k0 < KE; k0 += NT * 8 * GU) {
            uint4 t[GU];
#pragma unroll
            for (int u = 0; u < GU; ++u) {
                const int k = k0 + u * NT * 8;
                if (k < KE) t[u] = ldcg_u4(row + k);
            }
#pragma unroll
            for (int u = 0; u < GU; ++u) {
                const int k = k0 + u * NT * 8;
                if (k < KE) {
                    *reinterpret_cast<uint4*>(ysb + k) = t[u];
                    if (write_side && p.hid_out && p.mode == GL_EMBED) *reinterpret_cast<uint4*>(p.hid_out + (size_t)m * K + k) = t[u];
                    const uint32_t w[4] = {t[u].x, t[u].y, t[u].z, t[u].w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const float a = bf16lo(w[i]), b = bf16hi(w[i]); ss += a * a + b * b; }
                }
            }
        }
    } else if (p.mode == GL_RESID) {
        const int rs = p.acc_rowsum[m];
        const float sa = p.acc_ascale[m], ws = p.acc_wscale[0];
        const float* arow = p.acc + (size_t)m * p.acc_ld;
        const __nv_bfloat16* hrow = p.hid_in + (size_t)m * K;
        for (int k0 = kb + tid * 8; k0 < KE; k0 += NT * 8 * GU) {
            uint4 t[GU];
            float4 a0[GU], a1[GU];
#pragma unroll
            for (int u = 0; u < GU; ++u) {
                const int k = k0 + u * NT * 8;
                if (k < KE) {
                    const float* ak = (p.acc_own_cols > 0 ? p.acc_peer[k / p.acc_own_cols] + (size_t)m * p.acc_ld : arow) + k;
                    t[u] = ldcg_u4(hrow + k); a0[u] = ldcg_f4(ak); a1[u] = ldcg_f4(ak + 4);
                }
            }
#pragma unroll
            for (int u = 0; u < GU; ++u) {
                const int k = k0 + u * NT * 8;
                if (k < KE) {
                    const float S[8] = {a0[u].x, a0[u].y, a0[u].z, a0[u].w, a1[u].x, a1[u].y, a1[u].z, a1[u].w};
                    const uint32_t hw[4] = {t[u].x, t[u].y, t[u].z, t[u].w};
                    uint32_t o[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        // hidden + proj, both bf16 tensors: one rounding of the projection, one of the sum
                        const float x0 = bf16_round(bf16lo(hw[i]) + bf16_round(lin_value(S[2 * i], rs, ws, sa)));
                        const float x1 = bf16_round(bf16hi(hw[i]) + bf16_round(lin_value(S[2 * i + 1], rs, ws, sa)));
                        ss += x0 * x0 + x1 * x1;
                        o[i] = pack_bf16x2(x0, x1);
                    }
                    const uint4 ov = make_uint4(o[0], o[1], o[2], o[3]);
                    *reinterpret_cast<uint4*>(ysb + k) = ov;
                    if (write_side && p.hid_out) *reinterpret_cast<uint4*>(p.hid_out + (size_t)m * K + k) = ov;
                }
            }
        }
    } else {  // GL_RELU2 (4 independent 16-byte loads per iteration already; batching two iterations would spill)
        const int rs = p.acc_rowsum[m];
        const float sa = p.acc_ascale[m], wg = p.acc_wscale[0], wu = p.acc_wscale[1];
        const float* arow = p.acc + (size_t)m * p.acc_ld;
        for (int k = kb + tid * 8; k < KE; k += NT * 8) {
            const int ng = (k >> 4) * 32 + (k & 15);   // gate rows of this 8-group; up rows at +16
            const float4 g0 = ldcg_f4(arow + ng), g1 = ldcg_f4(arow + ng + 4);
            const float4 u0 = ldcg_f4(arow + ng + 16), u1 = ldcg_f4(arow + ng + 20);
            const float G[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float U[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
            float x[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float g = bf16_round(lin_value(G[i], rs, wg, sa));
                const float uu = bf16_round(lin_value(U[i], rs, wu, sa));
                const float r = fmaxf(g, 0.f);
                x[i] = bf16_round(bf16_round(r * r) * uu);
                ss += x[i] * x[i];
            }
            *reinterpret_cast<uint4*>(ysb + k) =
                make_uint4(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]), pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
        }
    }
    static_assert(NT / 32 <= 16, "scratch layout: 16 slots per reduction");
    if (tr) tr[0] = gtimer();
    ss = xr.sum(group_sum1<ID, NT>(ss, fs));
    if (tr) tr[1] = gtimer();
    const float var = ss / (float)K;
    const float r = bf16_round(1.0f / sqrtf(var + p.eps));
    // ---- pass 2: y = bf16(bf16(x*r)*gamma) in place, absmax
    float amax = 0.f;
    for (int k = kb + tid * 8; k < KE; k += NT * 8) {
        const uint4 t = *reinterpret_cast<const uint4*>(ysb + k);
        const uint4 gq = *reinterpret_cast<const uint4*>(p.gamma + k);
        const uint32_t xw[4] = {t.x, t.y, t.z, t.w}, gw[4] = {gq.x, gq.y, gq.z, gq.w};
        uint32_t o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float y0 = bf16_round(bf16_round(bf16lo(xw[i]) * r) * bf16lo(gw[i]));
            const float y1 = bf16_round(bf16_round(bf16hi(xw[i]) * r) * bf16hi(gw[i]));
            amax = fmaxf(amax, fmaxf(fabsf(y0), fabsf(y1)));
            o[i] = pack_bf16x2(y0, y1);
        }
        const uint4 ov = make_uint4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<uint4*>(ysb + k) = ov;
        if (write_side && p.y_out) *reinterpret_cast<uint4*>(p.y_out + (size_t)m * K + k) = ov;
    }
    if (tr) tr[2] = gtimer();
    amax = xr.max(group_max1<ID, NT>(amax, fs + 16));
    if (tr) tr[3] = gtimer();
    const float scale = act_scale_from_absmax(amax);
    const QDiv qd = make_qdiv(scale);
    // ---- pass 3: int8 codes (each thread re-reads exactly what it wrote)
    int psum = 0;
    for (int k = kb + tid * 8; k < KE; k += NT * 8) {
        const uint4 t = *reinterpret_cast<const uint4*>(ysb + k);
        const uint32_t yw[4] = {t.x, t.y, t.z, t.w};
        uint32_t w0 = 0, w1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int q0 = quant_i8(bf16lo(yw[i]), qd), q1 = quant_i8(bf16hi(yw[i]), qd);
            psum += q0 + q1;
            const uint32_t pr = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8);
            if (i < 2) w0 |= pr << (16 * i); else w1 |= pr << (16 * (i - 2));
        }
        if (qmap == QMAP_PLAIN) {
            *reinterpret_cast<uint2*>(qdst + k) = make_uint2(w0, w1);
        } else if (qmap == QMAP_CHUNKED) {
            *reinterpret_cast<uint2*>(qdst + (size_t)(k >> 9) * QCHUNK_BYTES + (k & 511)) = make_uint2(w0, w1);
        } else {
            // dp4a staging order (gemv_stream.cu): word (8g + 4h + t) of a 128-k block goes to word (8t + 2g + h)
            const int w = k >> 2, wi = w & 31;
            const int d = (w & ~31) + 8 * (wi & 3) + 2 * (wi >> 3) + ((wi >> 2) & 1);
            reinterpret_cast<uint32_t*>(qdst)[d] = w0;
            reinterpret_cast<uint32_t*>(qdst)[d + 8] = w1;   // w + 1: t + 1
        }
    }
    if (tr) tr[4] = gtimer();
    const int total = xr.sumi(group_sum_i1<ID, NT>(psum, is));
    scale_ret = scale;
    rowsum_ret = total;
}

// The same glue with the row held in REGISTERS between the passes (no bf16 scratch round trips through shared memory): each
// thread owns the 8-element granules k = (it*NT + tid)*8, it < MAXIT — the same thread -> k map and the same per-thread
// summation order as glue_row, so every reduction (and with it every output bit) is identical.  All global loads of the row
// (source, accumulator, residual, gamma) are issued before the first use.  Modes EMBED / PLAIN / RESID, local accumulator only;
// requires K <= NT * 8 * MAXIT.
//
template <int ID, int NT, int MAXIT>
__device__ __forceinline__ void glue_row_regs(const GlueParams& p, int m, int8_t* qdst, float* fs, int* is, bool write_side,
                                              float& scale_ret, int& rowsum_ret, int qmap, const HandOff* ho = nullptr) {
    const int K = p.K;
    const int tid = threadIdx.x;
    uint4 xv[MAXIT], gv[MAXIT];
    float ss = 0.f;
#pragma unroll
    for (int it = 0; it < MAXIT; ++it) {
        const int k = (it * NT + tid) * 8;
        if (k < K) gv[it] = *reinterpret_cast<const uint4*>(p.gamma + k);
    }
    if (p.mode == GL_RESID) {
        const bool prev = ho && ho->use_prev;
        const int rs = prev ? ho->prev_rowsum : p.acc_rowsum[m];
        const float sa = prev ? ho->prev_scale : p.acc_ascale[m], ws = p.acc_wscale[0];
        const float* arow = p.acc + (size_t)m * p.acc_ld;
        const __nv_bfloat16* hrow = p.hid_in + (size_t)m * K;
        float4 a0[MAXIT], a1[MAXIT];
#pragma unroll
        for (int it = 0; it < MAXIT; ++it) {
            const int k = (it * NT + tid) * 8;
            if (k < K) {
                xv[it] = (ho && ho->hid_sm_in) ? *reinterpret_cast<const uint4*>(ho->hid_sm_in + k) : ldcg_u4(hrow + k);
                if (ho && ho->poll) { a0[it] = handoff_poll_f4(arow + k, ho->err, ho->tag + 1000 * (k >> 3)); a1[it] = handoff_poll_f4(arow + k + 4, ho->err, ho->tag + 10 + 1000 * (k >> 3)); }
                else { a0[it] = ldcg_f4(arow + k); a1[it] = ldcg_f4(arow + k + 4); }
            }
        }
#pragma unroll
        for (int it = 0; it < MAXIT; ++it) {
            const int k = (it * NT + tid) * 8;
            if (k < K) {
                const float S[8] = {a0[it].x, a0[it].y, a0[it].z, a0[it].w, a1[it].x, a1[it].y, a1[it].z, a1[it].w};
                const uint32_t hw[4] = {xv[it].x, xv[it].y, xv[it].z, xv[it].w};
                uint32_t o[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float x0 = bf16_round(bf16lo(hw[i]) + bf16_round(lin_value(S[2 * i], rs, ws, sa)));
                    const float x1 = bf16_round(bf16hi(hw[i]) + bf16_round(lin_value(S[2 * i + 1], rs, ws, sa)));
                    ss += x0 * x0 + x1 * x1;
                    o[i] = pack_bf16x2(x0, x1);
                }
                xv[it] = make_uint4(o[0], o[1], o[2], o[3]);
                if (write_side && p.hid_out) *reinterpret_cast<uint4*>(p.hid_out + (size_t)m * K + k) = xv[it];
                if (ho && ho->hid_sm_out) *reinterpret_cast<uint4*>(ho->hid_sm_out + k) = xv[it];
            }
        }
    } else {
        const __nv_bfloat16* row = (p.mode == GL_EMBED) ? p.embed + (size_t)p.tokens[m] * K : p.src + (size_t)m * K;
#pragma unroll
        for (int it = 0; it < MAXIT; ++it) {
            const int k = (it * NT + tid) * 8;
            if (k < K) xv[it] = (ho && ho->poll) ? handoff_poll_u4(row + k, ho->err, ho->tag + 1000 * (k >> 3)) : ldcg_u4(row + k);
        }
#pragma unroll
        for (int it = 0; it < MAXIT; ++it) {
            const int k = (it * NT + tid) * 8;
            if (k < K) {
                if (write_side && p.hid_out && p.mode == GL_EMBED) *reinterpret_cast<uint4*>(p.hid_out + (size_t)m * K + k) = xv[it];
                const uint32_t w[4] = {xv[it].x, xv[it].y, xv[it].z, xv[it].w};
#pragma unroll
                for (int i = 0; i < 4; ++i) { const float a = bf16lo(w[i]), b = bf16hi(w[i]); ss += a * a + b * b; }
            }
        }
    }
    ss = group_sum1<ID, NT>(ss, fs);
    const float r = bf16_round(1.0f / sqrtf(ss / (float)K + p.eps));
    float amax = 0.f;
#pragma unroll
    for (int it = 0; it < MAXIT; ++it) {
        const int k = (it * NT + tid) * 8;
        if (k < K) {
            const uint32_t xw[4] = {xv[it].x, xv[it].y, xv[it].z, xv[it].w}, gw[4] = {gv[it].x, gv[it].y, gv[it].z, gv[it].w};
            uint32_t o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float y0 = bf16_round(bf16_round(bf16lo(xw[i]) * r) * bf16lo(gw[i]));
                const float y1 = bf16_round(bf16_round(bf16hi(xw[i]) * r) * bf16hi(gw[i]));
                amax = fmaxf(amax, fmaxf(fabsf(y0), fabsf(y1)));
                o[i] = pack_bf16x2(y0, y1);
            }
            xv[it] = make_uint4(o[0], o[1], o[2], o[3]);
            if (write_side && p.y_out) *reinterpret_cast<uint4*>(p.y_out + (size_t)m * K + k) = xv[it];
        }
    }
    amax = group_max1<ID, NT>(amax, fs + 16);
    const float scale = act_scale_from_absmax(amax);
    const QDiv qd = make_qdiv(scale);
    int psum = 0;
#pragma unroll
    for (int it = 0; it < MAXIT; ++it) {
        const int k = (it * NT + tid) * 8;
        if (k < K) {
            const uint32_t yw[4] = {xv[it].x, xv[it].y, xv[it].z, xv[it].w};
            uint32_t w0 = 0, w1 = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int q0 = quant_i8(bf16lo(yw[i]), qd), q1 = quant_i8(bf16hi(yw[i]), qd);
                psum += q0 + q1;
                const uint32_t pr = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8);
                if (i < 2) w0 |= pr << (16 * i); else w1 |= pr << (16 * (i - 2));
            }
            if (qmap == QMAP_PLAIN) {
                *reinterpret_cast<uint2*>(qdst + k) = make_uint2(w0, w1);
            } else if (qmap == QMAP_CHUNKED) {
                *reinterpret_cast<uint2*>(qdst + (size_t)(k >> 9) * QCHUNK_BYTES + (k & 511)) = make_uint2(w0, w1);
            } else {   // dp4a staging order (see glue_row)
                const int w = k >> 2, wi = w & 31;
                const int d = (w & ~31) + 8 * (wi & 3) + 2 * (wi >> 3) + ((wi >> 2) & 1);
                reinterpret_cast<uint32_t*>(qdst)[d] = w0;
                reinterpret_cast<uint32_t*>(qdst)[d + 8] = w1;
            }
        }
    }
    const int total = group_sum_i1<ID, NT>(psum, is);
    scale_ret = scale;
    rowsum_ret = total;
}

// grid-strided clearing of the accumulators a phase is responsible for (by the NT-thread group of every CTA)
template <int NT>
__device__ __forceinline__ void glue_zero_duty(const GlueParams& p) {
    const int gtid = blockIdx.x * NT + threadIdx.x, gstride = gridDim.x * NT;
    if (p.zero0)
        for (int i = gtid * 4; i < p.zero0_n; i += gstride * 4) *reinterpret_cast<float4*>(p.zero0 + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.zero1)
        for (int i = gtid * 4; i < p.zero1_n; i += gstride * 4) *reinterpret_cast<float4*>(p.zero1 + i) = make_float4(0.f, 0.f, 0.f, 0.f);
}

// launchers (host)
struct SmallLinArgs {
    GlueParams g;            // g.K = length of the (full) activation vector the prologue normalises / quantises
    const uint8_t* w;        // b200_tile_v1 [N][K]
    int N, K;                // K = columns of THIS rank's weight slice (== g.K without tensor parallelism)
    int k_off;               // first activation column of the slice
    TpDest dest;             // fp32 [M][acc_ld] accumulators, zero on entry
    int acc_ld;
    float* lin_ascale;       // [M] out: activation scale of THIS linear's input
    int32_t* lin_rowsum;     // [M] out: Sigma a
};
int launch_linear_small(const SmallLinArgs& a, cudaStream_t s, bool pdl);
bool linear_small_fits(int M, int K);
int launch_glue_rows(const GlueParams& g, cudaStream_t s, bool pdl);
struct OneshotArgs {
    const uint8_t* w;
    int N, K;
    const int8_t* q;         // [M][ldq], the slice starts at column k_off
    int M, ldq, k_off;
    TpDest dest;
    int acc_ld;
    TpComm tp;
};
// relu^2(gate)*up of the local ffn slice, stored into h_full of every rank (tensor parallel phase D')
struct HExchangeArgs {
    const float* acc;        // [M][2*ffn_local] gate|up interleaved in 16-row blocks
    int M, ffn_local, ffn_full, col0;
    const int32_t* rowsum;
    const float* ascale;
    const float* wscale;     // device [2]
    __nv_bfloat16* dst[TP_MAX];
    int n_dst, signal_flag;
    TpComm tp;
};
int launch_h_exchange(const HExchangeArgs& a, cudaStream_t s, bool pdl);
int launch_gemm_oneshot(const OneshotArgs& a, cudaStream_t s, bool pdl);
// streaming batched GEMM for 2..32 rows on large matrices (gemm_stream.cu); same contract as the one-shot GEMM
bool gemm_stream_fits(int M, int N, int K);
int launch_gemm_stream(const OneshotArgs& a, cudaStream_t s, bool pdl);
void gemm_stream_init_attrs();
void fused_init_attrs();

// persistent layer-tail kernel (layer_tail.cu): all linear phases between two attention kernels in one launch
constexpr int LT_MAX_PHASES = 4;
struct TailPhaseArgs {
    GlueParams g;            // the phase's input glue (M is filled in by the launcher)
    const uint8_t* w;        // b200_tile_v1 [N][K]
    int N, K;
    bool relu2;              // gate|up: the epilogue writes h = relu(gate)^2 * up (bf16) instead of raw accumulators
    float* acc;              // [M][acc_ld] raw fp32 accumulators (plain stores: the CTA owns every k of its rows)
    int acc_ld;
    float* lin_ascale;       // [M] out
    int32_t* lin_rowsum;     // [M] out
    __nv_bfloat16* h_out;    // relu2: [M][h_ld]
    int h_ld;
    const float* h_wscale;   // relu2: device [2]
};
struct TailArgs {
    int nph, M;
    TailPhaseArgs ph[LT_MAX_PHASES];
    unsigned long long* bar; // [LT_MAX_PHASES] arrival counters of THIS launch site (zero-initialised once, monotonic)
    int* epoch;              // device counter of tail-path decode steps: barrier target = grid * epoch
    int bump_epoch;          // this launch (the first of a step, single phase) increments the epoch
    int lam;                 // batch 1: flag-in-data hand-offs between the phases instead of grid barriers (the engine keeps the sentinels armed)
    int* err;                // device error word (bounded spins)
    unsigned long long* trace;   // WF_TRACE timeline taps of this launch site (32 slots) or null
};
int launch_layer_tail(const TailArgs& a, cudaStream_t s, bool pdl);          // M <= 8  (layer_tail.cu)
bool layer_tail_fits(int M, int Kmax, int max_rb_per_cta);
void layer_tail_init_attrs();
int launch_layer_tail_batch(const TailArgs& a, cudaStream_t s, bool pdl);    // 8 < M <= 64 (layer_tail_batch.cu); bar: 2 * LT_MAX_PHASES slots
bool layer_tail_batch_fits(int M, int Kmax);
void layer_tail_batch_init_attrs();

struct FusedAttnArgs {
    void* pool;
    int layers, layer, Hkv, page_size;
    const int32_t* block_table;
    int max_blocks;
    const int32_t* seq_lens;
    const int32_t* positions;
    int B, H;
    float scale;
    void* out;                 // bf16 [B][out_ld], this rank's heads start at column out_col0
    void* out_peer[TP_MAX];    // the same buffer on every rank (n_out entries; [0] unused when n_out <= 1)
    int n_out, out_ld, out_col0, signal_flag;
    TpComm tp;
    const float* qkv_acc;
    int qkv_ld;
    const int32_t* qkv_rowsum;
    const float* qkv_ascale;
    const float* qkv_wscale;   // device [3]
    const void *cos_tab, *sin_tab;
    int max_pos;
    int* merge_cnt;            // [B][Hkv], zero
    unsigned long long* trace; // WF_TRACE timeline taps or null
    float* zero0;
    int zero0_n;
    int max_seq_hint;
    void* ws;
    size_t ws_bytes;
    float* kv_scales;          // int8 pool: scale region (null = bf16 pool)
};
int attn_decode_fused(const FusedAttnArgs& a, cudaStream_t s, bool pdl);

}  // namespace wf
