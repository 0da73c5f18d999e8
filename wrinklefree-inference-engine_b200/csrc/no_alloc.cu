// wf_bitnet_gemm_static: the bitnet_gemv / bitnet_gemm computation of wf_bitnet_gemm with NO caller workspace and no allocation.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it).  The split-K accumulators and tile counters that
// wf_bitnet_gemm keeps in the caller's workspace live here in a file-scope __device__ array of WF_STATIC_WS_BYTES (4 MiB) which
// every GPU that loads the library holds until exit.
//
// In host code the array's NAME is a host-side stand-in, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current and cached below.
//
// Every call shares the array.  A call first clears the part it uses on its own stream (the kernels also leave it zero, but a
// call that failed half-way must not poison the next one), so calls that overlap on the GPU — two streams, or two host threads on
// different streams — RACE: serialise them on one stream.
#include "common.cuh"
#include "wf_bitnet.h"
#include <mutex>
#include <string.h>

static __device__ __align__(256) unsigned char wf_static_ws[WF_STATIC_WS_BYTES];

namespace wf {

static std::mutex g_static_mu;
static void* g_static_ptr[64] = {};   // device address of wf_static_ws, per device ordinal

// the device address of the scratch array on the CURRENT device (looked up once per device)
static int static_ws_ptr(void** out) {
    int dev = -1;
    if (cuda_check(cudaGetDevice(&dev), "wf_bitnet_gemm_static: cudaGetDevice")) return WF_ERR_CUDA;
    if (dev < 0 || dev >= 64) { set_error("wf_bitnet_gemm_static: device ordinal %d out of range", dev); return WF_ERR_UNSUPPORTED; }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        if (cuda_check(cudaGetSymbolAddress(&p, wf_static_ws), "wf_bitnet_gemm_static: cudaGetSymbolAddress")) return WF_ERR_CUDA;
        if (!p || ((uintptr_t)p & 255)) { set_error("wf_bitnet_gemm_static: bad scratch address %p", p); return WF_ERR_CUDA; }
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return WF_OK;
}

}  // namespace wf

using namespace wf;

extern "C" size_t wf_static_workspace_max_bytes(void) { return (size_t)WF_STATIC_WS_BYTES; }

extern "C" int wf_bitnet_gemm_static(const uint8_t* packed_i2s, const int8_t* x, float scale, int M, int N, int K, float* y,
                                     void* stream) {
    if (M < 0 || N <= 0 || K <= 0) { set_error("wf_bitnet_gemm_static: bad shape M=%d N=%d K=%d", M, N, K); return WF_ERR_INVALID; }
    if (M == 0) return WF_OK;
    const size_t need = wf_linear_workspace_bytes(M, N, K);
    if (need > (size_t)WF_STATIC_WS_BYTES) {
        set_error("wf_bitnet_gemm_static: M=%d N=%d needs %zu bytes of scratch, the static array holds %zu "
                  "(use wf_bitnet_gemm with a caller workspace)", M, N, need, (size_t)WF_STATIC_WS_BYTES);
        return WF_ERR_UNSUPPORTED;
    }
    void* ws = nullptr;
    int rc = static_ws_ptr(&ws);
    if (rc) return rc;
    if (cuda_check(cudaMemsetAsync(ws, 0, need, (cudaStream_t)stream), "wf_bitnet_gemm_static: clear scratch")) return WF_ERR_CUDA;

    wf_linear_args a;
    memset(&a, 0, sizeof a);
    a.w = packed_i2s; a.w_layout = WF_W_I2S; a.N = N; a.K = K;
    a.q = x; a.M = M; a.ldq = K;
    a.host_scale = scale;
    a.out_mode = WF_OUT_COMPAT_F32;
    a.out = y; a.ldo = N;
    a.workspace = ws; a.workspace_bytes = (size_t)WF_STATIC_WS_BYTES;
    a.algo = 2;                        // the kernel wf_bitnet_gemv / wf_bitnet_gemm run
    return wf_ternary_linear(&a, stream);
}
