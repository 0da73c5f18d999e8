// C-ABI plumbing: error state, device probe, argument validation and algorithm dispatch for the
// ternary linear entry points declared in include/wf_bitnet.h.
#include "linear_common.cuh"
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

namespace wf {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int cuda_check(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return WF_OK;
    set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return WF_ERR_CUDA;
}

static thread_local long long g_launches = 0;
int launch_check(cudaError_t e, const char* what) {
    ++g_launches;
    return cuda_check(e, what);
}
long long launch_count() { return g_launches; }

int sm_count() {
    static int per_dev[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev > 63) dev = 0;
    int& n = per_dev[dev];
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

int fill_epi(const wf_linear_args* a, EpiParams* p) {
    memset(p, 0, sizeof *p);
    p->M = a->M;
    p->N = a->N;
    p->rowsum = a->rowsum;
    p->act_scale = a->act_scale;
    p->act_scale_per_token = a->act_scale_per_token;
    p->w_scale = a->w_scale;
    p->nseg = a->nseg <= 0 ? 1 : a->nseg;
    for (int i = 0; i < 4; ++i) p->seg_end[i] = a->seg_end[i];
    p->host_scale = a->host_scale;
    p->out_mode = a->out_mode;
    p->epi_flags = a->epi_flags;
    p->out_dtype = a->out_dtype;
    p->out = a->out;
    p->ldo = a->ldo;
    p->residual = (const __nv_bfloat16*)a->residual;

    // an empty batch (M == 0) is a no-op: torch hands out null data pointers for zero-size tensors
    if (!a->w || (a->M > 0 && (!a->q || !a->out))) { set_error("wf_ternary_linear: null w/q/out"); return WF_ERR_INVALID; }
    if (a->N <= 0 || a->M < 0 || a->K <= 0) { set_error("wf_ternary_linear: bad shape M=%d N=%d K=%d", a->M, a->N, a->K); return WF_ERR_INVALID; }
    if (a->K % WF_QK != 0) { set_error("in_features (%d) must be multiple of %d", a->K, WF_QK); return WF_ERR_INVALID; }
    if (a->ldq < a->K || a->ldq % 16 != 0) { set_error("wf_ternary_linear: ldq (%d) must be >= K and a multiple of 16", a->ldq); return WF_ERR_INVALID; }
    if (((uintptr_t)a->w & 15) || ((uintptr_t)a->q & 15)) { set_error("wf_ternary_linear: w and q must be 16-byte aligned"); return WF_ERR_INVALID; }
    if (a->w_layout != WF_W_I2S && a->w_layout != WF_W_TILE) { set_error("wf_ternary_linear: bad w_layout %d", a->w_layout); return WF_ERR_INVALID; }
    if (a->out_mode < WF_OUT_I32_PLUS1 || a->out_mode > WF_OUT_LINEAR) { set_error("wf_ternary_linear: bad out_mode %d", a->out_mode); return WF_ERR_INVALID; }
    if ((a->out_mode == WF_OUT_I32 || a->out_mode == WF_OUT_LINEAR) && !a->rowsum) { set_error("wf_ternary_linear: rowsum required for this out_mode"); return WF_ERR_INVALID; }
    if (a->out_mode == WF_OUT_LINEAR) {
        if (!a->act_scale || !a->w_scale) { set_error("wf_ternary_linear: act_scale and w_scale required for WF_OUT_LINEAR"); return WF_ERR_INVALID; }
        if (a->nseg < 1 || a->nseg > 4) { set_error("wf_ternary_linear: nseg must be 1..4"); return WF_ERR_INVALID; }
        if (a->out_dtype != WF_F32 && a->out_dtype != WF_BF16) { set_error("wf_ternary_linear: out_dtype must be f32 or bf16"); return WF_ERR_INVALID; }
        if ((a->epi_flags & WF_EPI_RESIDUAL) && !a->residual) { set_error("wf_ternary_linear: residual pointer missing"); return WF_ERR_INVALID; }
        if ((a->epi_flags & WF_EPI_RESIDUAL) && (a->epi_flags & WF_EPI_RELU2_MUL)) { set_error("wf_ternary_linear: RESIDUAL and RELU2_MUL are exclusive"); return WF_ERR_INVALID; }
        if ((a->epi_flags & WF_EPI_RELU2_MUL) && (a->N % 32 != 0)) { set_error("wf_ternary_linear: RELU2_MUL needs N %% 32 == 0"); return WF_ERR_INVALID; }
    } else if (a->epi_flags) {
        set_error("wf_ternary_linear: epilogue flags need WF_OUT_LINEAR");
        return WF_ERR_INVALID;
    }
    const int out_cols = (a->epi_flags & WF_EPI_RELU2_MUL) ? a->N / 2 : a->N;
    if (a->ldo < out_cols) { set_error("wf_ternary_linear: ldo (%d) < output columns (%d)", a->ldo, out_cols); return WF_ERR_INVALID; }
    return WF_OK;
}

}  // namespace wf

using namespace wf;

extern "C" const char* wf_last_error(void) { return g_err; }
extern "C" const char* wf_version(void) { return "wf_bitnet b200 r1"; }
extern "C" int wf_sm_count(void) { return sm_count(); }
extern "C" int wf_copy_d2d(void* dst, const void* src, size_t bytes, void* stream) {
    if (!dst || !src) { set_error("wf_copy_d2d: null pointer"); return WF_ERR_INVALID; }
    return cuda_check(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream), "wf_copy_d2d");
}

extern "C" int wf_bitnet_available(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { cudaGetLastError(); return 0; }
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
    return major == 10 ? 1 : 0;
}

extern "C" size_t wf_linear_workspace_bytes(int M, int N, int K) {
    if (M <= 0 || N <= 0) return 256;
    return gemm_mma_workspace_bytes(M, N, K);
}

extern "C" int wf_ternary_linear(const wf_linear_args* a, void* stream) {
    if (!a) { set_error("wf_ternary_linear: null args"); return WF_ERR_INVALID; }
    EpiParams ep;
    int rc = fill_epi(a, &ep);
    if (rc) return rc;
    if (a->M == 0) return WF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    int algo = a->algo;
    if (algo == 0) {
        if (a->M <= 8 && a->w_layout == WF_W_TILE && a->workspace && gemv_stream_fits(a->M, a->K)) algo = 1;   // else (70B down at M = 8: 8 x 28672 int8 rows) the tiled GEMM
        else if (a->M >= 128) algo = 3;   // a real dense contraction: tcgen05 int8
        else algo = 2;
    }
    switch (algo) {
        case 1: return launch_gemv_stream(a, ep, s, false);
        case 2: return launch_gemm_mma(a, ep, s);
        case 3: return launch_gemm_tc(a, ep, s);
    }
    set_error("wf_ternary_linear: unknown algo %d", a->algo);
    return WF_ERR_INVALID;
}

static int compat_call(const uint8_t* packed, const int8_t* x, float scale, int M, int N, int K, float* y, void* ws,
                       size_t ws_bytes, void* stream) {
    wf_linear_args a;
    memset(&a, 0, sizeof a);
    a.w = packed; a.w_layout = WF_W_I2S; a.N = N; a.K = K;
    a.q = x; a.M = M; a.ldq = K;
    a.host_scale = scale;
    a.out_mode = WF_OUT_COMPAT_F32;
    a.out = y; a.ldo = N;
    a.workspace = ws; a.workspace_bytes = ws_bytes;
    a.algo = 2;
    return wf_ternary_linear(&a, stream);
}

extern "C" int wf_bitnet_gemv(const uint8_t* packed_i2s, const int8_t* x, float scale, int N, int K, float* y, void* workspace,
                              size_t workspace_bytes, void* stream) {
    return compat_call(packed_i2s, x, scale, 1, N, K, y, workspace, workspace_bytes, stream);
}

extern "C" int wf_bitnet_gemm(const uint8_t* packed_i2s, const int8_t* x, float scale, int M, int N, int K, float* y,
                              void* workspace, size_t workspace_bytes, void* stream) {
    return compat_call(packed_i2s, x, scale, M, N, K, y, workspace, workspace_bytes, stream);
}
