// TEST INFRASTRUCTURE ONLY — never linked into or called from the product path.
//
// Row-parallel driver around the UNMODIFIED reference kernel
// `ggml_vec_dot_i2_i8_s` (/root/reference/extern/BitNet/src/ggml-bitnet-mad.cpp:94-186,
// AVX2 branch), which is compiled from where it lies into oracle/_ref/libwf_ref.so
// by oracle/Makefile.  The driver reproduces the reference's threading model —
// an OpenMP `parallel for` over output rows (cf. the reference's embedded scalar
// kernel, src/wrinklefree_inference/kernels/native.py:60 and :84) — and the
// `y = scale * sum` epilogue of the sgl-kernel op `bitnet_gemv(packed, x_i8, scale)`
// (call sites: src/wrinklefree_inference/kernels/bitnet_patch.py:69,
// scripts/direct_inference_full.py:202).
//
// Used (a) by tests to validate oracle/wf_oracle.c against the real reference
// arithmetic and (b) by bench.py's cpu_baseline / --impl reference legs.
#include <stdint.h>
#include <stddef.h>
#include <omp.h>
#include "ggml-bitnet.h"

extern "C" {

int wfref_num_threads(void) {
    int n = 1;
#pragma omp parallel
    {
#pragma omp single
        n = omp_get_num_threads();
    }
    return n;
}

void wfref_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }

// packed: [N, K/4] I2_S rows; act: [K] int8; out[n] = scale * (float)sum_k q[n,k]*a[k]
void wfref_gemv(const uint8_t* packed, const int8_t* act, float scale, int N, int K, float* out) {
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n) {
        float s = 0.f;
        ggml_vec_dot_i2_i8_s(K, &s, 0, packed + (size_t)n * (K / 4), 0, act, 0, 1);
        out[n] = s * scale;
    }
}

// act: [M, K]; out: [M, N]
void wfref_gemm(const uint8_t* packed, const int8_t* act, float scale, int M, int N, int K, float* out) {
#pragma omp parallel for collapse(2) schedule(static)
    for (int m = 0; m < M; ++m) {
        for (int n = 0; n < N; ++n) {
            float s = 0.f;
            ggml_vec_dot_i2_i8_s(K, &s, 0, packed + (size_t)n * (K / 4), 0, act + (size_t)m * K, 0, 1);
            out[(size_t)m * N + n] = s * scale;
        }
    }
}

// raw (float-valued integer) sums, no scale: for accumulator parity checks
void wfref_gemm_raw(const uint8_t* packed, const int8_t* act, int M, int N, int K, float* out) {
    wfref_gemm(packed, act, 1.0f, M, N, K, out);
}

// quantize_i2_s wrapper: dst must hold N*K/4 + 64 bytes. Returns the f32 scale written at N*K/4.
float wfref_quantize_i2_s(const float* src, uint8_t* dst, int N, int K) {
    quantize_i2_s(src, dst, N, K, NULL);
    return *(const float*)(dst + (size_t)N * K / 4);
}

}  // extern "C"
