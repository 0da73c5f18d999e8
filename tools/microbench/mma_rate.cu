// Issue-rate microbenchmark of the legacy warp-level tensor instructions on sm_100a:
// IMMA.16832 (s8), HMMA.16816 (f16 / bf16, fp32 accumulate).  Prints MACs per clock per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int KIND>
__global__ void __launch_bounds__(256) rate_kernel(int iters, float* sink, long long* clk) {
    float acc[8][4];
    int iacc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        for (int j = 0; j < 4; ++j) { acc[i][j] = 0.f; iacc[i][j] = 0; }
    uint32_t a0 = threadIdx.x * 2654435761u, a1 = a0 ^ 0x3c003c00u, a2 = a0 + 7, a3 = a1 + 11, b0 = a0 * 3, b1 = a1 * 5;
    if (KIND != 0) { a0 = 0x3c003c00u; a1 = 0x3c003c00u; a2 = 0x3c003c00u; a3 = 0x3c003c00u; b0 = 0x3c003c00u; b1 = 0x38003800u; }
    if (KIND == 2) { a0 = a1 = a2 = a3 = 0x3f803f80u; b0 = b1 = 0x3f003f00u; }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (KIND == 0)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(iacc[i][0]), "+r"(iacc[i][1]), "+r"(iacc[i][2]), "+r"(iacc[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 1)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        for (int j = 0; j < 4; ++j) s += acc[i][j] + (float)iacc[i][j];
    if (s == 123.456f) sink[0] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int KIND>
static void run(const char* name, int warps_per_sm, double macs_per_mma) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* sink; long long* clk;
    cudaMalloc(&sink, 4); cudaMalloc(&clk, sizeof(long long) * sms * 4);
    const int iters = 4096;
    const int ctas_per_sm = warps_per_sm / 8 > 0 ? warps_per_sm / 8 : 1;
    const int threads = warps_per_sm >= 8 ? 256 : warps_per_sm * 32;
    rate_kernel<KIND><<<sms * ctas_per_sm, threads>>>(16, sink, clk);
    rate_kernel<KIND><<<sms * ctas_per_sm, threads>>>(iters, sink, clk);
    cudaDeviceSynchronize();
    long long h[8];
    cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    const double mmas_per_sm = (double)iters * 8 * warps_per_sm;
    printf("%-10s warps/SM %2d: %8.0f clk  -> %.3f mma/clk/SM, %.0f MAC/clk/SM  (%s)\n", name, warps_per_sm, (double)h[0], mmas_per_sm / h[0],
           mmas_per_sm * macs_per_mma / h[0], cudaGetErrorString(cudaGetLastError()));
    cudaFree(sink); cudaFree(clk);
}

int main() {
    for (int w : {4, 8, 16}) {
        if (w == 4) { run<0>("imma.s8", 4, 16 * 8 * 32); run<1>("hmma.f16", 4, 16 * 8 * 16); run<2>("hmma.bf16", 4, 16 * 8 * 16); }
        if (w == 8) { run<0>("imma.s8", 8, 16 * 8 * 32); run<1>("hmma.f16", 8, 16 * 8 * 16); run<2>("hmma.bf16", 8, 16 * 8 * 16); }
        if (w == 16) { run<0>("imma.s8", 16, 16 * 8 * 32); run<1>("hmma.f16", 16, 16 * 8 * 16); run<2>("hmma.bf16", 16, 16 * 8 * 16); }
    }
    return 0;
}
