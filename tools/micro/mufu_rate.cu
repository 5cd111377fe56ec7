// MUFU.EX2 throughput on this GPU: nvcc -arch=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_fp16.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ unsigned ex2h2(unsigned x) { unsigned y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ float ex2_poly(float x) {
    x = fmaxf(x, -126.0f);
    const float t = __fadd_rn(x, 12582912.0f);
    const float f = __fsub_rn(x, __fsub_rn(t, 12582912.0f));
    float p = fmaf(5.517164990e-02f, f, 2.426111251e-01f);
    p = fmaf(p, f, 6.932609677e-01f);
    p = fmaf(p, f, 9.999280572e-01f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
template <int MODE>
__global__ void k(float* out, int iters) {
    float a[8];
    for (int i = 0; i < 8; ++i) a[i] = -0.001f * (threadIdx.x + i);
    unsigned h[8];
    for (int i = 0; i < 8; ++i) h[i] = 0x80008000u + threadIdx.x + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) a[i] = ex2(a[i]) - 1.0f;
            else if (MODE == 1) h[i] = ex2h2(h[i]) ^ 0x80008000u;
            else a[i] = ex2_poly(a[i]) - 1.0f;
        }
    }
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(h[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, int threads, int per_elem) {
    float* out; cudaMalloc(&out, 148 * 8 * 1024 * 4);
    const int iters = 4096, blocks = 148 * (1024 / threads > 0 ? 2048 / threads : 1);
    k<MODE><<<blocks, threads>>>(out, 16);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(out, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double n = (double)blocks * threads * iters * 8 * per_elem;
    printf("%-28s threads/block %4d: %.3f ms, %.1f results per ns per GPU = %.2f per SM per clock at %d MHz (nominal)\n", name, threads, ms,
           n / ms / 1e6, n / (ms * 1e-3) / 148 / (clk * 1e3), clk / 1000);
    cudaFree(out);
}
int main() {
    run<0>("ex2.approx.ftz.f32", 256, 1);
    run<0>("ex2.approx.ftz.f32", 64, 1);
    run<1>("ex2.approx.f16x2", 256, 2);
    run<2>("ex2 polynomial (FMA pipe)", 256, 1);
    return 0;
}
