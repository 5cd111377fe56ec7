// K14: FreeInit frequency-domain re-initialisation (utils/freeinit_utils.py).
//
//   freq_mix_3d (freeinit_utils.py:7-47):  y = Re IFFT3(ifftshift(fftshift(FFT3 x)*LPF + fftshift(FFT3 n)*(1-LPF)))
// By linearity this equals  y = n + Re IFFT3( LPF' * FFT3(x - n) )  with LPF'[k] = LPF[(k + L/2) mod L] per axis,
// so ONE forward and ONE inverse transform are needed instead of the reference's three (cuFFT) transforms.
// H = 40 is not a power of two; the three axes are tiny (T<=32, H=40, W=64), so each axis is a direct DFT against a
// twiddle table held in shared memory (separable, O(N*(T+H+W)) — 2.5 GFLOP for config 5, far below the HBM time).
//   pass A (per (volume,t) slice): d = x-n ; DFT over W, then over H  -> Z [NV,T,H,W] complex in `work`
//   pass B (per (volume,h,w) column): DFT over T, multiply by LPF', inverse DFT over T (in place)
//   pass C (per slice): inverse DFT over H then W (real part), y = n + result/(T*H*W)
// Filters (get_freq_filter, freeinit_utils.py:51-157) are evaluated in double like the reference's python floats
// and stored as fp32.
#include "common.cuh"

namespace moca {

__global__ void freeinit_filter_kernel(int type, int n, double d_s, double d_t, int T, int H, int W, float* __restrict__ lpf) {
    const int total = T * H * W;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int w = i % W, h = (i / W) % H, t = i / (W * H);
    float v = 0.0f;
    if (d_s != 0.0 && d_t != 0.0) {
        if (type == 3) {                                              // box, freeinit_utils.py:137-157
            // python round() is round-half-to-even == rint
            const int thr_s = (int)rint((double)(H / 2) * d_s);
            const int thr_t = (int)rint((double)(T / 2) * d_t);
            const int cf = T / 2, cr = H / 2, cc = W / 2;
            const bool in = t >= cf - thr_t && t < cf + thr_t && h >= cr - thr_s && h < cr + thr_s && w >= cc - thr_s &&
                            w < cc + thr_s;
            v = in ? 1.0f : 0.0f;
        } else {
            const double a = (d_s / d_t) * (2.0 * t / T - 1.0);
            const double b = 2.0 * h / H - 1.0;
            const double c = 2.0 * w / W - 1.0;
            const double d2 = a * a + b * b + c * c;
            if (type == 0) v = (float)exp(-1.0 / (2.0 * d_s * d_s) * d2);              // gaussian :73-91
            else if (type == 1) v = (float)(1.0 / (1.0 + pow(d2 / (d_s * d_s), (double)n)));   // butterworth :94-113
            else v = d2 <= d_s * 2.0 ? 1.0f : 0.0f;                                    // ideal :116-134
        }
    }
    lpf[i] = v;
}

__device__ __forceinline__ void fill_twiddle(float2* tw, int L) {
    for (int j = threadIdx.x; j < L; j += blockDim.x) {
        double s, c;
        sincospi(2.0 * (double)j / (double)L, &s, &c);
        tw[j] = make_float2((float)c, (float)s);
    }
}

// pass A: forward W then H.  smem: d[H*W] float, c1[H*W] float2, twW[W], twH[H]
__global__ void __launch_bounds__(256) freeinit_fwd_hw_kernel(const float* __restrict__ x, const float* __restrict__ nz,
                                                              float2* __restrict__ Z, int H, int W) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* c1 = (float2*)smem_raw;
    float2* twW = c1 + H * W;
    float2* twH = twW + W;
    float* d = (float*)(twH + H);
    const size_t off = (size_t)blockIdx.x * H * W;
    fill_twiddle(twW, W);
    fill_twiddle(twH, H);
    for (int i = threadIdx.x; i < H * W; i += blockDim.x) d[i] = x[off + i] - nz[off + i];
    __syncthreads();
    for (int i = threadIdx.x; i < H * W; i += blockDim.x) {
        const int kw = i % W, h = i / W;
        float re = 0.f, im = 0.f;
        int j = 0;
        for (int w = 0; w < W; ++w) {
            const float2 t = twW[j];
            const float v = d[h * W + w];
            re += v * t.x;
            im -= v * t.y;                 // exp(-i theta)
            j += kw; if (j >= W) j -= W;
        }
        c1[i] = make_float2(re, im);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < H * W; i += blockDim.x) {
        const int kw = i % W, kh = i / W;
        float re = 0.f, im = 0.f;
        int j = 0;
        for (int h = 0; h < H; ++h) {
            const float2 t = twH[j];
            const float2 v = c1[h * W + kw];
            re += v.x * t.x + v.y * t.y;   // v * (c - i s)
            im += v.y * t.x - v.x * t.y;
            j += kh; if (j >= H) j -= H;
        }
        Z[off + i] = make_float2(re, im);
    }
}

constexpr int kMaxT = 32;

// pass B: T-axis DFT, filter, inverse T-axis DFT.  One thread per (volume, h, w).
__global__ void __launch_bounds__(128) freeinit_t_filter_kernel(float2* __restrict__ Z, const float* __restrict__ lpf, int NV,
                                                                int T, int H, int W) {
    __shared__ float2 twT[kMaxT];
    fill_twiddle(twT, T);
    __syncthreads();
    const int HW = H * W;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NV * HW) return;
    const int hw = i % HW, v = i / HW;
    const int kw = hw % W, kh = hw / W;
    const int sh = (kh + H / 2) % H, sw = (kw + W / 2) % W;
    float2* col = Z + (size_t)v * T * HW + hw;
    float2 z[kMaxT], f[kMaxT];
    for (int t = 0; t < T; ++t) z[t] = col[(size_t)t * HW];
    for (int kt = 0; kt < T; ++kt) {
        float re = 0.f, im = 0.f;
        int j = 0;
        for (int t = 0; t < T; ++t) {
            const float2 tw = twT[j];
            re += z[t].x * tw.x + z[t].y * tw.y;
            im += z[t].y * tw.x - z[t].x * tw.y;
            j += kt; if (j >= T) j -= T;
        }
        const float g = lpf[((size_t)((kt + T / 2) % T) * H + sh) * W + sw];
        f[kt] = make_float2(re * g, im * g);
    }
    for (int t = 0; t < T; ++t) {
        float re = 0.f, im = 0.f;
        int j = 0;
        for (int kt = 0; kt < T; ++kt) {
            const float2 tw = twT[j];
            re += f[kt].x * tw.x - f[kt].y * tw.y;     // * exp(+i theta)
            im += f[kt].y * tw.x + f[kt].x * tw.y;
            j += t; if (j >= T) j -= T;
        }
        col[(size_t)t * HW] = make_float2(re, im);
    }
}

// pass C: inverse H then inverse W (real part), y = n + r / (T*H*W)
__global__ void __launch_bounds__(256) freeinit_inv_hw_kernel(const float2* __restrict__ Z, const float* __restrict__ nz,
                                                              float* __restrict__ y, int H, int W, float inv_n) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* c0 = (float2*)smem_raw;
    float2* c1 = c0 + H * W;
    float2* twW = c1 + H * W;
    float2* twH = twW + W;
    const size_t off = (size_t)blockIdx.x * H * W;
    fill_twiddle(twW, W);
    fill_twiddle(twH, H);
    for (int i = threadIdx.x; i < H * W; i += blockDim.x) c0[i] = Z[off + i];
    __syncthreads();
    for (int i = threadIdx.x; i < H * W; i += blockDim.x) {
        const int kw = i % W, h = i / W;
        float re = 0.f, im = 0.f;
        int j = 0;
        for (int kh = 0; kh < H; ++kh) {
            const float2 t = twH[j];
            const float2 v = c0[kh * W + kw];
            re += v.x * t.x - v.y * t.y;
            im += v.y * t.x + v.x * t.y;
            j += h; if (j >= H) j -= H;
        }
        c1[i] = make_float2(re, im);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < H * W; i += blockDim.x) {
        const int w = i % W, h = i / W;
        float re = 0.f;
        int j = 0;
        for (int kw = 0; kw < W; ++kw) {
            const float2 t = twW[j];
            const float2 v = c1[h * W + kw];
            re += v.x * t.x - v.y * t.y;
            j += w; if (j >= W) j -= W;
        }
        y[off + i] = nz[off + i] + re * inv_n;
    }
}

}  // namespace moca

using namespace moca;

extern "C" int32_t moca_freeinit_filter(moca_ctx* ctx, int32_t type, int32_t n, double d_s, double d_t, int32_t T, int32_t H,
                                        int32_t W, float* lpf, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && lpf != nullptr, MOCA_E_ARG, "moca_freeinit_filter: NULL argument");
    MOCA_REQUIRE(type >= 0 && type <= 3, MOCA_E_ARG, "moca_freeinit_filter: unknown filter type %d", type);
    MOCA_REQUIRE(T > 0 && H > 0 && W > 0, MOCA_E_SHAPE, "moca_freeinit_filter: bad shape");
    const int total = T * H * W;
    freeinit_filter_kernel<<<cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(type, n, d_s, d_t, T, H, W, lpf);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

extern "C" int64_t moca_freeinit_work_bytes(int32_t NV, int32_t T, int32_t H, int32_t W) {
    return (int64_t)NV * T * H * W * (int64_t)sizeof(float2);
}

extern "C" int32_t moca_freeinit_mix(moca_ctx* ctx, const float* x, const float* noise, const float* lpf, int32_t NV, int32_t T,
                                     int32_t H, int32_t W, float* y, void* work, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && x && noise && lpf && y && work, MOCA_E_ARG, "moca_freeinit_mix: NULL argument");
    MOCA_REQUIRE(NV > 0 && T > 0 && T <= kMaxT && H > 0 && W > 0, MOCA_E_SHAPE, "moca_freeinit_mix: bad shape NV%d T%d H%d W%d (T<=%d)",
                 NV, T, H, W, kMaxT);
    const size_t smemA = (size_t)H * W * (sizeof(float2) + sizeof(float)) + (size_t)(H + W) * sizeof(float2);
    const size_t smemC = (size_t)H * W * 2 * sizeof(float2) + (size_t)(H + W) * sizeof(float2);
    MOCA_REQUIRE(smemC <= 200 * 1024, MOCA_E_SHAPE, "moca_freeinit_mix: slice %dx%d does not fit shared memory", H, W);
    cudaStream_t st = (cudaStream_t)stream;
    // the attribute is per device and the call is cheap: set it on every launch (a process-wide "done" flag would leave the
    // second device of a process without the opt-in)
    MOCA_CHECK_CUDA(cudaFuncSetAttribute(freeinit_fwd_hw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    MOCA_CHECK_CUDA(cudaFuncSetAttribute(freeinit_inv_hw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    float2* Z = (float2*)work;
    freeinit_fwd_hw_kernel<<<NV * T, 256, smemA, st>>>(x, noise, Z, H, W);
    MOCA_LAUNCH_CHECK();
    freeinit_t_filter_kernel<<<cdiv((int64_t)NV * H * W, 128), 128, 0, st>>>(Z, lpf, NV, T, H, W);
    MOCA_LAUNCH_CHECK();
    freeinit_inv_hw_kernel<<<NV * T, 256, smemC, st>>>(Z, noise, y, H, W, 1.0f / (float)((int64_t)T * H * W));
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

// ---- allocation-free variant of moca_freeinit_mix -------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate (code under CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc).  The spectrum scratch that moca_freeinit_mix takes from the
// caller lives in a file-scope device array instead: kStaticSpectrum complex values = 2 MiB, held by every GPU that loads the
// library until the process exits.  All calls (all streams, all contexts of a device) share it and it keeps the last call's
// contents: the forward kernel rewrites every element the later kernels read, so nothing has to be cleared, but two calls that
// overlap on the GPU race on it — keep them on one stream, or order them with events.
constexpr int64_t kStaticSpectrum = 262144;          // float2 elements: NV * T * H * W <= 262144 (FIFO shift: 4 * 1 * 40 * 64 = 10240)
static __device__ __align__(256) float2 g_freeinit_spectrum[kStaticSpectrum];

extern "C" int64_t moca_freeinit_mix_static_max_elems() { return kStaticSpectrum; }

extern "C" int32_t moca_freeinit_mix_static(moca_ctx* ctx, const float* x, const float* noise, const float* lpf, int32_t NV, int32_t T,
                                            int32_t H, int32_t W, float* y, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && x && noise && lpf && y, MOCA_E_ARG, "moca_freeinit_mix_static: NULL argument");
    MOCA_REQUIRE(NV > 0 && T > 0 && T <= kMaxT && H > 0 && W > 0, MOCA_E_SHAPE, "moca_freeinit_mix_static: bad shape NV%d T%d H%d W%d (T<=%d)",
                 NV, T, H, W, kMaxT);
    const int64_t elems = (int64_t)NV * T * H * W;
    MOCA_REQUIRE(elems <= kStaticSpectrum, MOCA_E_SHAPE,
                 "moca_freeinit_mix_static: NV*T*H*W = %lld exceeds the built-in scratch of %lld complex values (2 MiB); use moca_freeinit_mix "
                 "with caller-provided work memory", (long long)elems, (long long)kStaticSpectrum);
    const size_t smemA = (size_t)H * W * (sizeof(float2) + sizeof(float)) + (size_t)(H + W) * sizeof(float2);
    const size_t smemC = (size_t)H * W * 2 * sizeof(float2) + (size_t)(H + W) * sizeof(float2);
    MOCA_REQUIRE(smemC <= 200 * 1024, MOCA_E_SHAPE, "moca_freeinit_mix_static: slice %dx%d does not fit shared memory", H, W);
    // device address of the array, resolved once per device with that device current (the symbol's name is a host-side stand-in
    // here: handing it to a kernel would fault or touch the wrong memory)
    static float2* dev_ptr[64] = {nullptr};
    int dev = -1;
    MOCA_CHECK_CUDA(cudaGetDevice(&dev));
    MOCA_REQUIRE(dev >= 0 && dev < 64, MOCA_E_ARG, "moca_freeinit_mix_static: device index %d outside 0..63", dev);
    if (!dev_ptr[dev]) {
        float2* p = nullptr;
        MOCA_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_freeinit_spectrum));
        MOCA_REQUIRE(p != nullptr, MOCA_E_CUDA, "moca_freeinit_mix_static: cudaGetSymbolAddress returned NULL");
        dev_ptr[dev] = p;
    }
    float2* Z = dev_ptr[dev];
    cudaStream_t st = (cudaStream_t)stream;
    MOCA_CHECK_CUDA(cudaFuncSetAttribute(freeinit_fwd_hw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    MOCA_CHECK_CUDA(cudaFuncSetAttribute(freeinit_inv_hw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    freeinit_fwd_hw_kernel<<<NV * T, 256, smemA, st>>>(x, noise, Z, H, W);
    MOCA_LAUNCH_CHECK();
    freeinit_t_filter_kernel<<<cdiv((int64_t)NV * H * W, 128), 128, 0, st>>>(Z, lpf, NV, T, H, W);
    MOCA_LAUNCH_CHECK();
    freeinit_inv_hw_kernel<<<NV * T, 256, smemC, st>>>(Z, noise, y, H, W, 1.0f / (float)((int64_t)T * H * W));
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}
