// HBM-bound latent kernels of the sampler / FIFO queue.
//
// This translation unit is compiled with -fmad=false: the reference runs these updates as separate torch
// elementwise ops (no FMA contraction), and the parity bar for this arithmetic is bit equality with
// oracle/step_ref.py on identical inputs.  Division and sqrt are IEEE round-to-nearest (nvcc defaults).
//
//   cfg_ddim_step_kernel   DDIMSampler.unet CFG combine (lvdm/models/samplers/ddim.py:372) +
//                          DDIMSampler.ddim_step (ddim.py:405-430, 557-609)      [SURVEY Appendix B]
//   p_sample_kernel        DDIMSampler.p_sample_ddim (ddim.py:339-357)
//   renoise_kernel         prepare_latents (scripts/evaluation/funcs.py:55-80)
//   copy_frames_kernel     window slicing / look-ahead write-back / shift (funcs.py:307-316, 351-352, 89-91)
#include "common.cuh"

namespace moca {

// ------------------------------------------------------------------------------------------------------------
// mask pre-pass: sum of every mask frame over (B,H,W), accumulated in double in a fixed order (ddim.py:587 `.sum() != 0`,
// :826 coverage test).  One block per mask frame.
__global__ void mask_sum_kernel(const float* __restrict__ mask, int B, int Fm, int HW, double* __restrict__ sums) {
    const int j = blockIdx.x;
    double acc = 0.0;
    for (int b = 0; b < B; ++b) {
        const float* p = mask + ((size_t)b * Fm + j) * HW;
        for (int i = threadIdx.x; i < HW; i += blockDim.x) acc += (double)p[i];
    }
    __shared__ double sh[32];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.0;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) sums[j] = v;
    }
}

struct StepParams {
    float A[MOCA_MAX_FRAMES], Ap[MOCA_MAX_FRAMES], Sg[MOCA_MAX_FRAMES], R[MOCA_MAX_FRAMES];
    int ts[MOCA_MAX_FRAMES];
    int mask_idx[MOCA_MAX_FRAMES];
    float cfg;
    int use_cfg;
    float omg, g;             // (float)(1 - gamma), (float)gamma
    float enh_lo, enh_hi;
    int t_thresh;
    double cover_limit;       // cover_max * B*H*W
    int B, C, F, H, W, Fm;
};

constexpr int kStepChunk = 4;

// One thread per (b,c,h,w); frames are visited in order because the momentum correction is a scan over frames
// (SURVEY Q4).  Loads of a chunk of frames are issued before the dependent arithmetic of that chunk.
__global__ void __launch_bounds__(128) cfg_ddim_step_kernel(
    const float* __restrict__ x, const float* __restrict__ e_c, const float* __restrict__ e_u,
    const float* __restrict__ noise, const float* __restrict__ mask, const float* __restrict__ cond,
    const double* __restrict__ mask_sums, float* __restrict__ x_prev, float* __restrict__ pred_x0,
    const __grid_constant__ StepParams P) {
    __shared__ float s_sqrtA[MOCA_MAX_FRAMES], s_sqrtAp[MOCA_MAX_FRAMES], s_cdir[MOCA_MAX_FRAMES], s_corr[MOCA_MAX_FRAMES];
    __shared__ float s_k[MOCA_MAX_FRAMES];          // blend gain of the frame, 0 = no blend
    if (threadIdx.x < P.F) {
        const int i = threadIdx.x;
        s_sqrtA[i] = sqrtf(P.A[i]);                                   // ddim.py:415 a_t.sqrt()
        s_sqrtAp[i] = sqrtf(P.Ap[i]);                                 // :562
        s_cdir[i] = sqrtf((1.0f - P.Ap[i]) - P.Sg[i] * P.Sg[i]);      // :418
        s_corr[i] = 2.0f * (1.0f - (float)P.ts[i] / 1000.0f);         // :428
        float k = 0.0f;
        const int j = P.mask_idx[i];
        if (mask != nullptr && j >= 0 && j < P.Fm) {
            k = P.ts[i] <= P.t_thresh ? P.enh_lo : P.enh_hi;          // :584 / :592,:845
            const double s = mask_sums[j];
            if (!(s != 0.0 && (P.cover_limit <= 0.0 || s <= P.cover_limit))) k = 0.0f;     // :587 / :826 (limit <= 0: no coverage test)
        }
        s_k[i] = k;
    }
    __syncthreads();
    const int HW = P.H * P.W;
    const int total = P.B * P.C * HW;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= total) return;
    const int hw = tid % HW;
    const int bc = tid / HW;
    const int b = bc / P.C;
    const size_t base = (size_t)bc * P.F * HW + hw;
    const float ci = (mask != nullptr) ? cond[(size_t)bc * HW + hw] : 0.0f;
    const float* mrow = (mask != nullptr) ? mask + (size_t)b * P.Fm * HW + hw : nullptr;

    float m_prev = 0.0f, p_prev = 0.0f;
    for (int f0 = 0; f0 < P.F; f0 += kStepChunk) {
        float vx[kStepChunk], vc[kStepChunk], vu[kStepChunk], vn[kStepChunk], vm[kStepChunk];
#pragma unroll
        for (int u = 0; u < kStepChunk; ++u) {
            const int i = f0 + u;
            if (i < P.F) {
                const size_t o = base + (size_t)i * HW;
                vx[u] = x[o];
                vc[u] = e_c[o];
                vu[u] = P.use_cfg ? e_u[o] : 0.0f;
                vn[u] = noise[o];
                vm[u] = (s_k[i] != 0.0f) ? mrow[(size_t)P.mask_idx[i] * HW] : 0.0f;
            }
        }
#pragma unroll
        for (int u = 0; u < kStepChunk; ++u) {
            const int i = f0 + u;
            if (i < P.F) {
                float e = vc[u];
                if (P.use_cfg) e = vu[u] + P.cfg * (vc[u] - vu[u]);                 // ddim.py:372
                float p = (vx[u] - P.R[i] * e) / s_sqrtA[i];                        // :415
                const float d = s_cdir[i] * e;                                      // :418
                if (i > 0) {
                    const float g = (p - p_prev) + 1.5f * d;                        // :422-423
                    const float m = 0.9f * m_prev + 0.1f * g;                       // :424-427
                    p = p + s_corr[i] * m;                                          // :428-430, 557
                    m_prev = m;
                }
                p_prev = p;                                                         // :559
                const float n = P.Sg[i] * vn[u];                                    // :561
                const size_t o = base + (size_t)i * HW;
                x_prev[o] = (s_sqrtAp[i] * p + d) + n;                              // :562
                const float k = s_k[i];
                if (k != 0.0f && vm[u] > 0.5f) p = ci * k;                          // :588-590 / :851
                pred_x0[o] = P.omg * p + P.g * n;                                   // :609
            }
        }
    }
}

__global__ void p_sample_kernel(const float* __restrict__ x, const float* __restrict__ e_c, const float* __restrict__ e_u,
                                const float* __restrict__ noise, float cfg, int use_cfg, float sqrtA, float R, float cdir,
                                float Sg, float xcoef, float scale, int use_scale, int64_t n, float* __restrict__ x_prev,
                                float* __restrict__ pred_x0) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float e = e_c[i];
        if (use_cfg) {
            const float u = e_u[i];
            e = u + cfg * (e - u);                       // ddim.py:304
        }
        float p = (x[i] - R * e) / sqrtA;                // :339
        const float d = cdir * e;                        // :343
        const float nz = Sg * noise[i];                  // :345
        if (use_scale) p = p / scale;                    // :353
        x_prev[i] = (xcoef * p + d) + nz;                // :355 / :357
        pred_x0[i] = p;
    }
}

// ------------------------------------------------------------------------------------------------------------
// SAM branch of ddim_step (no DAVIS masks; lvdm/models/samplers/ddim.py:591-606 -> _apply_segmentation :739-903)
//
//   latent_to_u8_kernel   :745-761  x0 prediction of one frame -> the uint8 RGB image shown to Grounding-DINO / SAM-2
//   mask_iou_kernel       :905-943  intersection / union counts of mask pairs (the `iou < 0.5` fallback gate, :809-813)
//   sam_blend_kernel      :818-901 + :609  per frame: masks in order, a mask covering > cover_max of the frame RESETS the
//                         result to the unblended prediction, otherwise where(mask > 0.5, cond * gain, cur); then the
//                         gamma noise blend.  (As shipped the reference's torch.where returns C identical copies of the
//                         frame along dim 2; one copy is produced here.)
// One block; the frame is 3*H*W <= a few 10k values.
__global__ void __launch_bounds__(1024) latent_to_u8_kernel(const float* __restrict__ p, int C, int F, int HW, int frame,
                                                            uint8_t* __restrict__ img) {
    __shared__ float red[32];
    __shared__ float s_max;
    const int nch = C == 1 ? 1 : 3;
    float mx = -INFINITY;
    for (int i = threadIdx.x; i < nch * HW; i += blockDim.x) {
        const int c = i / HW, hw = i - c * HW;
        mx = fmaxf(mx, p[((size_t)c * F + frame) * HW + hw]);
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = red[0];
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m = fmaxf(m, red[w]);
        s_max = m;
    }
    __syncthreads();
    const bool scale = floorf(s_max) <= 1.0f;                               // :757 `np.floor(frame.max()) <= 1.0`
    for (int i = threadIdx.x; i < 3 * HW; i += blockDim.x) {
        const int hw = i / 3, c = i - hw * 3;
        float v = p[((size_t)(C == 1 ? 0 : c) * F + frame) * HW + hw];
        if (scale) v = v * 255.0f;                                          // :758
        // ndarray.astype(np.uint8) of a float: truncate toward zero, keep the low 8 bits (x86-64 numpy; pinned by the golden)
        const long long t = (long long)truncf(v);
        img[i] = (uint8_t)(t & 255);
    }
}

// pair j: masks a[j], b[j] (fp32 [HW]) binarised at 0.5 -> counts[2j] = |a & b|, counts[2j+1] = |a | b|
__global__ void __launch_bounds__(256) mask_iou_kernel(const float* __restrict__ a, const float* __restrict__ b, int HW,
                                                       int* __restrict__ counts) {
    const int j = blockIdx.x;
    int inter = 0, uni = 0;
    for (int i = threadIdx.x; i < HW; i += blockDim.x) {
        const bool x = a[(size_t)j * HW + i] > 0.5f, y = b[(size_t)j * HW + i] > 0.5f;
        inter += (x && y) ? 1 : 0;
        uni += (x || y) ? 1 : 0;
    }
    __shared__ int sh[2][8];
    for (int o = 16; o > 0; o >>= 1) { inter += __shfl_xor_sync(0xffffffffu, inter, o); uni += __shfl_xor_sync(0xffffffffu, uni, o); }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = inter; sh[1][threadIdx.x >> 5] = uni; }
    __syncthreads();
    if (threadIdx.x == 0) {
        int ti = 0, tu = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { ti += sh[0][w]; tu += sh[1][w]; }
        counts[2 * j] = ti; counts[2 * j + 1] = tu;
    }
}

struct SamParams {
    float Sg[MOCA_MAX_FRAMES];
    int nmask[MOCA_MAX_FRAMES];       // masks to apply for the frame (0: none, also for t > t_thresh)
    float omg, g, gain;
    double cover_limit;               // cover_max * H * W
    int B, C, F, HW, Nmax;
};

__global__ void __launch_bounds__(128) sam_blend_kernel(const float* __restrict__ p_corr, const float* __restrict__ noise,
                                                        const float* __restrict__ cond, const float* __restrict__ masks,
                                                        const double* __restrict__ mask_sums, float* __restrict__ pred_x0,
                                                        const __grid_constant__ SamParams P) {
    const int total = P.B * P.C * P.HW;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= total) return;
    const int hw = tid % P.HW;
    const int bc = tid / P.HW;
    const size_t base = (size_t)bc * P.F * P.HW + hw;
    const float cg = cond[(size_t)bc * P.HW + hw] * P.gain;                 // :851 `cond_image * enhancement_factor`
    for (int i = 0; i < P.F; ++i) {
        const size_t o = base + (size_t)i * P.HW;
        const float p = p_corr[o];
        float cur = p;
        for (int m = 0; m < P.nmask[i]; ++m) {
            const int j = i * P.Nmax + m;
            if (mask_sums[j] > P.cover_limit) { cur = p; continue; }        // :826-828
            if (masks[(size_t)j * P.HW + hw] > 0.5f) cur = cg;             // :897-901
        }
        const float n = P.Sg[i] * noise[o];                                 // :561
        pred_x0[o] = P.omg * cur + P.g * n;                                 // :609
    }
}

struct RenoiseParams {
    float sa[96], sb[96];
    int src[96];
};

__global__ void renoise_kernel(const float* __restrict__ z, const float* __restrict__ noise, float* __restrict__ queue,
                               int BC, int Fz, int Q, int HW, const __grid_constant__ RenoiseParams P) {
    const int64_t total = (int64_t)BC * Q * HW;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int hw = (int)(i % HW);
        const int q = (int)((i / HW) % Q);
        const int bc = (int)(i / ((int64_t)HW * Q));
        const float zz = z[((size_t)bc * Fz + P.src[q]) * HW + hw];
        queue[i] = P.sa[q] * zz + P.sb[q] * noise[i];     // funcs.py:61 / :75
    }
}

// dst[bc, d0+f, :] = src[bc, s0+f, :] for f < nf.  V = float4 when HW % 4 == 0 and pointers are 16B aligned.
template <typename V>
__global__ void copy_frames_kernel(const V* __restrict__ src, int srcF, int s0, V* __restrict__ dst, int dstF, int d0,
                                   int nf, int BC, int HWv) {
    const int64_t total = (int64_t)BC * nf * HWv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int hw = (int)(i % HWv);
        const int f = (int)((i / HWv) % nf);
        const int bc = (int)(i / ((int64_t)HWv * nf));
        dst[((size_t)bc * dstF + d0 + f) * HWv + hw] = src[((size_t)bc * srcF + s0 + f) * HWv + hw];
    }
}

static int copy_frames(const float* src, int srcF, int s0, float* dst, int dstF, int d0, int nf, int BC, int HW,
                       cudaStream_t st) {
    if (nf <= 0 || BC <= 0) return MOCA_OK;
    const bool vec = (HW % 4 == 0) && (((uintptr_t)src | (uintptr_t)dst) % 16 == 0);
    if (vec) {
        const int64_t total = (int64_t)BC * nf * (HW / 4);
        const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
        copy_frames_kernel<float4><<<blocks, 256, 0, st>>>((const float4*)src, srcF, s0, (float4*)dst, dstF, d0, nf, BC,
                                                           HW / 4);
    } else {
        const int64_t total = (int64_t)BC * nf * HW;
        const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
        copy_frames_kernel<float><<<blocks, 256, 0, st>>>(src, srcF, s0, dst, dstF, d0, nf, BC, HW);
    }
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

}  // namespace moca

using namespace moca;

extern "C" int32_t moca_cfg_ddim_step(moca_ctx* ctx, const float* x, const float* e_c, const float* e_u,
                                      const float* noise, float cfg_scale, const float* coef, const int32_t* ts,
                                      const float* mask, int32_t Fm, const int32_t* mask_idx, const float* cond_img,
                                      double gamma, float enh_lo, float enh_hi, int32_t t_thresh, float cover_max,
                                      int32_t B, int32_t C, int32_t F, int32_t H, int32_t W, float* x_prev,
                                      float* pred_x0, void* stream) {
    MOCA_REQUIRE(ctx != nullptr, MOCA_E_ARG, "moca_cfg_ddim_step: ctx is NULL");
    MOCA_REQUIRE(x && e_c && noise && coef && ts && x_prev && pred_x0, MOCA_E_ARG, "moca_cfg_ddim_step: NULL tensor");
    MOCA_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, MOCA_E_SHAPE, "moca_cfg_ddim_step: bad shape B%d C%d H%d W%d", B, C, H, W);
    MOCA_REQUIRE(F > 0 && F <= MOCA_MAX_FRAMES, MOCA_E_SHAPE, "moca_cfg_ddim_step: F=%d outside 1..%d", F, MOCA_MAX_FRAMES);
    MOCA_REQUIRE(mask == nullptr || (cond_img != nullptr && mask_idx != nullptr && Fm > 0 && Fm <= MOCA_MAX_FRAMES),
                 MOCA_E_ARG, "moca_cfg_ddim_step: mask given without cond_img/mask_idx or Fm=%d outside 1..%d", Fm,
                 MOCA_MAX_FRAMES);
    MOCA_REQUIRE((int64_t)B * C * H * W < (1ll << 31), MOCA_E_SHAPE, "moca_cfg_ddim_step: tensor too large");
    cudaStream_t st = (cudaStream_t)stream;
    StepParams P;
    memset(&P, 0, sizeof(P));
    for (int i = 0; i < F; ++i) {
        P.A[i] = coef[i];
        P.Ap[i] = coef[F + i];
        P.Sg[i] = coef[2 * F + i];
        P.R[i] = coef[3 * F + i];
        P.ts[i] = ts[i];
        P.mask_idx[i] = (mask && mask_idx) ? mask_idx[i] : -1;
    }
    P.cfg = cfg_scale;
    P.use_cfg = (e_u != nullptr && cfg_scale != 1.0f) ? 1 : 0;      // ddim.py:366
    P.omg = (float)(1.0 - gamma);
    P.g = (float)gamma;
    P.enh_lo = enh_lo;
    P.enh_hi = enh_hi;
    P.t_thresh = t_thresh;
    P.cover_limit = (double)cover_max * (double)B * H * W;
    P.B = B; P.C = C; P.F = F; P.H = H; P.W = W; P.Fm = mask ? Fm : 0;
    if (mask) {
        mask_sum_kernel<<<Fm, 256, 0, st>>>(mask, B, Fm, H * W, ctx->c.d_mask_sums);
        MOCA_LAUNCH_CHECK();
    }
    const int total = B * C * H * W;
    cfg_ddim_step_kernel<<<cdiv(total, 128), 128, 0, st>>>(x, e_c, e_u, noise, mask, cond_img, ctx->c.d_mask_sums, x_prev,
                                                           pred_x0, P);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

extern "C" int32_t moca_p_sample_ddim(moca_ctx* ctx, const float* x, const float* e_c, const float* e_u,
                                      const float* noise, float cfg_scale, float A, float A_prev, float sigma,
                                      float sqrt_one_minus_A, int32_t use_scale, float scale, float scale_prev, int64_t n,
                                      float* x_prev, float* pred_x0, void* stream) {
    MOCA_REQUIRE(ctx != nullptr, MOCA_E_ARG, "moca_p_sample_ddim: ctx is NULL");
    MOCA_REQUIRE(x && e_c && noise && x_prev && pred_x0, MOCA_E_ARG, "moca_p_sample_ddim: NULL tensor");
    MOCA_REQUIRE(n > 0, MOCA_E_SHAPE, "moca_p_sample_ddim: n=%lld", (long long)n);
    // scalar coefficients exactly as the reference forms them from torch.full tensors (ddim.py:329-345, 353-355);
    // volatile keeps the host compiler from contracting or reassociating
    volatile float sqrtA = sqrtf(A);
    volatile float sg2 = sigma * sigma;
    volatile float om = 1.0f - A_prev;
    volatile float cdir = sqrtf(om - sg2);
    volatile float sAp = sqrtf(A_prev);
    volatile float xcoef = use_scale ? sAp * scale_prev : sAp;
    const int use_cfg = (e_u != nullptr && cfg_scale != 1.0f) ? 1 : 0;
    const int blocks = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
    p_sample_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, e_c, e_u, noise, cfg_scale, use_cfg, sqrtA, sqrt_one_minus_A,
                                                              cdir, sigma, xcoef, scale, use_scale, n, x_prev, pred_x0);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

extern "C" int32_t moca_renoise(moca_ctx* ctx, const float* z, const float* noise, const int32_t* src_frame,
                                const float* alpha, int32_t B, int32_t C, int32_t Fz, int32_t Q, int32_t H, int32_t W,
                                float* queue, void* stream) {
    MOCA_REQUIRE(ctx != nullptr, MOCA_E_ARG, "moca_renoise: ctx is NULL");
    MOCA_REQUIRE(z && noise && src_frame && alpha && queue, MOCA_E_ARG, "moca_renoise: NULL argument");
    MOCA_REQUIRE(Q > 0 && Q <= 96, MOCA_E_SHAPE, "moca_renoise: Q=%d outside 1..96", Q);
    MOCA_REQUIRE(B > 0 && C > 0 && Fz > 0 && H > 0 && W > 0, MOCA_E_SHAPE, "moca_renoise: bad shape");
    RenoiseParams P;
    for (int q = 0; q < Q; ++q) {
        MOCA_REQUIRE(src_frame[q] >= 0 && src_frame[q] < Fz, MOCA_E_ARG, "moca_renoise: src_frame[%d]=%d outside 0..%d", q,
                     src_frame[q], Fz - 1);
        volatile float a = alpha[q];
        volatile float om = 1.0f - a;
        P.sa[q] = sqrtf(a);          // alpha**0.5: torch.pow special-cases 0.5 to sqrt
        P.sb[q] = sqrtf(om);
        P.src[q] = src_frame[q];
    }
    const int64_t total = (int64_t)B * C * Q * H * W;
    const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    renoise_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(z, noise, queue, B * C, Fz, Q, H * W, P);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

extern "C" int32_t moca_queue_shift(moca_ctx* ctx, const float* queue, const float* tail, int32_t B, int32_t C, int32_t Q,
                                    int32_t H, int32_t W, float* out, void* stream) {
    MOCA_REQUIRE(ctx != nullptr, MOCA_E_ARG, "moca_queue_shift: ctx is NULL");
    MOCA_REQUIRE(queue && tail && out && queue != out, MOCA_E_ARG, "moca_queue_shift: NULL or aliased tensors");
    MOCA_REQUIRE(B > 0 && C > 0 && Q > 0 && H > 0 && W > 0, MOCA_E_SHAPE, "moca_queue_shift: bad shape");
    int rc = copy_frames(queue, Q, 1, out, Q, 0, Q - 1, B * C, H * W, (cudaStream_t)stream);
    if (rc) return rc;
    return copy_frames(tail, 1, 0, out, Q, Q - 1, 1, B * C, H * W, (cudaStream_t)stream);
}

extern "C" int32_t moca_window_gather(moca_ctx* ctx, const float* queue, int32_t B, int32_t C, int32_t Q, int32_t H,
                                      int32_t W, int32_t start, int32_t Fw, float* win, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && queue && win, MOCA_E_ARG, "moca_window_gather: NULL argument");
    MOCA_REQUIRE(start >= 0 && Fw > 0 && start + Fw <= Q, MOCA_E_SHAPE, "moca_window_gather: window [%d,%d) outside queue of %d",
                 start, start + Fw, Q);
    return copy_frames(queue, Q, start, win, Fw, 0, Fw, B * C, H * W, (cudaStream_t)stream);
}

extern "C" int32_t moca_window_scatter(moca_ctx* ctx, const float* win, int32_t Fwin, int32_t win_off, int32_t B, int32_t C,
                                       int32_t Q, int32_t H, int32_t W, int32_t start, int32_t Fw, float* queue,
                                       void* stream) {
    MOCA_REQUIRE(ctx != nullptr && queue && win, MOCA_E_ARG, "moca_window_scatter: NULL argument");
    MOCA_REQUIRE(start >= 0 && Fw > 0 && start + Fw <= Q && win_off >= 0 && win_off + Fw <= Fwin, MOCA_E_SHAPE,
                 "moca_window_scatter: bad window start=%d Fw=%d Q=%d win_off=%d Fwin=%d", start, Fw, Q, win_off, Fwin);
    return copy_frames(win, Fwin, win_off, queue, Q, start, Fw, B * C, H * W, (cudaStream_t)stream);
}

extern "C" int32_t moca_halfwin_pack(moca_ctx* ctx, const float* xprev, int32_t Fwin, int32_t half, int32_t B, int32_t C,
                                     int32_t H, int32_t W, float* packed, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && xprev && packed, MOCA_E_ARG, "moca_halfwin_pack: NULL argument");
    MOCA_REQUIRE(half > 0 && half <= Fwin, MOCA_E_SHAPE, "moca_halfwin_pack: half=%d Fwin=%d", half, Fwin);
    return copy_frames(xprev, Fwin, Fwin - half, packed, half, 0, half, B * C, H * W, (cudaStream_t)stream);
}

extern "C" int32_t moca_halfwin_unpack(moca_ctx* ctx, const float* packed, int32_t n_windows, const int32_t* window_start,
                                       int32_t half, int32_t B, int32_t C, int32_t Q, int32_t H, int32_t W, float* queue,
                                       void* stream) {
    MOCA_REQUIRE(ctx != nullptr && packed && queue && window_start, MOCA_E_ARG, "moca_halfwin_unpack: NULL argument");
    const size_t per = (size_t)B * C * half * H * W;
    for (int i = 0; i < n_windows; ++i) {
        const int dst0 = window_start[i] + half;        // later half of the window (funcs.py:351-352)
        MOCA_REQUIRE(window_start[i] >= 0 && dst0 + half <= Q, MOCA_E_SHAPE, "moca_halfwin_unpack: window %d start %d outside queue %d",
                     i, window_start[i], Q);
        int rc = copy_frames(packed + per * i, half, 0, queue, Q, dst0, half, B * C, H * W, (cudaStream_t)stream);
        if (rc) return rc;
    }
    return MOCA_OK;
}

// ---- SAM branch glue (ddim.py:739-761, 787-822, 905-943) ----------------------------------------------
extern "C" int32_t moca_latent_to_u8(moca_ctx* ctx, const float* pred_x0, int32_t C, int32_t F, int32_t H, int32_t W,
                                     int32_t frame, uint8_t* image, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && pred_x0 && image, MOCA_E_ARG, "moca_latent_to_u8: NULL argument");
    MOCA_REQUIRE((C == 1 || C >= 3) && F > 0 && H > 0 && W > 0 && frame >= 0 && frame < F, MOCA_E_SHAPE,
                 "moca_latent_to_u8: C=%d (1 or >= 3) F=%d frame=%d", C, F, frame);
    latent_to_u8_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(pred_x0, C, F, H * W, frame, image);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

extern "C" int32_t moca_mask_iou(moca_ctx* ctx, const float* masks_a, int32_t n_a, const float* masks_b, int32_t n_b,
                                 int32_t H, int32_t W, float* iou_mean, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && masks_a && masks_b && iou_mean, MOCA_E_ARG, "moca_mask_iou: NULL argument");
    const int n = n_a < n_b ? n_a : n_b;                                   // zip(masks1, masks2), ddim.py:925
    MOCA_REQUIRE(n >= 0 && n <= MOCA_MAX_MASKS && H > 0 && W > 0, MOCA_E_SHAPE, "moca_mask_iou: %d pairs (max %d)", n, MOCA_MAX_MASKS);
    if (n == 0) { *iou_mean = nanf(""); return MOCA_OK; }                  // torch.tensor([]).mean()
    cudaStream_t st = (cudaStream_t)stream;
    int* d_counts = reinterpret_cast<int*>(ctx->c.d_mask_sums);
    mask_iou_kernel<<<n, 256, 0, st>>>(masks_a, masks_b, H * W, d_counts);
    MOCA_LAUNCH_CHECK();
    int h[2 * MOCA_MAX_MASKS];
    MOCA_CHECK_CUDA(cudaMemcpyAsync(h, d_counts, sizeof(int) * 2 * n, cudaMemcpyDeviceToHost, st));
    MOCA_CHECK_CUDA(cudaStreamSynchronize(st));       // the caller branches on the value (ddim.py:811)
    volatile float acc = 0.0f;
    for (int j = 0; j < n; ++j) {
        volatile float inter = (float)h[2 * j], uni = (float)h[2 * j + 1];
        volatile float iou = h[2 * j + 1] == 0 ? 1.0f : inter / uni;        // :931-938
        acc = acc + iou;
    }
    *iou_mean = acc / (float)n;                                            // :943
    return MOCA_OK;
}

extern "C" int32_t moca_sam_blend(moca_ctx* ctx, const float* p_corr, const float* noise, const float* sigma,
                                  const float* cond_img, const float* masks, const int32_t* n_masks, int32_t Nmax,
                                  double gamma, float gain, float cover_max, int32_t B, int32_t C, int32_t F, int32_t H,
                                  int32_t W, float* pred_x0, void* stream) {
    MOCA_REQUIRE(ctx != nullptr && p_corr && noise && sigma && n_masks && pred_x0, MOCA_E_ARG, "moca_sam_blend: NULL argument");
    MOCA_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && F > 0 && F <= MOCA_MAX_FRAMES, MOCA_E_SHAPE, "moca_sam_blend: bad shape");
    MOCA_REQUIRE(Nmax >= 0 && (int64_t)F * Nmax <= MOCA_MAX_MASKS, MOCA_E_SHAPE, "moca_sam_blend: F*Nmax=%lld > %d", (long long)F * Nmax,
                 MOCA_MAX_MASKS);
    cudaStream_t st = (cudaStream_t)stream;
    SamParams P;
    memset(&P, 0, sizeof(P));
    bool any = false;
    for (int i = 0; i < F; ++i) {
        MOCA_REQUIRE(n_masks[i] >= 0 && n_masks[i] <= Nmax, MOCA_E_ARG, "moca_sam_blend: n_masks[%d]=%d outside 0..%d", i, n_masks[i], Nmax);
        P.Sg[i] = sigma[i];
        P.nmask[i] = n_masks[i];
        any = any || n_masks[i] > 0;
    }
    MOCA_REQUIRE(!any || (masks && cond_img), MOCA_E_ARG, "moca_sam_blend: masks listed without mask / cond_img tensors");
    P.omg = (float)(1.0 - gamma); P.g = (float)gamma; P.gain = gain;
    P.cover_limit = (double)cover_max * (double)H * W;
    P.B = B; P.C = C; P.F = F; P.HW = H * W; P.Nmax = Nmax;
    if (any) {
        mask_sum_kernel<<<F * Nmax, 256, 0, st>>>(masks, 1, F * Nmax, H * W, ctx->c.d_mask_sums);
        MOCA_LAUNCH_CHECK();
    }
    const int total = B * C * H * W;
    sam_blend_kernel<<<cdiv(total, 128), 128, 0, st>>>(p_corr, noise, any ? cond_img : p_corr, masks, ctx->c.d_mask_sums, pred_x0, P);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}
