// HBM-bound glue kernels of the UNet: weight packing, GroupNorm statistics and
// apply(+SiLU) with layout scatter into the zero-bordered conv buffers, LayerNorm, embeddings, layout converts.
#pragma once
#include "common.cuh"

namespace moca {

enum ApplyDst { DST_LINEAR = 0, DST_SPATIAL_PAD = 1, DST_SPATIAL_PAD_UP2 = 2, DST_TIME_PAD = 3, DST_UP2 = 4 /* nearest x2, no border */ };

// fp32 [N, K, taps] (taps innermost, the torch conv weight layout [O, I, kh, kw] / [O, I, kt, 1, 1]; taps = 1 for Linear)
// -> bf16 dst[(row0 + perm(n)), tap*Kp + k], zero for k >= K.  geglu_bn > 0 permutes rows so that each geglu_bn-row
// tile holds geglu_bn/2 value rows followed by their geglu_bn/2 gate rows (GEGLU: attention.py:376-383).
int pack_weight(const float* src, int N, int K, int taps, int Kp, op_t* dst, int64_t dst_pitch, int row0, int geglu_bn,
                cudaStream_t st);

// Sum / sum of squares per (stat group, channel group) in double.  Tokens are channels-last fp32 [rows, C1] (+ optional
// second source [rows, C2] = virtual channel concat, openaimodel3d.py:571).  rows_per_stat = H*W (per-frame GroupNorm)
// or T*H*W (GroupNorm on 5-D input: TemporalConvBlock / TemporalTransformer).  acc: [nstat, 32, 2] doubles, zeroed.
// The last block of each statistics group to finish turns the sums into (mean, rstd) pairs: table [nstat, 32] float2.
// counter: [nstat] unsigned, zero before the first launch, reset by the kernel itself.
int gn_stats(const float* x1, int C1, const float* x2, int C2, int64_t rows, int rows_per_stat, int groups, double* acc,
             float eps, float2* table, unsigned* counter, cudaStream_t st);

// Statistics a GEMM epilogue left behind for the tensor it wrote (gemm_tc.cuh): {sum, sumsq} doubles per (entry, channel),
// cs [n, C, 2]; entry e covers output rows [e * rows, (e + 1) * rows) — a frame (rows = H*W) or a sample (rows = T*H*W),
// whichever the GroupNorm that consumes the tensor normalises over (a per-frame source can also feed a per-sample norm).
struct StatSrc {
    const double* cs = nullptr;
    int C = 0;
    int rows = 0;
};

// (scale, shift) = (rstd * gamma, beta - mean * rstd * gamma) per (statistics group, channel) of GroupNorm(32) over the virtual
// channel concat of s1 and s2 (s2.C == 0: single source), rows_per_stat rows per group (a multiple of the sources' rows per entry):
// ab [nstat, C1 + C2] float2, so that gn_apply is one FMA per element with no group arithmetic.
int gn_finalize(const StatSrc& s1, const StatSrc& s2, int nstat, int rows_per_stat, float eps, const float* gamma, const float* beta,
                float2* ab, cudaStream_t st);

struct ApplyArgs {
    const float* x1 = nullptr; int C1 = 0;
    const float* x2 = nullptr; int C2 = 0;
    int64_t rows = 0;
    int rows_per_stat = 0;          // 0: no normalisation (plain convert)
    int groups = 32;
    const float2* table = nullptr;  // (mean, rstd) per (statistics group, channel group), written by gn_stats
    const float* gamma = nullptr; const float* beta = nullptr;
    const float2* ab = nullptr;     // optional [nstat, C1+C2] precomputed (scale, shift); replaces table/gamma/beta
    int in16 = 0;                   // x1 points to 16-bit (op_t) rows instead of fp32 (affine-table path, single source, no raw copy)
    float eps = 1e-5f;
    int silu = 0;
    int dst_mode = DST_LINEAR;
    int NF = 0, H = 0, W = 0;       // spatial geometry of the SOURCE tokens (DST_SPATIAL_*)
    int B = 0, T = 0, HW = 0;       // DST_TIME_PAD
    op_t* dst = nullptr;   // [.., C1+C2]
    op_t* raw = nullptr;   // optional un-normalised bf16 copy [rows, C1+C2] (ResBlock.skip_connection input)
};
int gn_apply(const ApplyArgs& a, cudaStream_t st);

// nn.LayerNorm(C), eps 1e-5 (attention.py:199-201): fp32 [rows, C] -> bf16 [rows, C]
int layernorm_op16(const float* x, const float* gamma, const float* beta, int64_t rows, int C, op_t* out, cudaStream_t st);
// same with a 16-bit input (the transformer's inner residual stream is carried in op_t, unet_plan.cu); out may not alias x
int layernorm_op16_in16(const op_t* x, const float* gamma, const float* beta, int64_t rows, int C, op_t* out, cudaStream_t st);

// timestep_embedding (utils_diffusion.py:8-28): out bf16 [n, dim] = [cos(t*f), sin(t*f)]
int timestep_embedding_bf16(const int* t_dev, int n, int dim, op_t* out, cudaStream_t st);
// out bf16 = silu(a (+ b)); optionally also writes the fp32 sum
int silu_add_bf16(const float* a, const float* b, int64_t n, op_t* out, cudaStream_t st);
int f32_to_bf16(const float* a, int64_t n, op_t* out, cudaStream_t st);

// x [B, C, T, H, W] fp32 (reference layout) -> zero-bordered bf16 [B*T, H+2, W+2, Cp] (channels >= C zero)
int ncthw_to_pad(const float* x, int B, int C, int T, int H, int W, int Cp, op_t* dst, cudaStream_t st);
// same without the border: tokens [B*T, H, W, Cp] (the 4-D tensor-map convolution reads its halo as TMA zero fill)
int ncthw_to_tokens(const float* x, int B, int C, int T, int H, int W, int Cp, op_t* dst, cudaStream_t st);
// tokens fp32 [B*T*H*W, ld] (first C columns) -> [B, C, T, H, W] fp32
int tokens_to_ncthw(const float* tok, int ld, int B, int C, int T, int H, int W, float* y, cudaStream_t st);
// NHWC fp32 <-> padded bf16 helpers for the conv unit test
int nhwc_to_pad(const float* x, int NF, int H, int W, int C, op_t* dst, cudaStream_t st);

}  // namespace moca
