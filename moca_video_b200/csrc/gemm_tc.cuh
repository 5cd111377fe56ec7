// tcgen05 / TMEM / TMA GEMM core of the UNet.
//
//   acc[m, n] = sum_tap sum_k A[m + shift[tap], k] * Wt[n, tap*Kp + k]          (fp32 accumulate in TMEM)
//
// A is a bf16 channels-last token matrix [rows, K]; Wt the packed bf16 weight [N, ntaps*Kp] (K contiguous).
// One kernel serves
//   * nn.Linear / 1x1 conv           ntaps = 1
//   * 3x3 conv (ResBlock, Up/Down)   ntaps = 9 over a zero-bordered [NF, H+2, W+2, C] buffer: tap (dy,dx) is the row
//                                    shift (dy-1)*(W+2) + (dx-1); border rows are computed and dropped by the epilogue
//   * Conv3d (3,1,1) (TemporalConv)  ntaps = 3 over a [B, T+2, H*W, C] buffer with zero first/last frame: shift (dt-1)*H*W
// so the convolutions are implicit GEMMs whose "im2col" is nothing but the TMA row coordinate (rows outside the
// buffer are zero-filled by TMA).  The epilogue maps an input-space row to its output token row, adds bias /
// per-frame embedding bias / fp32 residual, or applies GEGLU, and stores fp32 or bf16.
//
// Kernel shape: persistent, 192 threads = warp 0 TMA producer, warp 1 MMA issuer (+TMEM alloc), warps 2-5 epilogue.
// 128 x BN output tile, BK = 64 (one 128-byte swizzled row), STAGES-deep smem ring, two TMEM accumulator stages so the
// epilogue of tile i overlaps the main loop of tile i+1.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace moca {

// GEMM_SPATIAL_4D: 3x3 conv on the UNPADDED token matrix [NF, H, W, C] through a 4-D tensor map (C, W, H, frame): an M tile is
// bf frames x bh image rows x Wo columns = 128 output pixels, tap (dy, dx) is the same box moved by (dx - 1, dy - 1) with the
// halo zero-filled by TMA, stride 2 is the tensor map's element stride — no border rows are computed, no 3/4 of the rows dropped.
enum GemmMode { GEMM_LINEAR = 0, GEMM_SPATIAL_PAD = 1, GEMM_TIME_PAD = 2, GEMM_SPATIAL_4D = 3 };
enum GemmEpi { EPI_F32 = 0, EPI_BF16 = 1, EPI_GEGLU = 2 };

struct GemmParams {
    CUtensorMap tmA, tmB;
    CUtensorMap tmOut, tmRes;  // TMA-store epilogue only: fp32 [out_rows, N] output / residual, 32 x 32 boxes
    int tepi;                 // 1: TMA-store epilogue (gemm_tc.cu)
    int ncta;                 // 2: CTA-pair shape (cluster of 2, tcgen05 cta_group::2, 256-row tiles)
    int M, N;                 // rows computed (input-row space), accumulator columns
    int ntaps, kblocks, Kp;   // kblocks = Kp / 64 per tap
    int tap_shift[9];
    int mode;
    int g_rows;               // SPATIAL: (H+2)*(W+2)   TIME: (T+2)*HW
    int g_w;                  // SPATIAL: W+2           TIME: HW
    int g_H, g_W;             // SPATIAL: H, W          TIME: g_H = T
    int stride;               // SPATIAL: 1 or 2 (Downsample)
    int c4_bh, c4_bf, c4_th, c4_tw, c4_pad;  // SPATIAL_4D: output image rows / frames per M tile, M tiles per frame group (Ho / bh); g_H, g_W = Ho, Wo
    float* outf;
    op_t* outb;
    int ldo;                  // output row pitch (elements)
    const float* bias;        // [N] or null
    const float* rowbias;     // [frames, rb_ld] or null (ResBlock emb_layers output, openaimodel3d.py:221-226)
    int rb_ld, rows_per_frame;
    const float* residual;    // fp32 [out_rows, ldo] or null
    const op_t* res16;        // EPI_BF16 only: 16-bit residual [out_rows, ldo] or null (may alias the output: in-place stream update)
    int epi;
    int out16;                // EPI_F32 epilogues: store the result as op_t (outb) instead of fp32; statistics unchanged
    int num_m_tiles, num_n_tiles;
    double* cstat;            // optional per-(frame, output channel) {sum, sum of squares} of the stored fp32 values
    int stat_rpf;             // output rows per frame (frame = out_row / stat_rpf)
    int stat_fast;
};

struct GemmDesc {
    const void* A = nullptr;          // bf16 [a_rows, a_cols], pitch a_pitch elements
    int64_t a_rows = 0;
    int a_cols = 0, a_pitch = 0;
    const void* Wt = nullptr;         // bf16 [N, ntaps*Kp]
    int N = 0, ntaps = 1, Kp = 0;
    int tap_shift[9] = {0};
    int mode = GEMM_LINEAR;
    int g_rows = 0, g_w = 0, g_H = 0, g_W = 0, stride = 1;
    int M = 0;
    int epi = EPI_F32;
    int out16 = 0;                    // EPI_F32 only: the fp32 epilogue (bias / row bias / residual / statistics) stores op_t
    void* out = nullptr;
    int ldo = 0;
    const float* bias = nullptr;
    const float* rowbias = nullptr;
    int rb_ld = 0, rows_per_frame = 1;
    const float* residual = nullptr;
    const op_t* residual16 = nullptr; // EPI_BF16 only: out = op16(acc + bias + residual16); may alias out
    // EPI_F32 only: accumulate {sum, sumsq} of the stored values into chan_stats[frame][N][2] (doubles, zeroed by the
    // caller) — the GroupNorm that consumes this tensor needs no statistics pass of its own.
    double* chan_stats = nullptr;
    int stat_rows_per_frame = 0;
    int pad_lo = 1;                   // SPATIAL_4D: zero rows / columns before the image (1: padding 1; 0: the VAE Downsample's pad (0,1,0,1))
    int stat_sample_rows = 0;         // rows of one sample in the GEMM's input-row space (0: unknown -> exact double partials)
};

struct GemmLaunch {
    GemmParams p;
    int BN = 128;
    int grid = 0;
    size_t smem = 0;
    double flops = 0;
};

// 2-D bf16 tensor map, box = 64 columns x box_rows rows, SWIZZLE_128B, out-of-bounds rows/columns read as zero
int make_tmap_bf16(CUtensorMap* tm, const void* ptr, int64_t rows, int64_t cols, int64_t pitch, int box_rows);
// GEMM_SPATIAL_4D geometry: W / stride a power of two <= 128, an image-row count bh | Ho with Wo * bh | 128, NF % (128 / (Wo bh)) == 0
bool conv4d_supported(int NF, int H, int W, int stride);
int gemm_plan(const Ctx& ctx, const GemmDesc& d, GemmLaunch* out);
int gemm_run(const GemmLaunch& l, cudaStream_t st);

}  // namespace moca
