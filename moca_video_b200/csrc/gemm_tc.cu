// See gemm_tc.cuh for the contract.  sm_100a only: tcgen05.mma (UTCHMMA), tcgen05.ld (LDTM), TMA (UTMALDG).
#include <stdlib.h>
#include "gemm_tc.cuh"

namespace moca {

using namespace ptx;

constexpr int kGemmThreads = 320;        // 2 control warps + 8 epilogue warps
constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kABytes = kBM * kBK * 2;        // 16 KB, 128 rows of 128 B

// TEPI: TMA-store epilogue variants.  fp32: 3 stages, per epilogue warp 4 KB output staging + two 4 KB residual buffers.
// GEGLU: 16 epilogue warps (one 32-column output chunk each, the GELU arithmetic needs the extra warps to hide its
// latencies), 2 KB bf16 staging per warp.
// RES (bf16 thread = row epilogue only): a 16-bit residual chunk arrives by TMA in two more 2 KB boxes per warp.
template <int BN, bool TEPI = false, int EPI = EPI_F32, int NCTA = 1, bool RES = false>
struct GemmCfg {
    static constexpr bool GT = TEPI && EPI == EPI_GEGLU;
    static constexpr int B_BYTES = (BN / NCTA) * kBK * 2;                   // CTA pair: each CTA holds half of the B rows
    static constexpr int STAGE_BYTES = kABytes + B_BYTES;                   // multiple of 1024 for BN in {32, 128, 160, 192, 256}
    static constexpr bool BT = TEPI && EPI == EPI_BF16;      // bf16 thread = row epilogue: classic ring, two 2 KB staging boxes per warp
    static constexpr int STAGES = NCTA == 2 ? 6 : (GT ? 4 : ((TEPI && !BT) ? (BN == 128 ? 4 : 3) : ((BN == 256) ? (RES ? 3 : 4) : (BN == 192 ? (RES ? 3 : 4) : (BN == 160 ? (RES ? 4 : 5) : (BN == 32 ? 8 : (RES ? 5 : 6)))))));
    static constexpr int TMEM_COLS = (2 * BN <= 256) ? 256 : 512;
    static constexpr int EPI_WARPS = GT ? 16 : 8;
    static constexpr int THREADS = 64 + 32 * EPI_WARPS;
    static constexpr int WARP_STG = GT ? 2048 : ((TEPI && !BT) ? 12288 : (RES ? 8192 : 4096));      // epilogue staging per warp
    static constexpr int OFF_STG = STAGES * STAGE_BYTES;                    // 1024-byte aligned
    static constexpr int OFF_BAR = OFF_STG + EPI_WARPS * WARP_STG;
    static constexpr size_t SMEM = (size_t)OFF_BAR + 512 /*barriers*/ + 1024 /*align slack*/;
};

// packed fp32 pairs (sm_100 f32x2 arithmetic): the thread = row epilogues are bound by their instruction streams, and adds / multiplies /
// FMAs on two adjacent columns are one instruction this way (bit-identical to the scalar forms)
__device__ __forceinline__ uint64_t f2_pack(float a, float b) {
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float gelu_erf(float g) { return 0.5f * g * (1.0f + erff(g * 0.70710678118654752f)); }

// erf-GELU (attention.py:383 uses F.gelu's default) for the GEGLU epilogues.  ncu showed the 16-warp GEGLU kernel issue-bound
// (70 % issue-active, 32 warp instructions per output), so the activation is the cheapest form that stays below the bf16
// rounding of the output: 0.5 g (1 + tanh(g (a0 + a1 g^2 + a2 g^4))) with the three coefficients fitted (minimax over
// [-8, 8], g^2 clamped at 64 so the odd polynomial stays monotone) to the erf form: |error| <= 2.6e-5 for the fit plus
// 2^-11 relative for tanh.approx — 8 instructions and one MUFU op, against 17 and two for the Abramowitz-Stegun
// erfc polynomial (|error| 2e-7; build with -DMOCA_GELU_ERFC_POLY to get it back).
#ifndef MOCA_GELU_ERFC_POLY
__device__ __forceinline__ float gelu_fast(float g) {
    const float g2 = fminf(g * g, 64.0f);
    float u = fmaf(g2, -3.51516789e-04f, 3.70056460e-02f);
    u = fmaf(g2, u, 7.97507884e-01f);
    float th;
    asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(g * u));
    const float h = 0.5f * g;
    return fmaf(h, th, h);
}
// (value + bias_v) * gelu_fast(gate + bias_g) for two adjacent columns, on packed pairs: all but the clamp and the tanh are one
// instruction per two outputs (the 16-warp GEGLU epilogue is bound by its instruction stream)
__device__ __forceinline__ void geglu_pair(float v0, float v1, float g0, float g1, float bv0, float bv1, float bg0, float bg1, float& o0, float& o1) {
    const uint64_t g = f2_add(f2_pack(g0, g1), f2_pack(bg0, bg1));
    float q0, q1;
    f2_unpack(f2_mul(g, g), q0, q1);
    const uint64_t g2 = f2_pack(fminf(q0, 64.0f), fminf(q1, 64.0f));
    uint64_t u = f2_fma(g2, f2_pack(-3.51516789e-04f, -3.51516789e-04f), f2_pack(3.70056460e-02f, 3.70056460e-02f));
    u = f2_fma(g2, u, f2_pack(7.97507884e-01f, 7.97507884e-01f));
    float a0, a1, t0, t1;
    f2_unpack(f2_mul(g, u), a0, a1);
    asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(a0));
    asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(a1));
    const uint64_t h = f2_mul(g, f2_pack(0.5f, 0.5f));
    const uint64_t ge = f2_fma(h, f2_pack(t0, t1), h);
    f2_unpack(f2_mul(f2_add(f2_pack(v0, v1), f2_pack(bv0, bv1)), ge), o0, o1);
}
#else
__device__ __forceinline__ float gelu_fast(float g) {
    const float z = g * 0.70710678118654752f;
    const float az = fabsf(z);
    float t;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, az, 1.0f)));
    float p = fmaf(1.061405429f, t, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    float ex;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(az * az * -1.4426950408889634f));
    const float e = p * t * ex;
    const float one_plus_erf = z < 0.f ? e : 2.0f - e;
    return 0.5f * g * one_plus_erf;
}
#endif

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    op2_t v = floats2op2(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

// Maps an input-space row to its output row; returns false for rows that produce no output.
__device__ __forceinline__ bool map_row(const GemmParams& P, int row, int& out_row) {
    if (row >= P.M) return false;
    if (P.mode == GEMM_LINEAR) {
        out_row = row;
        return true;
    }
    if (P.mode == GEMM_SPATIAL_4D) {
        // row = (tile, frame in tile, image row in tile, column); tile = (frame group, row group)
        const int mt = row >> 7, r = row & 127;
        const int fg = mt / P.c4_th, th = mt - fg * P.c4_th;
        const int per_f = P.c4_bh * P.g_W;
        const int fl = r / per_f, rr = r - fl * per_f;
        const int hl = rr / P.g_W, w = rr - hl * P.g_W;
        out_row = ((fg * P.c4_bf + fl) * P.g_H + th * P.c4_bh + hl) * P.g_W + w;
        return true;
    }
    const int f = row / P.g_rows;
    const int r = row - f * P.g_rows;
    const int a = r / P.g_w;
    const int b = r - a * P.g_w;
    if (P.mode == GEMM_SPATIAL_PAD) {
        if (a < 1 || a > P.g_H || b < 1 || b > P.g_W) return false;
        const int h = a - 1, w = b - 1;
        if (P.stride == 2) {
            if ((h | w) & 1) return false;
            out_row = f * ((P.g_H >> 1) * (P.g_W >> 1)) + (h >> 1) * (P.g_W >> 1) + (w >> 1);
        } else {
            out_row = f * (P.g_H * P.g_W) + h * P.g_W + w;
        }
        return true;
    }
    // GEMM_TIME_PAD: a = padded frame index, b = position in frame
    if (a < 1 || a > P.g_H) return false;
    out_row = (f * P.g_H + (a - 1)) * P.g_w + b;
    return true;
}

// STATS: GroupNorm statistics of the stored values (EPI_F32); for the bf16 thread = row epilogue (TEPI && EPI_BF16) the same
// flag selects the variant that adds a 16-bit residual (RES16).
template <int BN, int EPI, bool STATS, bool TEPI = false, int NCTA = 1>
__global__ void __launch_bounds__((GemmCfg<BN, TEPI, EPI, NCTA, TEPI && EPI == EPI_BF16 && STATS>::THREADS), 1) gemm_tc_kernel(const __grid_constant__ GemmParams P) {
    using Cfg = GemmCfg<BN, TEPI, EPI, NCTA, TEPI && EPI == EPI_BF16 && STATS>;
    static_assert(EPI == EPI_F32 || EPI == EPI_GEGLU || EPI == EPI_BF16, "unknown epilogue");
    static_assert(NCTA == 1 || (NCTA == 2 && !TEPI && EPI == EPI_F32), "the CTA-pair shape exists for the staged fp32 epilogue");
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + Cfg::OFF_BAR);
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + 2;
    uint64_t* rfull = tempty + 2;                          // [8 warps][2] residual chunk landed (TEPI)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rfull + 16);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    // Tile sequence of this CTA.  NCTA == 2 (launched as clusters of 2): a tile index names a PAIR tile of 256 rows, pair
    // pm = tile / num_n_tiles; this CTA computes m-tile 2 pm + rank of it (rank 1 may fall off the end of an odd M: its rows are
    // >= M, TMA zero-fills them and the epilogue stores nothing, but it still takes part in every barrier).
    const int cta_rank = NCTA == 2 ? (int)cluster_ctarank() : 0;
    const int tile0 = NCTA == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int tstep = NCTA == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int ntiles = NCTA == 2 ? ((P.num_m_tiles + 1) >> 1) * P.num_n_tiles : P.num_m_tiles * P.num_n_tiles;
    auto MT = [&](int tile) { return NCTA == 2 ? 2 * (tile / P.num_n_tiles) + cta_rank : tile / P.num_n_tiles; };
    const int kiters = P.ntaps * P.kblocks;

    if (threadIdx.x == 0) {
        prefetch_tmap(&P.tmA);
        prefetch_tmap(&P.tmB);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull[i], 1);
            mbar_init(&tempty[i], Cfg::EPI_WARPS * NCTA);       // pair: the leader's barrier collects the epilogue warps of both CTAs
        }
        if constexpr (TEPI) {
            prefetch_tmap(&P.tmOut);
            prefetch_tmap(&P.tmRes);
            for (int i = 0; i < 16; ++i) mbar_init(&rfull[i], 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        if constexpr (NCTA == 2) { tmem_alloc2(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish2(); }
        else { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
    }
    tc_fence_before();
    if constexpr (NCTA == 2) cluster_sync_all();             // the peer's barriers exist before anything arrives on them
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // everything above overlaps the tail of the previous kernel (PDL, common.cuh); global memory is touched only below
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = tile0; tile < ntiles; tile += tstep) {
                const int mt = MT(tile), nt = tile % P.num_n_tiles;
                // GEMM_SPATIAL_4D: first frame / first input image row of the tile
                const int c4_fg = P.mode == GEMM_SPATIAL_4D ? mt / P.c4_th : 0;
                const int c4_t = mt - c4_fg * P.c4_th, c4_hq = c4_t / P.c4_tw;       // c4_tw > 1: 128-pixel segments of wide rows
                const int c4_f0 = c4_fg * P.c4_bf, c4_h0 = c4_hq * P.c4_bh * P.stride, c4_w0 = (c4_t - c4_hq * P.c4_tw) * kBM * P.stride - P.c4_pad;
                for (int tap = 0; tap < P.ntaps; ++tap) {
                    const int arow = mt * kBM + P.tap_shift[tap];
                    const int dy = tap / 3, dx = tap - dy * 3;
                    for (int kb = 0; kb < P.kblocks; ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        uint8_t* sa = base + (size_t)stage * Cfg::STAGE_BYTES;
                        if (NCTA == 1 && P.mode == GEMM_SPATIAL_4D) {
                            mbar_expect_tx(&full[stage], Cfg::STAGE_BYTES);
                            tma_load_4d(sa, &P.tmA, &full[stage], kb * kBK, c4_w0 + dx, c4_h0 + dy - P.c4_pad, c4_f0);
                            tma_load_2d(sa + kABytes, &P.tmB, &full[stage], tap * P.Kp + kb * kBK, nt * BN);
                        } else if constexpr (NCTA == 2) {
                            // both CTAs load their A rows and their half of the B rows; all bytes are counted on the leader's barrier
                            if (cta_rank == 0) mbar_expect_tx(&full[stage], 2 * Cfg::STAGE_BYTES);
                            tma_load_2d_pair(sa, &P.tmA, &full[stage], kb * kBK, arow);
                            tma_load_2d_pair(sa + kABytes, &P.tmB, &full[stage], tap * P.Kp + kb * kBK, nt * BN + cta_rank * (BN / 2));
                        } else {
                            mbar_expect_tx(&full[stage], Cfg::STAGE_BYTES);
                            tma_load_2d(sa, &P.tmA, &full[stage], kb * kBK, arow);
                            tma_load_2d(sa + kABytes, &P.tmB, &full[stage], tap * P.Kp + kb * kBK, nt * BN);
                        }
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && cta_rank == 0) {
            // ===== MMA issuer (pair: the leader CTA issues 256 x BN x 16 MMAs that read A / B halves from both CTAs' smem) =====
            constexpr uint32_t idesc = umma_idesc_bf16(kBM * NCTA, BN);
            int stage = 0;
            uint32_t phase = 0;
            int as = 0;
            uint32_t aphase = 0;
            for (int tile = tile0; tile < ntiles; tile += tstep) {
                mbar_wait(&tempty[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
                for (int it = 0; it < kiters; ++it) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(base + (size_t)stage * Cfg::STAGE_BYTES);
                    const uint64_t da = umma_desc_sw128(sa);
                    const uint64_t db = umma_desc_sw128(sa + kABytes);
#pragma unroll
                    for (int k = 0; k < kBK / 16; ++k) {
                        // advance 16 bf16 = 32 bytes inside the 128-byte swizzle atom: +2 in the (addr >> 4) field
                        if constexpr (NCTA == 2) umma_bf16_pair(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((it | k) != 0));
                        else umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((it | k) != 0));
                    }
                    if constexpr (NCTA == 2) umma_commit_pair(&empty[stage], 3);
                    else umma_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if constexpr (NCTA == 2) umma_commit_pair(&tfull[as], 3);
                else umma_commit(&tfull[as]);
                as ^= 1;
                if (as == 0) aphase ^= 1;
            }
        }
    } else if constexpr (TEPI && EPI == EPI_GEGLU) {
        // ===== GEGLU, thread = row, 16 warps: warp (q, c) owns output columns c*32 .. +31 of the tile's 128 for the rows of
        // TMEM lane quarter q: value and gate straight from TMEM (16 columns at a time), x * gelu(gate) in registers, bf16 row
        // segment into a 32 x 64-byte staging box, one TMA store per tile and warp.
        const int w = warp - 2, q = warp & 3, c = w >> 2;
        uint8_t* stgb = base + Cfg::OFF_STG + (size_t)w * Cfg::WARP_STG;
        const int ncols_total = P.N >> 1;
        int as = 0;
        uint32_t aphase = 0;
        for (int tile = tile0; tile < ntiles; tile += tstep) {
            const int mt = MT(tile), nt = tile % P.num_n_tiles;
            const int row0 = mt * kBM + q * 32;
            const int n0 = nt * (BN / 2) + c * 32;
            mbar_wait(&tfull[as], aphase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
            if (row0 < P.M && n0 < ncols_total) {
                if (lane == 0) bulk_wait_read0();
                __syncwarp();
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t rv[16], rg[16];
                    tmem_ld16(t_row + (uint32_t)(c * 32 + h * 16), rv);
                    tmem_ld16(t_row + (uint32_t)(BN / 2 + c * 32 + h * 16), rg);
                    const float4* bv = reinterpret_cast<const float4*>(P.bias + (size_t)nt * BN + c * 32 + h * 16);
                    const float4* bg = reinterpret_cast<const float4*>(P.bias + (size_t)nt * BN + BN / 2 + c * 32 + h * 16);
                    tmem_ld_wait();
                    uint32_t pk[8];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const float4 b1 = __ldg(bv + u), b2 = __ldg(bg + u);
#ifndef MOCA_GELU_ERFC_POLY
                        float o0, o1, o2, o3;
                        geglu_pair(__uint_as_float(rv[4 * u]), __uint_as_float(rv[4 * u + 1]), __uint_as_float(rg[4 * u]), __uint_as_float(rg[4 * u + 1]),
                                   b1.x, b1.y, b2.x, b2.y, o0, o1);
                        geglu_pair(__uint_as_float(rv[4 * u + 2]), __uint_as_float(rv[4 * u + 3]), __uint_as_float(rg[4 * u + 2]), __uint_as_float(rg[4 * u + 3]),
                                   b1.z, b1.w, b2.z, b2.w, o2, o3);
#else
                        const float o0 = (__uint_as_float(rv[4 * u]) + b1.x) * gelu_fast(__uint_as_float(rg[4 * u]) + b2.x);
                        const float o1 = (__uint_as_float(rv[4 * u + 1]) + b1.y) * gelu_fast(__uint_as_float(rg[4 * u + 1]) + b2.y);
                        const float o2 = (__uint_as_float(rv[4 * u + 2]) + b1.z) * gelu_fast(__uint_as_float(rg[4 * u + 2]) + b2.z);
                        const float o3 = (__uint_as_float(rv[4 * u + 3]) + b1.w) * gelu_fast(__uint_as_float(rg[4 * u + 3]) + b2.w);
#endif
                        pk[2 * u] = pack_bf16(o0, o1);
                        pk[2 * u + 1] = pack_bf16(o2, o3);
                    }
                    uint4* dst = reinterpret_cast<uint4*>(stgb + lane * 64 + h * 32);
                    dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&P.tmOut, stgb, n0, row0);
                    bulk_commit_group();
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[as]);
            as ^= 1;
            if (as == 0) aphase ^= 1;
        }
        if (lane == 0) bulk_wait0();
    } else if constexpr (TEPI && EPI == EPI_BF16) {
        // ===== 16-bit output, thread = row: TMEM -> (+ bias) (+ 16-bit residual) -> 16-bit row segment in one of two 32 x 64-byte
        // staging boxes -> TMA store.  RES16: the residual chunk of the same rows / columns arrives by TMA in one of two more boxes,
        // requested two chunks ahead (the output may alias the residual: a chunk is read before it is stored).
        constexpr bool RES16 = STATS;
        const int w = warp - 2, q = warp & 3, half = w >> 2;
        uint8_t* stgb = base + Cfg::OFF_STG + (size_t)w * Cfg::WARP_STG;
        uint64_t* rbar = rfull + 2 * w;
        constexpr int kChunks = BN / 32;
        // the two warps of a lane quarter take alternate 32-column chunks; with an odd chunk count (BN = 160: 3 + 2) they swap
        // roles from one tile of this CTA to the next, so that over two tiles (= the two accumulator stages) both do 5 chunks
        auto first_chunk = [&](int t) { return (kChunks & 1) ? (half ^ (((t - tile0) / tstep) & 1)) : half; };
        int pt = tile0, pc = first_chunk(tile0);
        uint32_t n_issued = 0, n_waited = 0;
        auto norm = [&](int& t, int& c) {
            while (t < ntiles && !(c < kChunks && (t % P.num_n_tiles) * BN + c * 32 < P.N)) { t += tstep; c = first_chunk(t); }
        };
        auto issue_res = [&]() {                               // lane 0 only
            while (pt < ntiles) {
                const int prow0 = MT(pt) * kBM + q * 32;
                const int pn0 = (pt % P.num_n_tiles) * BN + pc * 32;
                pc += 2; norm(pt, pc);
                if (prow0 >= P.M) continue;
                const uint32_t slot = n_issued & 1u;
                mbar_expect_tx(&rbar[slot], 2048);
                tma_load_2d(stgb + 4096 + slot * 2048, &P.tmRes, &rbar[slot], pn0, prow0);
                ++n_issued;
                return;
            }
        };
        if constexpr (RES16) {
            norm(pt, pc);
            if (lane == 0) { issue_res(); issue_res(); }
        }
        int as = 0;
        uint32_t aphase = 0, nst = 0;
        for (int tile = tile0; tile < ntiles; tile += tstep) {
            const int mt = MT(tile), nt = tile % P.num_n_tiles;
            const int row0 = mt * kBM + q * 32;
            mbar_wait(&tfull[as], aphase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
            if (row0 < P.M) {
                for (int c = first_chunk(tile); c < kChunks; c += 2) {
                    const int n0 = nt * BN + c * 32;
                    if (n0 >= P.N) break;
                    uint32_t r[32];
                    tmem_ld32(t_row + (uint32_t)(c * 32), r);
                    const uint4* rsrc = reinterpret_cast<const uint4*>(stgb + 4096 + (n_waited & 1u) * 2048 + lane * 64);
                    if constexpr (RES16) mbar_wait(&rbar[n_waited & 1u], (n_waited >> 1) & 1u);
                    if (lane == 0) bulk_wait_read1();          // the store issued two chunks ago has released this box
                    __syncwarp();
                    tmem_ld_wait();
                    uint4* dst = reinterpret_cast<uint4*>(stgb + (nst & 1u) * 2048 + lane * 64);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        // packed fp32 pairs: one add per two columns for the bias and for the residual (same values as scalar adds)
                        uint64_t v[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) v[e] = f2_pack(__uint_as_float(r[8 * u + 2 * e]), __uint_as_float(r[8 * u + 2 * e + 1]));
                        if (P.bias) {
                            const float4 b1 = __ldg(reinterpret_cast<const float4*>(P.bias + n0) + 2 * u);
                            const float4 b2 = __ldg(reinterpret_cast<const float4*>(P.bias + n0) + 2 * u + 1);
                            v[0] = f2_add(v[0], f2_pack(b1.x, b1.y)); v[1] = f2_add(v[1], f2_pack(b1.z, b1.w));
                            v[2] = f2_add(v[2], f2_pack(b2.x, b2.y)); v[3] = f2_add(v[3], f2_pack(b2.z, b2.w));
                        }
                        if constexpr (RES16) {
                            const uint4 t4 = rsrc[u];
                            const float2 f0 = op22float2(*reinterpret_cast<const op2_t*>(&t4.x)), f1 = op22float2(*reinterpret_cast<const op2_t*>(&t4.y));
                            const float2 f2 = op22float2(*reinterpret_cast<const op2_t*>(&t4.z)), f3 = op22float2(*reinterpret_cast<const op2_t*>(&t4.w));
                            v[0] = f2_add(v[0], f2_pack(f0.x, f0.y)); v[1] = f2_add(v[1], f2_pack(f1.x, f1.y));
                            v[2] = f2_add(v[2], f2_pack(f2.x, f2.y)); v[3] = f2_add(v[3], f2_pack(f3.x, f3.y));
                        }
                        float o[8];
#pragma unroll
                        for (int e = 0; e < 4; ++e) f2_unpack(v[e], o[2 * e], o[2 * e + 1]);
                        dst[u] = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_2d(&P.tmOut, stgb + (nst & 1u) * 2048, n0, row0);
                        bulk_commit_group();
                        if constexpr (RES16) issue_res();      // the residual box just read is free again
                    }
                    ++nst;
                    if constexpr (RES16) ++n_waited;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[as]);
            as ^= 1;
            if (as == 0) aphase ^= 1;
        }
        if (lane == 0) bulk_wait0();
    } else if constexpr (TEPI) {
        // ===== TMA-store epilogue (linear / time-padded rows, 32-row slices that map to 32 consecutive output rows) =====
        // Warp = (lane quarter q, half): alternate 32-column chunks of the tile.  Everything stays in the thread = row layout
        // of tcgen05.ld: the residual chunk arrives by TMA in a 128-byte-swizzled 32 x 32 fp32 box (two buffers per warp, the
        // load for chunk k + 2 is issued as soon as chunk k has been read), the thread adds bias + residual to its row and
        // writes it into a swizzled staging box that one lane hands to a TMA store.  GroupNorm partial sums are column sums
        // of the staging box (lane = column).  ~10x fewer instructions per chunk than the staged-and-coalesced epilogue below.
        const int w = warp - 2, q = warp & 3, half = w >> 2;
        uint8_t* wbuf = base + Cfg::OFF_STG + (size_t)w * Cfg::WARP_STG;          // [out 4 KB][res0 4 KB][res1 4 KB]
        uint64_t* rbar = rfull + 2 * w;
        constexpr int kChunks = BN / 32;
        const bool has_res = P.residual != nullptr;
        const bool out16 = P.out16 != 0;
        // slice of 32 input-space rows -> first output row (or -1: padding frame / beyond M)
        auto slice_row = [&](int tile) -> int {
            const int row0 = MT(tile) * kBM + q * 32;
            if (row0 >= P.M) return -1;
            if (P.mode == GEMM_LINEAR) return row0;
            const int f = row0 / P.g_rows, r = row0 - f * P.g_rows;
            const int a = r / P.g_w, b = r - a * P.g_w;
            return (a < 1 || a > P.g_H) ? -1 : (f * P.g_H + (a - 1)) * P.g_w + b;
        };
        // chunk cursor: (tile, c) with c = c0, c0 + 2, ..; normalised so that the chunk exists.  With an odd chunk count
        // (BN = 160: 3 + 2) the two warps of a lane quarter swap roles from one tile of this CTA to the next.
        auto first_chunk = [&](int t) { return (kChunks & 1) ? (half ^ (((t - tile0) / tstep) & 1)) : half; };
        auto norm = [&](int& t, int& c) {
            while (t < ntiles && !(c < kChunks && (t % P.num_n_tiles) * BN + c * 32 < P.N)) { t += tstep; c = first_chunk(t); }
        };
        auto next = [&](int& t, int& c) { c += 2; norm(t, c); };
        // producer cursor runs two loaded chunks ahead
        int pt = tile0, pc = first_chunk(tile0);
        uint32_t n_issued = 0, n_waited = 0;
        norm(pt, pc);
        auto issue_res = [&]() {                               // lane 0 only
            while (pt < ntiles) {
                const int orow = slice_row(pt);
                const int n0 = (pt % P.num_n_tiles) * BN + pc * 32;
                next(pt, pc);
                if (orow < 0) continue;
                const uint32_t slot = n_issued & 1u;
                mbar_expect_tx(&rbar[slot], 4096);
                tma_load_2d(wbuf + 4096 + slot * 4096, &P.tmRes, &rbar[slot], n0, orow);
                ++n_issued;
                return;
            }
        };
        if (has_res && lane == 0) { issue_res(); issue_res(); }
        int as = 0;
        uint32_t aphase = 0;
        for (int tile = tile0; tile < ntiles; tile += tstep) {
            const int orow0 = slice_row(tile);
            const int col_tile0 = (tile % P.num_n_tiles) * BN;
            mbar_wait(&tfull[as], aphase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
            if (orow0 >= 0) {
                for (int c = first_chunk(tile); c < kChunks; c += 2) {
                    const int n0 = col_tile0 + c * 32;
                    if (n0 >= P.N) break;
                    uint32_t r[32];
                    tmem_ld32(t_row + (uint32_t)(c * 32), r);
                    const uint8_t* rsrc = wbuf + 4096 + (n_waited & 1u) * 4096 + lane * 128;
                    if (has_res) mbar_wait(&rbar[n_waited & 1u], (n_waited >> 1) & 1u);
                    // the previous store of this warp must have finished reading the staging box
                    if (lane == 0) bulk_wait_read0();
                    __syncwarp();
                    tmem_ld_wait();
                    uint8_t* dst = wbuf + (out16 ? lane * 64 : lane * 128);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        // bias and residual as packed-pair adds (same values as scalar adds)
                        uint64_t lo = f2_pack(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]));
                        uint64_t hi = f2_pack(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                        if (P.bias) {
                            const float4 b4 = __ldg(reinterpret_cast<const float4*>(P.bias + n0) + j);
                            lo = f2_add(lo, f2_pack(b4.x, b4.y)); hi = f2_add(hi, f2_pack(b4.z, b4.w));
                        }
                        const int sw = (j ^ (lane & 7)) << 4;
                        if (has_res) {
                            const float4 t4 = *reinterpret_cast<const float4*>(rsrc + sw);
                            lo = f2_add(lo, f2_pack(t4.x, t4.y)); hi = f2_add(hi, f2_pack(t4.z, t4.w));
                        }
                        float4 v;
                        f2_unpack(lo, v.x, v.y);
                        f2_unpack(hi, v.z, v.w);
                        // 16-bit output: plain 32 x 64-byte box (row = 32 op_t), else the 128-byte-swizzled fp32 box
                        if (out16) *reinterpret_cast<uint2*>(dst + j * 8) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
                        else *reinterpret_cast<float4*>(dst + sw) = v;
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_2d(&P.tmOut, wbuf, n0, orow0);
                        bulk_commit_group();
                        if (has_res) issue_res();              // the buffer just read is free again
                    }
                    if (has_res) ++n_waited;
                    if constexpr (STATS) {
                        // lane = column: sum / sum of squares over the 32 rows of the staging box (conflict-free: one row per step)
                        // (same association as the staged epilogue: rows 4i + g summed over i for g = 0..3, then (p0 + p1) + (p2 + p3),
                        // so both epilogues give bit-identical statistics and the result does not depend on the tile width)
                        // one statistics entry (frame or sample) per 32-row slice: stat_rpf % 32 == 0 (gemm_plan)
                        double* dbase = P.cstat + ((size_t)(orow0 / P.stat_rpf) * P.N + n0) * 2;
                        if (out16) {
                            // statistics of the stored (rounded) values: lane = (column pair, row parity), 16 rows each (conflict-free:
                            // even rows use banks 0-15, odd rows 16-31), parities combined by one shuffle
                            const int cp = lane & 15, rp = lane >> 4;
                            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const float2 x = op22float2(*reinterpret_cast<const op2_t*>(wbuf + (2 * i + rp) * 64 + cp * 4));
                                s0 += x.x; s1 += x.y;
                                q0 = fmaf(x.x, x.x, q0); q1 = fmaf(x.y, x.y, q1);
                            }
                            s0 += __shfl_xor_sync(0xffffffffu, s0, 16); s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
                            q0 += __shfl_xor_sync(0xffffffffu, q0, 16); q1 += __shfl_xor_sync(0xffffffffu, q1, 16);
                            // lanes 0-15 publish the sums, lanes 16-31 the sums of squares of their two columns
                            double* d = dbase + (size_t)cp * 4 + rp;
                            atomicAdd(d, (double)(rp ? q0 : s0));
                            atomicAdd(d + 2, (double)(rp ? q1 : s1));
                        } else {
                        float ps[4] = {0.f, 0.f, 0.f, 0.f}, pq[4] = {0.f, 0.f, 0.f, 0.f};
                        const int cj = lane >> 2, ce = (lane & 3) << 2;
#pragma unroll
                        for (int rr = 0; rr < 32; ++rr) {
                            const float x = *reinterpret_cast<const float*>(wbuf + rr * 128 + ((cj ^ (rr & 7)) << 4) + ce);
                            ps[rr & 3] += x;
                            pq[rr & 3] = fmaf(x, x, pq[rr & 3]);
                        }
                        double* d = dbase + (size_t)lane * 2;
                        atomicAdd(d, (double)((ps[0] + ps[1]) + (ps[2] + ps[3])));
                        atomicAdd(d + 1, (double)((pq[0] + pq[1]) + (pq[2] + pq[3])));
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[as]);
            as ^= 1;
            if (as == 0) aphase ^= 1;
        }
        if (lane == 0) bulk_wait0();
    } else {
        // ===== epilogue warps 2..9: TMEM lane quarter q = warp % 4, two warps per quarter take alternate 32-column chunks.
        // Each chunk goes TMEM -> registers (thread = row) -> XOR-swizzled smem staging -> registers (8 lanes = one row's
        // 128 bytes) so that residual reads and output stores are full 128-byte segments per row.  The residual (+ per-frame
        // embedding bias) of the NEXT chunk is loaded before the current one is processed, across tile boundaries, so a
        // warp always has a DRAM round trip in flight behind its TMEM / staging work.
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        float* stg = reinterpret_cast<float*>(base + Cfg::OFF_STG) + (size_t)(warp - 2) * 1024;
        constexpr bool kGeglu = EPI == EPI_GEGLU;
        constexpr int kChunks = kGeglu ? BN / 64 : BN / 32;
        const int ncols_total = kGeglu ? (P.N >> 1) : P.N;
        int as = 0;
        uint32_t aphase = 0;

        // frame_u: the frame all valid rows of this warp's 32-row slice belong to, or -1 when they straddle frames
        auto tile_rows = [&](int tile, int& out_row, uint32_t& vmask, int& frame_u) {
            const int mt = MT(tile);
            const int row = mt * kBM + q * 32 + lane;
            out_row = 0;
            const bool valid = map_row(P, row, out_row);
            vmask = __ballot_sync(0xffffffffu, valid);
            frame_u = -1;
            if constexpr (STATS) {
                const int fr = out_row / P.stat_rpf;
                const int lo = __reduce_min_sync(0xffffffffu, valid ? fr : 0x7fffffff);
                const int hi = __reduce_max_sync(0xffffffffu, valid ? fr : -1);
                if (lo == hi && P.stat_fast) frame_u = lo;
            }
        };
        auto col0_of = [&](int tile) {
            const int nt = tile % P.num_n_tiles;
            return kGeglu ? nt * (BN / 2) : nt * BN;
        };
        // fp32 path: lane covers columns cc*4..cc*4+3 of rows rr = i*4 + lane/8
        auto prefetch = [&](float4 (&pre)[8], int out_row, uint32_t vmask, int n0) {
            const int col = n0 + (lane & 7) * 4;
            const bool full = col + 4 <= P.N;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int rr = i * 4 + (lane >> 3);
                const int orow = __shfl_sync(0xffffffffu, out_row, rr);
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (full && ((vmask >> rr) & 1u)) {
                    if (P.residual) v = __ldg(reinterpret_cast<const float4*>(P.residual + (size_t)orow * P.ldo + col));
                    if (P.rowbias) {
                        const float4 t = __ldg(reinterpret_cast<const float4*>(P.rowbias + (size_t)(orow / P.rows_per_frame) * P.rb_ld + col));
                        v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
                    }
                }
                pre[i] = v;
            }
        };

        int tile = tile0;
        int out_row = 0, nxt_row = 0;
        int frame_u = -1, nxt_frame = -1;
        uint32_t vmask = 0, nxt_mask = 0;
        float4 pre[8];
        bool have_pre = false;
        if (tile < ntiles) {
            tile_rows(tile, out_row, vmask, frame_u);
            if (EPI == EPI_F32 && col0_of(tile) + half * 32 < ncols_total) {
                prefetch(pre, out_row, vmask, col0_of(tile) + half * 32);
                have_pre = true;
            }
        }
        for (; tile < ntiles; tile += tstep) {
            const int nt = tile % P.num_n_tiles;
            const int col_tile0 = col0_of(tile);
            const int ntile = tile + tstep;
            bool nxt_ready = false;
            mbar_wait(&tfull[as], aphase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
            for (int c = half; c < kChunks; c += 2) {
                const int n0 = col_tile0 + c * 32;                 // first output column of this chunk
                if (n0 >= ncols_total) break;
                uint32_t r[32];
                tmem_ld32(t_row + (uint32_t)(c * 32), r);
                if constexpr (EPI == EPI_F32) {
                    // issue the next chunk's residual loads before touching this chunk's data
                    float4 nxt[8];
                    bool have_nxt = false;
                    if (c + 2 < kChunks && n0 + 64 < ncols_total) {
                        prefetch(nxt, out_row, vmask, n0 + 64);
                        have_nxt = true;
                    } else if (ntile < ntiles) {
                        if (!nxt_ready) { tile_rows(ntile, nxt_row, nxt_mask, nxt_frame); nxt_ready = true; }
                        if (col0_of(ntile) + half * 32 < ncols_total) {
                            prefetch(nxt, nxt_row, nxt_mask, col0_of(ntile) + half * 32);
                            have_nxt = true;
                        }
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<uint4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                    __syncwarp();
                    const int cc = lane & 7;
                    const int col = n0 + cc * 4;
                    const int nvalid = min(4, P.N - col);          // <= 0: nothing to store for this lane
                    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (P.bias && nvalid >= 4) b4 = __ldg(reinterpret_cast<const float4*>(P.bias + col));
                    float4 vv[8];                                    // stored values (zero for rows / columns not stored)
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int rr = i * 4 + (lane >> 3);
                        const int orow = __shfl_sync(0xffffffffu, out_row, rr);
                        vv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (!((vmask >> rr) & 1u) || nvalid <= 0) continue;
                        float4 v = *reinterpret_cast<const float4*>(stg + rr * 32 + ((cc ^ (rr & 7)) << 2));
                        const size_t o = (size_t)orow * P.ldo + col;
                        if (nvalid >= 4) {
                            v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
                            if (have_pre) { v.x += pre[i].x; v.y += pre[i].y; v.z += pre[i].z; v.w += pre[i].w; }
                            if (P.out16) {
                                // statistics are those of the stored (rounded) values, as in the TMA-store epilogue
                                const uint2 h = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
                                *reinterpret_cast<uint2*>(P.outb + o) = h;
                                const float2 lo = op22float2(*reinterpret_cast<const op2_t*>(&h.x)), hi = op22float2(*reinterpret_cast<const op2_t*>(&h.y));
                                v = make_float4(lo.x, lo.y, hi.x, hi.y);
                            } else {
                                *reinterpret_cast<float4*>(P.outf + o) = v;
                            }
                            vv[i] = v;
                        } else {                                     // ragged last columns (N % 4 != 0): scalar path
                            const float sv[4] = {v.x, v.y, v.z, v.w};
                            for (int e = 0; e < nvalid; ++e) {
                                float t = sv[e] + (P.bias ? P.bias[col + e] : 0.f);
                                if (P.rowbias) t += P.rowbias[(size_t)(orow / P.rows_per_frame) * P.rb_ld + col + e];
                                if (P.residual) t += P.residual[o + e];
                                if (P.out16) P.outb[o + e] = float2op(t);
                                else P.outf[o + e] = t;
                            }
                        }
                    }
                    if constexpr (STATS) {
                        if (frame_u >= 0) {
                            // all valid rows of the slice are one frame: fp32 partials, 8 rows per lane -> 32 rows
                            // (lanes with equal (lane & 7) hold the same 4 columns)
                            float4 cs = make_float4(0.f, 0.f, 0.f, 0.f), cq = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                cs.x += vv[i].x; cs.y += vv[i].y; cs.z += vv[i].z; cs.w += vv[i].w;
                                cq.x = fmaf(vv[i].x, vv[i].x, cq.x); cq.y = fmaf(vv[i].y, vv[i].y, cq.y);
                                cq.z = fmaf(vv[i].z, vv[i].z, cq.z); cq.w = fmaf(vv[i].w, vv[i].w, cq.w);
                            }
#pragma unroll
                            for (int o = 8; o <= 16; o <<= 1) {
                                cs.x += __shfl_xor_sync(0xffffffffu, cs.x, o); cs.y += __shfl_xor_sync(0xffffffffu, cs.y, o);
                                cs.z += __shfl_xor_sync(0xffffffffu, cs.z, o); cs.w += __shfl_xor_sync(0xffffffffu, cs.w, o);
                                cq.x += __shfl_xor_sync(0xffffffffu, cq.x, o); cq.y += __shfl_xor_sync(0xffffffffu, cq.y, o);
                                cq.z += __shfl_xor_sync(0xffffffffu, cq.z, o); cq.w += __shfl_xor_sync(0xffffffffu, cq.w, o);
                            }
                            const int e = lane >> 3;                   // lane group e publishes column col + e
                            const float s1 = e == 0 ? cs.x : (e == 1 ? cs.y : (e == 2 ? cs.z : cs.w));
                            const float s2 = e == 0 ? cq.x : (e == 1 ? cq.y : (e == 2 ? cq.z : cq.w));
                            if (nvalid >= 4) {
                                double* d = P.cstat + ((size_t)frame_u * P.N + col + e) * 2;
                                atomicAdd(d, (double)s1);
                                atomicAdd(d + 1, (double)s2);
                            }
                        } else if (vmask != 0u && P.cstat) {
                            // slice straddles frames (or samples are not 32-row aligned): exact double partials per lane,
                            // flushed when the frame changes — the result does not depend on the row grouping
                            double ds[4] = {0.0, 0.0, 0.0, 0.0}, dq[4] = {0.0, 0.0, 0.0, 0.0};
                            int dfr = -1;
#pragma unroll 1
                            for (int i = 0; i <= 8; ++i) {
                                const int rr = (i & 7) * 4 + (lane >> 3);
                                const int orow = __shfl_sync(0xffffffffu, out_row, rr);
                                const bool ok = i < 8 && ((vmask >> rr) & 1u) && nvalid >= 4;
                                const int fr = ok ? orow / P.stat_rpf : -1;
                                if ((fr != dfr || i == 8) && dfr >= 0) {
                                    double* d = P.cstat + ((size_t)dfr * P.N + col) * 2;
#pragma unroll
                                    for (int e = 0; e < 4; ++e) {
                                        atomicAdd(d + 2 * e, ds[e]);
                                        atomicAdd(d + 2 * e + 1, dq[e]);
                                        ds[e] = 0.0; dq[e] = 0.0;
                                    }
                                }
                                dfr = fr;
                                if (ok) {
                                    float4 t = vv[0];
#pragma unroll
                                    for (int k = 1; k < 8; ++k)
                                        if (i == k) t = vv[k];
                                    const float sv[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                                    for (int e = 0; e < 4; ++e) {
                                        ds[e] += (double)sv[e];
                                        dq[e] = fma((double)sv[e], (double)sv[e], dq[e]);
                                    }
                                }
                            }
                        }
                    }
                    __syncwarp();
                    have_pre = have_nxt;
#pragma unroll
                    for (int i = 0; i < 8; ++i) pre[i] = nxt[i];
                } else {
                    // bf16 output: 4 lanes cover one row's 32 columns (64 bytes), 8 rows per pass
                    const int cc = lane & 3;
                    const int col = n0 + cc * 8;
                    const int nvalid = min(8, ncols_total - col);
                    float va[4][8];
                    float bb[8], bg[8];
                    if constexpr (kGeglu) {
                        uint32_t rg[32];
                        tmem_ld32(t_row + (uint32_t)(BN / 2 + c * 32), rg);
                        tmem_ld_wait();
                        // values first: stage, read back in the store layout, then the gates through the same buffer
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<uint4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                        __syncwarp();
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int rr = i * 8 + (lane >> 2);
                            const float4 a = *reinterpret_cast<const float4*>(stg + rr * 32 + (((2 * cc) ^ (rr & 7)) << 2));
                            const float4 b = *reinterpret_cast<const float4*>(stg + rr * 32 + (((2 * cc + 1) ^ (rr & 7)) << 2));
                            va[i][0] = a.x; va[i][1] = a.y; va[i][2] = a.z; va[i][3] = a.w;
                            va[i][4] = b.x; va[i][5] = b.y; va[i][6] = b.z; va[i][7] = b.w;
                        }
                        __syncwarp();
#pragma unroll
                        for (int j = 0; j < 32; ++j) r[j] = rg[j];
                        const float* bvp = P.bias + (size_t)nt * BN + c * 32 + cc * 8;
#pragma unroll
                        for (int e = 0; e < 8; ++e) { bb[e] = __ldg(bvp + e); bg[e] = __ldg(bvp + BN / 2 + e); }
                    } else {
                        tmem_ld_wait();
#pragma unroll
                        for (int e = 0; e < 8; ++e) bb[e] = (P.bias && e < nvalid) ? P.bias[col + e] : 0.f;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<uint4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int rr = i * 8 + (lane >> 2);
                        const int orow = __shfl_sync(0xffffffffu, out_row, rr);
                        if (!((vmask >> rr) & 1u) || nvalid <= 0) continue;
                        const float4 a = *reinterpret_cast<const float4*>(stg + rr * 32 + (((2 * cc) ^ (rr & 7)) << 2));
                        const float4 b = *reinterpret_cast<const float4*>(stg + rr * 32 + (((2 * cc + 1) ^ (rr & 7)) << 2));
                        float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                        if constexpr (kGeglu) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) v[e] = (va[i][e] + bb[e]) * gelu_fast(v[e] + bg[e]);
                        } else {
#pragma unroll
                            for (int e = 0; e < 8; ++e) v[e] += bb[e];
                            if (P.res16) {
                                const op_t* rr16 = P.res16 + (size_t)orow * P.ldo + col;
#pragma unroll
                                for (int e = 0; e < 8; ++e)
                                    if (e < nvalid) v[e] += op2float(rr16[e]);
                            }
                        }
                        op_t* o = P.outb + (size_t)orow * P.ldo + col;
                        if (nvalid >= 8) {
                            *reinterpret_cast<uint4*>(o) = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
                        } else {
                            for (int e = 0; e < nvalid; ++e) o[e] = float2op(v[e]);
                        }
                    }
                    __syncwarp();
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (NCTA == 2) mbar_arrive_cluster(&tempty[as], 0);       // the leader's MMA thread owns the accumulator ring
                else mbar_arrive(&tempty[as]);
            }
            as ^= 1;
            if (as == 0) aphase ^= 1;
            if (EPI == EPI_F32) {
                if (!nxt_ready && ntile < ntiles) tile_rows(ntile, nxt_row, nxt_mask, nxt_frame);
                out_row = nxt_row; vmask = nxt_mask; frame_u = nxt_frame;
            } else if (ntile < ntiles) {
                tile_rows(ntile, out_row, vmask, frame_u);
            }
        }
    }
    tc_fence_before();
    if constexpr (NCTA == 2) cluster_sync_all();             // no CTA of the pair exits while its peer can still reach its smem / TMEM
    else __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        if constexpr (NCTA == 2) tmem_dealloc2(tmem_base, Cfg::TMEM_COLS);
        else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

}  // namespace moca

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
namespace moca {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// bf16 row-major [rows, cols] (pitch elements) -> box of 64 columns x box_rows rows, 128-byte swizzle, zero OOB fill
int make_tmap_bf16(CUtensorMap* tm, const void* ptr, int64_t rows, int64_t cols, int64_t pitch, int box_rows) {
    EncodeTiledFn enc = get_encode();
    MOCA_REQUIRE(enc != nullptr, MOCA_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
    MOCA_REQUIRE(((uintptr_t)ptr & 15) == 0, MOCA_E_ALIGN, "TMA base pointer %p not 16-byte aligned", ptr);
    MOCA_REQUIRE((pitch * 2) % 16 == 0, MOCA_E_ALIGN, "TMA row pitch %lld bf16 is not a multiple of 16 bytes", (long long)pitch);
    MOCA_REQUIRE(box_rows >= 1 && box_rows <= 256, MOCA_E_SHAPE, "TMA box rows %d outside 1..256", box_rows);
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, MOCA_TMAP_OP16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MOCA_REQUIRE(r == CUDA_SUCCESS, MOCA_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows %lld cols %lld pitch %lld)",
                 (int)r, (long long)rows, (long long)cols, (long long)pitch);
    return MOCA_OK;
}

// GEMM_SPATIAL_4D geometry
static bool conv4d_geometry(int NF, int H, int W, int stride, int* bh, int* bf) {
    if (stride != 1 && stride != 2) return false;
    if (H % stride || W % stride) return false;
    const int Ho = H / stride, Wo = W / stride;
    if (Wo > 128) {
        // wide images: an M tile is 128 consecutive pixels of one image row (*bh = 1, *bf = 1, Wo / 128 tiles per row)
        if (Wo % 128 != 0) return false;
        *bh = 1; *bf = 1;
        return true;
    }
    if (Wo < 1 || (Wo & (Wo - 1)) != 0) return false;
    int best = 0;
    for (int h = 1; h <= Ho && Wo * h <= 128; ++h)
        if (Ho % h == 0 && 128 % (Wo * h) == 0) best = h;
    if (!best) return false;
    const int f = 128 / (Wo * best);
    if (NF % f) return false;
    *bh = best; *bf = f;
    return true;
}
bool conv4d_supported(int NF, int H, int W, int stride) {
    static const int on = [] { const char* e = getenv("MOCA_CONV4D"); return e ? atoi(e) : 1; }();
    int bh, bf;
    return on && conv4d_geometry(NF, H, W, stride, &bh, &bf);
}

// op_t tokens [NF, H, W, C] -> 4-D map (C, W, H, NF), box 64 channels x (Wo x bh x bf) pixels with element stride `stride` in w, h
static int make_tmap_conv4d(CUtensorMap* tm, const void* ptr, int NF, int H, int W, int C, int pitch, int stride, int bh, int bf) {
    EncodeTiledFn enc = get_encode();
    MOCA_REQUIRE(enc != nullptr, MOCA_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
    MOCA_REQUIRE(((uintptr_t)ptr & 15) == 0 && (pitch * 2) % 16 == 0, MOCA_E_ALIGN, "TMA conv tensor %p pitch %d not 16-byte aligned", ptr, pitch);
    const int Wo = W / stride;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)NF};
    cuuint64_t strides[3] = {(cuuint64_t)pitch * 2, (cuuint64_t)W * pitch * 2, (cuuint64_t)H * W * pitch * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)((Wo > 128 ? 128 : Wo) * stride), (cuuint32_t)(bh * stride), (cuuint32_t)bf};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = enc(tm, MOCA_TMAP_OP16, 4, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MOCA_REQUIRE(r == CUDA_SUCCESS, MOCA_E_CUDA, "cuTensorMapEncodeTiled (conv 4-D, NF %d H %d W %d C %d stride %d box %dx%dx%d) failed with CUresult %d",
                 NF, H, W, C, stride, Wo, bh, bf, (int)r);
    return MOCA_OK;
}

// fp32 row-major [rows, cols] (pitch elements) -> box of 32 columns (128 bytes) x 32 rows, 128-byte swizzle: the per-warp
// epilogue chunk of the TMA-store epilogue (residual load and output store)
int make_tmap_f32_chunk(CUtensorMap* tm, const void* ptr, int64_t rows, int64_t cols, int64_t pitch) {
    EncodeTiledFn enc = get_encode();
    MOCA_REQUIRE(enc != nullptr, MOCA_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
    MOCA_REQUIRE(((uintptr_t)ptr & 15) == 0 && (pitch * 4) % 16 == 0, MOCA_E_ALIGN, "TMA fp32 tensor %p pitch %lld not 16-byte aligned", ptr,
                 (long long)pitch);
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MOCA_REQUIRE(r == CUDA_SUCCESS, MOCA_E_CUDA, "cuTensorMapEncodeTiled (fp32 chunk) failed with CUresult %d", (int)r);
    return MOCA_OK;
}

// bf16 row-major [rows, cols] -> box of 32 columns (64 bytes) x 32 rows, no swizzle (GEGLU TMA-store epilogue)
int make_tmap_bf16_chunk(CUtensorMap* tm, const void* ptr, int64_t rows, int64_t cols, int64_t pitch) {
    EncodeTiledFn enc = get_encode();
    MOCA_REQUIRE(enc != nullptr, MOCA_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
    MOCA_REQUIRE(((uintptr_t)ptr & 15) == 0 && (pitch * 2) % 16 == 0, MOCA_E_ALIGN, "TMA bf16 tensor %p pitch %lld not 16-byte aligned", ptr,
                 (long long)pitch);
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, MOCA_TMAP_OP16, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MOCA_REQUIRE(r == CUDA_SUCCESS, MOCA_E_CUDA, "cuTensorMapEncodeTiled (bf16 chunk) failed with CUresult %d", (int)r);
    return MOCA_OK;
}

// Tile width: minimise (waves over the SMs) x (columns computed per tile) x (cost per column of that width).  The cost per
// column is the measured main-loop rate: 1412 TFLOP/s marginal with 128x256 tiles, ~1240 with 128x160, less with 128x128
// (the A tile is re-read from shared memory for fewer columns).  A sweep over the b=2 forward (per-op times, same box)
// put the ratios at 1.20 / 1.40: with the earlier 1.06 / 1.15 the 20x32 and 10x16 convolutions ran 748 / 432 tiles of
// 160 columns where 561 / 270 tiles of 256 are 15-20 % faster (1.35 ms per forward).  GEGLU weights are packed for 256.
static int pick_bn(const GemmDesc& d, int num_sms) {
    if (d.epi == EPI_GEGLU) return 256;
    // outputs of a few channels (the UNet's `out` conv, N = 4; the VAE's conv_out, N = 3): a 32-wide tile instead of 128 columns of
    // zeros; the launch is then bound by streaming A through shared memory (staged fp32 epilogue only)
    static const int bn32 = [] { const char* e = getenv("MOCA_GEMM_BN32"); return e ? atoi(e) : 1; }();
    if (bn32 && d.epi == EPI_F32 && d.N <= 32) return 32;
    // 192: 1280-wide outputs on 40 M tiles (the 10 x 16 level) are 200 tiles of 256 = 1.35 waves, but 280 tiles of 192 = 1.89
    const int cand[4] = {256, 192, 160, 128};
    static const double pen192 = [] { const char* e = getenv("MOCA_GEMM_PEN192"); return e ? atof(e) : 1.12; }();    // <= 0: never
    static const double pen160 = [] { const char* e = getenv("MOCA_GEMM_PEN160"); return e ? atof(e) : 1.20; }();
    static const double pen128 = [] { const char* e = getenv("MOCA_GEMM_PEN128"); return e ? atof(e) : 1.40; }();
    const double pen[4] = {1.0, pen192, pen160, pen128};
    const int64_t mt = (d.M + kBM - 1) / kBM;
    int best = 128;
    double best_cost = 1e30;
    for (int i = 0; i < 4; ++i) {
        const int bn = cand[i];
        if (pen[i] <= 0.0) continue;
        const int64_t tiles = mt * ((d.N + bn - 1) / bn);
        const double waves = (double)((tiles + num_sms - 1) / num_sms);
        const double cost = waves * bn * pen[i];
        if (cost < best_cost - 1e-9) { best_cost = cost; best = bn; }
    }
    return best;
}

// dynamic shared memory opt-in: a per-device attribute, so "done" is kept per device (16 devices per process at most)
static int cur_device() { int d = 0; cudaGetDevice(&d); return d < 0 || d >= 16 ? 0 : d; }

template <int BN, int EPI, bool STATS = false, bool TEPI = false>
static int set_attr() {
    static bool done_dev[16] = {false};
    bool& done = done_dev[cur_device()];
    if (!done) {
        MOCA_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI, STATS, TEPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)GemmCfg<BN, TEPI, EPI, 1, TEPI && EPI == EPI_BF16 && STATS>::SMEM));
        done = true;
    }
    return MOCA_OK;
}

template <int BN, bool STATS>
static int set_attr_pair() {
    static bool done_dev[16] = {false};
    bool& done = done_dev[cur_device()];
    if (!done) {
        MOCA_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI_F32, STATS, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)GemmCfg<BN, false, EPI_F32, 2>::SMEM));
        MOCA_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI_F32, STATS, false, 2>, cudaFuncAttributeNonPortableClusterSizeAllowed, 0));
        done = true;
    }
    return MOCA_OK;
}

// clusters of 2 CTAs (the pair is co-scheduled on one TPC)
template <typename K>
static cudaError_t launch_pair(K kernel, int grid, int threads, size_t smem, cudaStream_t st, const GemmParams& p) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_level() >= 1 ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, p);
}

template <int BN>
static int set_attr_bn(int epi, int tepi) {
    if (epi == EPI_F32) {
        if (tepi) {
            if constexpr (BN != 256) { int rc = set_attr<BN, EPI_F32, true, true>(); return rc ? rc : set_attr<BN, EPI_F32, false, true>(); }
        }
        int rc = set_attr<BN, EPI_F32, true>();
        return rc ? rc : set_attr<BN, EPI_F32>();
    }
    if (epi == EPI_BF16) {
        if (!tepi) return set_attr<BN, EPI_BF16>();
        int rc = set_attr<BN, EPI_BF16, false, true>();
        return rc ? rc : set_attr<BN, EPI_BF16, true, true>();
    }
    if (tepi) {
        if constexpr (BN == 256) return set_attr<BN, EPI_GEGLU, false, true>();
    }
    return set_attr<BN, EPI_GEGLU>();
}

template <int BN>
static void launch_bn(const GemmLaunch& l, cudaStream_t st) {
    const bool stats = l.p.cstat != nullptr;
    if (l.p.ncta == 2) {
        if constexpr (BN == 256 || BN == 160) {
            if (stats) launch_pair(gemm_tc_kernel<BN, EPI_F32, true, false, 2>, l.grid, kGemmThreads, l.smem, st, l.p);
            else launch_pair(gemm_tc_kernel<BN, EPI_F32, false, false, 2>, l.grid, kGemmThreads, l.smem, st, l.p);
        }
    } else if (l.p.epi == EPI_F32 && l.p.tepi) {
        if constexpr (BN != 256) {
            if (stats) launch_pdl_tc(gemm_tc_kernel<BN, EPI_F32, true, true>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
            else launch_pdl_tc(gemm_tc_kernel<BN, EPI_F32, false, true>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
        }
    } else if (l.p.epi == EPI_F32 && stats) launch_pdl_tc(gemm_tc_kernel<BN, EPI_F32, true>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
    else if (l.p.epi == EPI_F32) launch_pdl_tc(gemm_tc_kernel<BN, EPI_F32, false>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
    else if (l.p.epi == EPI_BF16 && l.p.tepi && l.p.res16) launch_pdl_tc(gemm_tc_kernel<BN, EPI_BF16, true, true>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
    else if (l.p.epi == EPI_BF16 && l.p.tepi) launch_pdl_tc(gemm_tc_kernel<BN, EPI_BF16, false, true>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
    else if (l.p.epi == EPI_BF16) launch_pdl_tc(gemm_tc_kernel<BN, EPI_BF16, false>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
    else if (l.p.tepi) {
        if constexpr (BN == 256)
            launch_pdl_tc(gemm_tc_kernel<BN, EPI_GEGLU, false, true>, dim3(l.grid), dim3(GemmCfg<BN, true, EPI_GEGLU>::THREADS), l.smem, st, l.p);
    } else launch_pdl_tc(gemm_tc_kernel<BN, EPI_GEGLU, false>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
}

int gemm_plan(const Ctx& ctx, const GemmDesc& d, GemmLaunch* out) {
    MOCA_REQUIRE(d.A && d.Wt && d.out, MOCA_E_ARG, "gemm_plan: NULL operand");
    MOCA_REQUIRE(d.M > 0 && d.N > 0 && d.Kp > 0 && d.Kp % 64 == 0, MOCA_E_SHAPE, "gemm_plan: bad M=%d N=%d Kp=%d", d.M, d.N, d.Kp);
    MOCA_REQUIRE(d.ntaps >= 1 && d.ntaps <= 9, MOCA_E_SHAPE, "gemm_plan: ntaps=%d", d.ntaps);
    MOCA_REQUIRE(d.a_cols <= d.Kp, MOCA_E_SHAPE, "gemm_plan: a_cols %d > Kp %d", d.a_cols, d.Kp);
    MOCA_REQUIRE(d.epi != EPI_GEGLU || (d.N % 256 == 0 && d.bias), MOCA_E_SHAPE, "gemm_plan: GEGLU needs N %% 256 == 0 and a bias");
    MOCA_REQUIRE(d.epi == EPI_GEGLU ? (d.ldo % 8 == 0) : (d.epi == EPI_BF16 ? d.ldo % 8 == 0 || d.N % 32 != 0 : d.ldo % 4 == 0 || d.N % 32 != 0),
                 MOCA_E_ALIGN, "gemm_plan: output pitch %d not vector aligned", d.ldo);
    GemmLaunch L;
    memset(&L.p, 0, sizeof(L.p));
    L.BN = pick_bn(d, ctx.num_sms);
    int rc = MOCA_OK;
    int c4_bh = 0, c4_bf = 0;
    if (d.mode == GEMM_SPATIAL_4D) {
        // d.g_H / d.g_W: INPUT image size, d.a_rows = NF * H * W, d.M = NF * Ho * Wo
        const int NF = (int)(d.a_rows / ((int64_t)d.g_H * d.g_W));
        MOCA_REQUIRE(d.ntaps == 9 && conv4d_geometry(NF, d.g_H, d.g_W, d.stride, &c4_bh, &c4_bf) && d.M % kBM == 0, MOCA_E_SHAPE,
                     "gemm_plan: unsupported 4-D conv geometry NF=%d H=%d W=%d stride=%d", NF, d.g_H, d.g_W, d.stride);
        rc = make_tmap_conv4d(&L.p.tmA, d.A, NF, d.g_H, d.g_W, d.a_cols, d.a_pitch, d.stride, c4_bh, c4_bf);
    } else {
        rc = make_tmap_bf16(&L.p.tmA, d.A, d.a_rows, d.a_cols, d.a_pitch, kBM);
    }
    if (rc) return rc;
    // CTA-pair shape (MOCA_GEMM_2CTA=1): staged fp32 epilogue, 160- / 256-wide tiles, long main loops
    static const int pair_on = [] { const char* e = getenv("MOCA_GEMM_2CTA"); return e ? atoi(e) : 0; }();
    static const int pair_mink = [] { const char* e = getenv("MOCA_GEMM_2CTA_MINK"); return e ? atoi(e) : 16; }();
    const int ncta = (pair_on && d.epi == EPI_F32 && (L.BN == 256 || L.BN == 160) && d.ntaps * (d.Kp / 64) >= pair_mink && ctx.num_sms >= 2 &&
                      d.mode != GEMM_SPATIAL_4D) ? 2 : 1;
    rc = make_tmap_bf16(&L.p.tmB, d.Wt, d.N, (int64_t)d.ntaps * d.Kp, (int64_t)d.ntaps * d.Kp, L.BN / ncta);
    if (rc) return rc;
    L.p.ncta = ncta;
    GemmParams& p = L.p;
    p.M = d.M; p.N = d.N; p.ntaps = d.ntaps; p.Kp = d.Kp; p.kblocks = d.Kp / 64;
    for (int i = 0; i < 9; ++i) p.tap_shift[i] = d.tap_shift[i];
    p.mode = d.mode; p.g_rows = d.g_rows; p.g_w = d.g_w; p.g_H = d.g_H; p.g_W = d.g_W; p.stride = d.stride;
    p.c4_tw = 1; p.c4_pad = d.pad_lo;
    if (d.mode == GEMM_SPATIAL_4D) {
        p.g_H = d.g_H / d.stride; p.g_W = d.g_W / d.stride;            // output image size from here on
        p.c4_bh = c4_bh; p.c4_bf = c4_bf; p.c4_th = p.g_H / c4_bh; p.c4_tw = 1;
        if (p.g_W > 128) {
            // wide rows: the output rows of a tile are consecutive, which map_row reproduces for an image of 128-pixel rows
            p.c4_tw = p.g_W / 128; p.c4_th = p.g_H * p.c4_tw;
            p.g_H = p.c4_th; p.g_W = 128;
        }
    }
    p.outf = (float*)d.out; p.outb = (op_t*)d.out; p.ldo = d.ldo;
    p.bias = d.bias; p.rowbias = d.rowbias; p.rb_ld = d.rb_ld; p.rows_per_frame = d.rows_per_frame > 0 ? d.rows_per_frame : 1;
    p.residual = d.residual; p.res16 = d.residual16; p.epi = d.epi; p.out16 = d.out16;
    MOCA_REQUIRE(!d.out16 || (d.epi == EPI_F32 && d.ldo % 4 == 0 && d.N % 4 == 0), MOCA_E_ARG, "gemm_plan: out16 needs the fp32 epilogue and 4-element aligned rows");
    MOCA_REQUIRE(!d.residual16 || (d.epi == EPI_BF16 && !d.residual), MOCA_E_ARG, "gemm_plan: a 16-bit residual needs the 16-bit epilogue");
    p.cstat = d.chan_stats; p.stat_rpf = d.stat_rows_per_frame > 0 ? d.stat_rows_per_frame : 1;
    // fp32 warp-slice partial sums only when every sample starts on a 32-row slice boundary, so that the grouping (and
    // with it every rounding) is the same whatever the batch size; otherwise exact double partials
    p.stat_fast = d.stat_sample_rows > 0 && d.stat_sample_rows % 32 == 0;
    MOCA_REQUIRE(!d.chan_stats || (d.epi == EPI_F32 && d.N % 4 == 0 && d.stat_rows_per_frame > 0), MOCA_E_ARG,
                 "gemm_plan: channel statistics need the fp32 epilogue, N %% 4 == 0 and rows per frame");
    p.num_m_tiles = cdiv(d.M, kBM);
    p.num_n_tiles = cdiv(d.N, L.BN);
    const int ntiles = p.num_m_tiles * p.num_n_tiles;
    L.grid = ntiles < ctx.num_sms ? ntiles : ctx.num_sms;
    if (ncta == 2) {
        const int npair = ((p.num_m_tiles + 1) / 2) * p.num_n_tiles, maxpair = ctx.num_sms / 2;
        L.grid = 2 * (npair < maxpair ? npair : maxpair);
    }
    // TMA-store epilogue: fp32 output whose 32-row slices map to 32 consecutive output rows, short K (the epilogue, not the
    // main loop, bounds these), slice-mode statistics only.  MOCA_GEMM_TEPI=0 pins the staged-and-coalesced epilogue.
    static const int tepi_on = [] { const char* e = getenv("MOCA_GEMM_TEPI"); return e ? atoi(e) : 1; }();
    static const int tepi_maxk160 = [] { const char* e = getenv("MOCA_GEMM_TEPI_MAXK"); return e ? atoi(e) : 15; }();       // 3-stage ring
    static const int tepi_maxk128 = [] { const char* e = getenv("MOCA_GEMM_TEPI_MAXK128"); return e ? atoi(e) : 15; }();    // 4-stage ring
    const int tepi_maxk = L.BN == 128 ? tepi_maxk128 : tepi_maxk160;
    const bool rows_ok = (d.mode == GEMM_LINEAR && d.M % 32 == 0) || (d.mode == GEMM_TIME_PAD && d.g_w % 32 == 0 && d.g_rows % 32 == 0);
    p.tepi = tepi_on && ncta == 1 && d.epi == EPI_F32 && L.BN != 256 && L.BN != 32 && rows_ok && d.N % 32 == 0 && d.ldo % 4 == 0 && (!d.out16 || d.ldo % 8 == 0) && !d.rowbias && (!d.chan_stats || d.stat_rows_per_frame % 32 == 0) &&
             d.ntaps * p.kblocks <= tepi_maxk && (((uintptr_t)d.out | (uintptr_t)d.residual | (uintptr_t)d.bias) & 15) == 0;
    const bool gt = tepi_on && d.epi == EPI_GEGLU && L.BN == 256 && d.mode == GEMM_LINEAR && d.M % 32 == 0 && (d.N / 2) % 32 == 0 &&
                    d.ldo % 8 == 0 && (((uintptr_t)d.out | (uintptr_t)d.bias) & 15) == 0;
    const bool bt = tepi_on && d.epi == EPI_BF16 && d.mode == GEMM_LINEAR && d.M % 32 == 0 && d.N % 32 == 0 && d.ldo % 8 == 0 &&
                    (((uintptr_t)d.out | (uintptr_t)d.bias | (uintptr_t)d.residual16) & 15) == 0;
    if (bt) {
        p.tepi = 1;
        rc = make_tmap_bf16_chunk(&p.tmOut, d.out, d.M, d.N, d.ldo);
        if (rc) return rc;
        if (d.residual16) {
            rc = make_tmap_bf16_chunk(&p.tmRes, d.residual16, d.M, d.N, d.ldo);
            if (rc) return rc;
            L.smem = L.BN == 256 ? GemmCfg<256, true, EPI_BF16, 1, true>::SMEM
                                 : (L.BN == 192 ? GemmCfg<192, true, EPI_BF16, 1, true>::SMEM
                                 : (L.BN == 160 ? GemmCfg<160, true, EPI_BF16, 1, true>::SMEM : GemmCfg<128, true, EPI_BF16, 1, true>::SMEM));
        } else
        L.smem = L.BN == 256 ? GemmCfg<256, true, EPI_BF16>::SMEM : (L.BN == 192 ? GemmCfg<192, true, EPI_BF16>::SMEM : (L.BN == 160 ? GemmCfg<160, true, EPI_BF16>::SMEM : GemmCfg<128, true, EPI_BF16>::SMEM));
    } else if (gt) {
        p.tepi = 1;
        rc = make_tmap_bf16_chunk(&p.tmOut, d.out, d.M, d.N / 2, d.ldo);
        if (rc) return rc;
        L.smem = GemmCfg<256, true, EPI_GEGLU>::SMEM;
    } else if (p.tepi) {
        const int64_t out_rows = d.mode == GEMM_LINEAR ? d.M : (int64_t)(d.M / d.g_rows) * d.g_H * d.g_w;
        rc = d.out16 ? make_tmap_bf16_chunk(&p.tmOut, d.out, out_rows, d.N, d.ldo) : make_tmap_f32_chunk(&p.tmOut, d.out, out_rows, d.N, d.ldo);
        if (rc) return rc;
        if (d.out16 && !d.residual) p.tmRes = p.tmOut;      // never loaded from; keeps the prefetched descriptor valid
        else rc = make_tmap_f32_chunk(&p.tmRes, d.residual ? (const void*)d.residual : d.out, out_rows, d.N, d.ldo);
        if (rc) return rc;
        L.smem = L.BN == 192 ? GemmCfg<192, true>::SMEM : (L.BN == 160 ? GemmCfg<160, true>::SMEM : GemmCfg<128, true>::SMEM);
    } else {
        L.smem = L.BN == 32 ? GemmCfg<32>::SMEM : L.BN == 256 ? GemmCfg<256>::SMEM : (L.BN == 192 ? GemmCfg<192>::SMEM : (L.BN == 160 ? GemmCfg<160>::SMEM : GemmCfg<128>::SMEM));
    }
    L.flops = 2.0 * (double)d.M * d.N * d.ntaps * d.a_cols;
    if (ncta == 2) {
        L.smem = L.BN == 256 ? GemmCfg<256, false, EPI_F32, 2>::SMEM : GemmCfg<160, false, EPI_F32, 2>::SMEM;
        rc = L.BN == 256 ? set_attr_pair<256, true>() : set_attr_pair<160, true>();
        if (!rc) rc = L.BN == 256 ? set_attr_pair<256, false>() : set_attr_pair<160, false>();
        if (rc) return rc;
    }
    if (L.BN == 32) { rc = set_attr<32, EPI_F32, true>(); if (!rc) rc = set_attr<32, EPI_F32>(); }
    else rc = L.BN == 256 ? set_attr_bn<256>(d.epi, p.tepi) : (L.BN == 192 ? set_attr_bn<192>(d.epi, p.tepi) : (L.BN == 160 ? set_attr_bn<160>(d.epi, p.tepi) : set_attr_bn<128>(d.epi, p.tepi)));
    if (rc) return rc;
    *out = L;
    return MOCA_OK;
}

int gemm_run(const GemmLaunch& l, cudaStream_t st) {
    if (l.BN == 32) {
        if (l.p.cstat) launch_pdl_tc(gemm_tc_kernel<32, EPI_F32, true>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
        else launch_pdl_tc(gemm_tc_kernel<32, EPI_F32, false>, dim3(l.grid), dim3(kGemmThreads), l.smem, st, l.p);
    } else if (l.BN == 256) launch_bn<256>(l, st);
    else if (l.BN == 192) launch_bn<192>(l, st);
    else if (l.BN == 160) launch_bn<160>(l, st);
    else launch_bn<128>(l, st);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

}  // namespace moca
