// See gemm_tc.cuh for the contract.  sm_100a only: tcgen05.mma (UTCHMMA), tcgen05.ld (LDTM), TMA (UTMALDG).
#include "gemm_tc.cuh"

namespace moca {

using namespace ptx;

constexpr int kGemmThreads = 192;
constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kABytes = kBM * kBK * 2;        // 16 KB, 128 rows of 128 B

template <int BN>
struct GemmCfg {
    static constexpr int B_BYTES = BN * kBK * 2;
    static constexpr int STAGE_BYTES = kABytes + B_BYTES;
    static constexpr int STAGES = (BN == 256) ? 4 : (BN == 160 ? 5 : 6);
    static constexpr int TMEM_COLS = (2 * BN <= 256) ? 256 : 512;
    static constexpr size_t SMEM = (size_t)STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ float gelu_erf(float g) { return 0.5f * g * (1.0f + erff(g * 0.70710678118654752f)); }

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

// Maps an input-space row to its output row; returns false for rows that produce no output.
__device__ __forceinline__ bool map_row(const GemmParams& P, int row, int& out_row) {
    if (row >= P.M) return false;
    if (P.mode == GEMM_LINEAR) {
        out_row = row;
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

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_tc_kernel(const __grid_constant__ GemmParams P) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + (size_t)STAGES * Cfg::STAGE_BYTES);
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int ntiles = P.num_m_tiles * P.num_n_tiles;
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
            mbar_init(&tempty[i], 4);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const int mt = tile / P.num_n_tiles, nt = tile - mt * P.num_n_tiles;
                for (int tap = 0; tap < P.ntaps; ++tap) {
                    const int arow = mt * kBM + P.tap_shift[tap];
                    for (int kb = 0; kb < P.kblocks; ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        uint8_t* sa = base + (size_t)stage * Cfg::STAGE_BYTES;
                        mbar_expect_tx(&full[stage], Cfg::STAGE_BYTES);
                        tma_load_2d(sa, &P.tmA, &full[stage], kb * kBK, arow);
                        tma_load_2d(sa + kABytes, &P.tmB, &full[stage], tap * P.Kp + kb * kBK, nt * BN);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int as = 0;
            uint32_t aphase = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
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
                        umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((it | k) != 0));
                    }
                    umma_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(&tfull[as]);
                as ^= 1;
                if (as == 0) aphase ^= 1;
            }
        }
    } else {
        // ===== epilogue warps 2..5: TMEM lane quarter = warp % 4 =====
        const int q = warp & 3;
        int as = 0;
        uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const int mt = tile / P.num_n_tiles, nt = tile - mt * P.num_n_tiles;
            mbar_wait(&tfull[as], aphase);
            tc_fence_after();
            const int row = mt * kBM + q * 32 + lane;
            int out_row = 0;
            const bool valid = map_row(P, row, out_row);
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
            if (P.epi == EPI_GEGLU) {
                constexpr int HALF = BN / 2;
                const int nout = P.N >> 1;
                for (int c = 0; c < HALF / 32; ++c) {
                    const int n0 = nt * HALF + c * 32;          // output column
                    if (n0 >= nout) break;
                    __syncwarp();
                    uint32_t rv[32], rg[32];
                    tmem_ld32(t_row + (uint32_t)(c * 32), rv);
                    tmem_ld32(t_row + (uint32_t)(HALF + c * 32), rg);
                    tmem_ld_wait();
                    if (valid) {
                        const float* bv = P.bias + (size_t)nt * BN + c * 32;
                        const float* bg = bv + HALF;
                        __nv_bfloat16* o = P.outb + (size_t)out_row * P.ldo + n0;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            uint32_t w[4];
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const int jj = j + 2 * u;
                                const float v0 = __uint_as_float(rv[jj]) + bv[jj], v1 = __uint_as_float(rv[jj + 1]) + bv[jj + 1];
                                const float g0 = __uint_as_float(rg[jj]) + bg[jj], g1 = __uint_as_float(rg[jj + 1]) + bg[jj + 1];
                                w[u] = pack_bf16(v0 * gelu_erf(g0), v1 * gelu_erf(g1));
                            }
                            *reinterpret_cast<uint4*>(o + j) = make_uint4(w[0], w[1], w[2], w[3]);
                        }
                    }
                }
            } else {
                for (int c = 0; c < BN / 32; ++c) {
                    const int n0 = nt * BN + c * 32;
                    if (n0 >= P.N) break;
                    __syncwarp();
                    uint32_t r[32];
                    tmem_ld32(t_row + (uint32_t)(c * 32), r);
                    tmem_ld_wait();
                    __syncwarp();
                    if (!valid) continue;
                    const int ncols = min(32, P.N - n0);
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    if (P.bias) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (j < ncols) v[j] += P.bias[n0 + j];
                    }
                    if (P.rowbias) {
                        const float* rb = P.rowbias + (size_t)(out_row / P.rows_per_frame) * P.rb_ld + n0;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (j < ncols) v[j] += rb[j];
                    }
                    if (P.epi == EPI_F32) {
                        float* o = P.outf + (size_t)out_row * P.ldo + n0;
                        if (ncols == 32) {
                            if (P.residual) {
                                const float4* rs = reinterpret_cast<const float4*>(P.residual + (size_t)out_row * P.ldo + n0);
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    const float4 t = rs[j];
                                    v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
                                }
                            }
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                reinterpret_cast<float4*>(o)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                        } else {
                            const float* rs = P.residual ? P.residual + (size_t)out_row * P.ldo + n0 : nullptr;
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (j < ncols) o[j] = v[j] + (rs ? rs[j] : 0.0f);
                        }
                    } else {
                        __nv_bfloat16* o = P.outb + (size_t)out_row * P.ldo + n0;
                        if (ncols == 32) {
#pragma unroll
                            for (int j = 0; j < 32; j += 8)
                                *reinterpret_cast<uint4*>(o + j) = make_uint4(pack_bf16(v[j], v[j + 1]), pack_bf16(v[j + 2], v[j + 3]),
                                                                              pack_bf16(v[j + 4], v[j + 5]), pack_bf16(v[j + 6], v[j + 7]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (j < ncols) o[j] = __float2bfloat16_rn(v[j]);
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
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
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
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MOCA_REQUIRE(r == CUDA_SUCCESS, MOCA_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows %lld cols %lld pitch %lld)",
                 (int)r, (long long)rows, (long long)cols, (long long)pitch);
    return MOCA_OK;
}

static int pick_bn(const GemmDesc& d) {
    if (d.epi == EPI_GEGLU) return 256;
    if (d.N % 256 == 0) return 256;
    if (d.N % 160 == 0) return 160;
    if (d.N > 128 && d.N % 128 != 0 && d.N <= 160) return 160;
    return 128;
}

template <int BN>
static int set_attr() {
    static bool done = false;
    if (!done) {
        MOCA_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GemmCfg<BN>::SMEM));
        done = true;
    }
    return MOCA_OK;
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
    L.BN = pick_bn(d);
    int rc = make_tmap_bf16(&L.p.tmA, d.A, d.a_rows, d.a_cols, d.a_pitch, kBM);
    if (rc) return rc;
    rc = make_tmap_bf16(&L.p.tmB, d.Wt, d.N, (int64_t)d.ntaps * d.Kp, (int64_t)d.ntaps * d.Kp, L.BN);
    if (rc) return rc;
    GemmParams& p = L.p;
    p.M = d.M; p.N = d.N; p.ntaps = d.ntaps; p.Kp = d.Kp; p.kblocks = d.Kp / 64;
    for (int i = 0; i < 9; ++i) p.tap_shift[i] = d.tap_shift[i];
    p.mode = d.mode; p.g_rows = d.g_rows; p.g_w = d.g_w; p.g_H = d.g_H; p.g_W = d.g_W; p.stride = d.stride;
    p.outf = (float*)d.out; p.outb = (__nv_bfloat16*)d.out; p.ldo = d.ldo;
    p.bias = d.bias; p.rowbias = d.rowbias; p.rb_ld = d.rb_ld; p.rows_per_frame = d.rows_per_frame > 0 ? d.rows_per_frame : 1;
    p.residual = d.residual; p.epi = d.epi;
    p.num_m_tiles = cdiv(d.M, kBM);
    p.num_n_tiles = cdiv(d.N, L.BN);
    const int ntiles = p.num_m_tiles * p.num_n_tiles;
    L.grid = ntiles < ctx.num_sms ? ntiles : ctx.num_sms;
    L.smem = L.BN == 256 ? GemmCfg<256>::SMEM : (L.BN == 160 ? GemmCfg<160>::SMEM : GemmCfg<128>::SMEM);
    L.flops = 2.0 * (double)d.M * d.N * d.ntaps * d.a_cols;
    rc = L.BN == 256 ? set_attr<256>() : (L.BN == 160 ? set_attr<160>() : set_attr<128>());
    if (rc) return rc;
    *out = L;
    return MOCA_OK;
}

int gemm_run(const GemmLaunch& l, cudaStream_t st) {
    if (l.BN == 256) gemm_tc_kernel<256><<<l.grid, kGemmThreads, l.smem, st>>>(l.p);
    else if (l.BN == 160) gemm_tc_kernel<160><<<l.grid, kGemmThreads, l.smem, st>>>(l.p);
    else gemm_tc_kernel<128><<<l.grid, kGemmThreads, l.smem, st>>>(l.p);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

}  // namespace moca
