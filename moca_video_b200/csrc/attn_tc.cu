// tcgen05 flash attention for the spatial self- and cross-attention of the UNet (SURVEY.md table 2.2 rows K6/K7;
// CrossAttention.forward, lvdm/modules/attention.py:92-114): head dim 64, contiguous sequences, bf16 operands,
// fp32 scores / softmax / output accumulation.  sm_100a only.
//
// One CTA = 128 query rows of one (frame, head); two CTAs per SM so that the softmax of one overlaps the MMAs of the other.
//   warp 4 (one lane)  TMA producer: Q once, then K tiles (two stages) and V tiles (one stage) of 128 keys
//   warp 5 (one lane)  MMA issuer:   S = Q K^T      128x128x64 into TMEM columns [0,128)
//                                    PV = P V       128x64x128 into TMEM columns [128,192)   (V is the MN-major B operand)
//   warps 0-3          softmax: thread = query row (TMEM lane).  Two passes over S in TMEM (row max, then exp2 / row sum /
//                      bf16 P written into a 128-byte-swizzled K-major smem tile), running (max, sum) and the 64 fp32 output
//                      accumulators stay in registers; PV of every key tile is read back from TMEM and added.
// Scores never leave the SM; keys >= Lk (sequence tail, cross-attention padding) get P = 0.
#include <stdlib.h>
#include "attn.cuh"
#include "gemm_tc.cuh"

namespace moca {

using namespace ptx;

namespace {

constexpr int kThreads = 192;
constexpr int kTileQ = 128, kTileK = 128, kD = 64;
constexpr int kTileBytes = 128 * kD * 2;                 // 16 KB: 128 rows of 128 B
// smem: Q | K0 | K1 | V | P (two 64-key halves) | barriers
constexpr int kOffQ = 0, kOffK = kTileBytes, kOffV = 3 * kTileBytes, kOffP = 4 * kTileBytes, kOffBar = 6 * kTileBytes;
constexpr size_t kSmem = 6 * kTileBytes + 128 + 1024;    // + barriers + alignment slack

struct AttnTcP {
    CUtensorMap tmQ, tmK, tmV;
    op_t* out; int ldo;
    SeqMap qmap, kvmap;
    int heads, Lq, Lk;
    const int* kv_len; int kv_len_div;
    float scale_log2;
};

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// packed fp32 pairs (sm_100 fma.rn.f32x2 / add.rn.f32x2): the softmax threads are bound by their own instruction stream, and the
// scale-and-subtract and the row sum are one instruction per two scores this way
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
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    op2_t v = floats2op2(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ int64_t first_row(const SeqMap& m, int o) { return (int64_t)(o / m.div) * m.mul_a + (int64_t)(o % m.div) * m.mul_b; }

}  // namespace

__global__ void __launch_bounds__(kThreads, 2) attn_tc_kernel(const __grid_constant__ AttnTcP P) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + kOffBar);
    uint64_t* q_full = bars + 0;
    uint64_t* k_full = bars + 1;     // [2]
    uint64_t* k_empty = bars + 3;    // [2]
    uint64_t* v_full = bars + 5;
    uint64_t* v_empty = bars + 6;
    uint64_t* s_full = bars + 7;
    uint64_t* p_full = bars + 8;
    uint64_t* pv_full = bars + 9;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int o = blockIdx.y / P.heads, h = blockIdx.y - o * P.heads;
    const int q0 = blockIdx.x * kTileQ;
    const int qrow0 = (int)first_row(P.qmap, o), kvrow0 = (int)first_row(P.kvmap, o);

    if (threadIdx.x == 0) {
        prefetch_tmap(&P.tmQ); prefetch_tmap(&P.tmK); prefetch_tmap(&P.tmV);
        mbar_init(q_full, 1);
        mbar_init(&k_full[0], 1); mbar_init(&k_full[1], 1);
        mbar_init(&k_empty[0], 1); mbar_init(&k_empty[1], 1);
        mbar_init(v_full, 1); mbar_init(v_empty, 1);
        mbar_init(s_full, 1);
        mbar_init(p_full, 4);
        mbar_init(pv_full, 1);
        fence_mbar_init();
    }
    if (warp == 5) {
        tmem_alloc(tmem_slot, 256);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                  // PDL (common.cuh): the prologue above overlaps the previous kernel's tail
    pdl_launch_dependents();
    const int Lk = P.kv_len ? P.kv_len[o / P.kv_len_div] : P.Lk;
    const int ntiles = (Lk + kTileK - 1) / kTileK;

    if (warp == 4) {
        if (lane == 0) {
            // ===== TMA producer =====
            mbar_expect_tx(q_full, kTileBytes);
            tma_load_2d(base + kOffQ, &P.tmQ, q_full, h * kD, qrow0 + q0);
            for (int j = 0; j < ntiles; ++j) {
                const int s = j & 1, n = j >> 1;
                mbar_wait(&k_empty[s], (uint32_t)((n & 1) ^ 1));
                mbar_expect_tx(&k_full[s], kTileBytes);
                tma_load_2d(base + kOffK + s * kTileBytes, &P.tmK, &k_full[s], h * kD, kvrow0 + j * kTileK);
                if (j > 0) {
                    // V of the previous tile goes in after its K: the single V stage is released by PV(j-2)
                    const int jv = j - 1;
                    mbar_wait(v_empty, (uint32_t)((jv & 1) ^ 1));
                    mbar_expect_tx(v_full, kTileBytes);
                    tma_load_2d(base + kOffV, &P.tmV, v_full, h * kD, kvrow0 + jv * kTileK);
                }
            }
            const int jv = ntiles - 1;
            mbar_wait(v_empty, (uint32_t)((jv & 1) ^ 1));
            mbar_expect_tx(v_full, kTileBytes);
            tma_load_2d(base + kOffV, &P.tmV, v_full, h * kD, kvrow0 + jv * kTileK);
        }
    } else if (warp == 5) {
        if (lane == 0) {
            // ===== MMA issuer =====
            constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128);
            constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64) | (1u << 16);      // B (= V) is MN-major
            const uint32_t t_s = tmem_base, t_pv = tmem_base + 128;
            const uint64_t dq = umma_desc_sw128(smem_u32(base + kOffQ));
            mbar_wait(q_full, 0);
            auto issue_s = [&](int j) {
                const int s = j & 1, n = j >> 1;
                mbar_wait(&k_full[s], (uint32_t)(n & 1));
                tc_fence_after();
                const uint64_t dk = umma_desc_sw128(smem_u32(base + kOffK + s * kTileBytes));
#pragma unroll
                for (int k = 0; k < kD / 16; ++k) umma_bf16(t_s, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc_s, (uint32_t)(k != 0));
                umma_commit(s_full);
                umma_commit(&k_empty[s]);
            };
            issue_s(0);
            for (int j = 0; j < ntiles; ++j) {
                mbar_wait(p_full, (uint32_t)(j & 1));          // P(j) is in smem and S(j) has been consumed
                mbar_wait(v_full, (uint32_t)(j & 1));
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < kTileK / 16; ++k) {
                    // A: P half (k / 4), 16 keys = 32 bytes along the swizzled row; B: V rows 16k.. (2048 bytes per 16 keys)
                    const uint64_t dp = umma_desc_sw128(smem_u32(base + kOffP + (k >> 2) * kTileBytes)) + (uint64_t)(2 * (k & 3));
                    const uint64_t dv = umma_desc_sw128(smem_u32(base + kOffV + k * 2048));
                    umma_bf16(t_pv, dp, dv, idesc_pv, (uint32_t)(k != 0));
                }
                umma_commit(pv_full);
                umma_commit(v_empty);
                if (j + 1 < ntiles) issue_s(j + 1);
            }
        }
    } else {
        // ===== softmax warps: thread = query row =====
        const int row = warp * 32 + lane;
        const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16);
        uint8_t* prow = base + kOffP + row * 128;               // this row inside either P half
        float m = -INFINITY, l = 0.f;
        float acc[kD];
#pragma unroll
        for (int i = 0; i < kD; ++i) acc[i] = 0.f;
        for (int j = 0; j < ntiles; ++j) {
            const int nvalid = min(kTileK, Lk - j * kTileK);
            mbar_wait(s_full, (uint32_t)(j & 1));
            tc_fence_after();
            // pass 1: row max over the valid keys
            float mx = m;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                if (c * 32 >= nvalid) break;
                uint32_t r[32];
                tmem_ld32(t_row + (uint32_t)(c * 32), r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (c * 32 + i < nvalid) mx = fmaxf(mx, __uint_as_float(r[i]));
            }
            const float ms = mx * P.scale_log2;
            const float alpha = ex2(m * P.scale_log2 - ms);      // m = -inf on the first tile -> 0
            m = mx;
            // pass 2: p = exp2(s * scale - max * scale), row sum, bf16 P into the swizzled smem tile
            const uint64_t sc2 = f2_pack(P.scale_log2, P.scale_log2), nms2 = f2_pack(-ms, -ms);
            uint64_t rs2[2] = {0ull, 0ull};
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t pk[16];
                if (c * 32 < nvalid) {
                    uint32_t r[32];
                    tmem_ld32(t_row + (uint32_t)(c * 32), r);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        float x0, x1;
                        f2_unpack(f2_fma(f2_pack(__uint_as_float(r[i]), __uint_as_float(r[i + 1])), sc2, nms2), x0, x1);
                        float p0 = ex2(x0), p1 = ex2(x1);
                        if (c * 32 + i >= nvalid) p0 = 0.f;
                        if (c * 32 + i + 1 >= nvalid) p1 = 0.f;
                        rs2[(i >> 1) & 1] = f2_add(rs2[(i >> 1) & 1], f2_pack(p0, p1));
                        pk[i >> 1] = pack2(p0, p1);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) pk[i] = 0u;
                }
                // keys c*32 .. c*32+31 = 16-byte chunks (c & 1) * 4 .. + 3 of this row in P half (c >> 1)
                uint8_t* dst = prow + (c >> 1) * kTileBytes;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int chunk = ((c & 1) * 4 + u) ^ (row & 7);
                    *reinterpret_cast<uint4*>(dst + chunk * 16) = make_uint4(pk[4 * u], pk[4 * u + 1], pk[4 * u + 2], pk[4 * u + 3]);
                }
            }
            float rs;
            {
                float a, b;
                f2_unpack(f2_add(rs2[0], rs2[1]), a, b);
                rs = a + b;
            }
            l = l * alpha + rs;
            tc_fence_before();
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
            // rescale the running output while the PV MMA runs, then add PV(j)
#pragma unroll
            for (int i = 0; i < kD; ++i) acc[i] *= alpha;
            mbar_wait(pv_full, (uint32_t)(j & 1));
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t r[32];
                tmem_ld32(t_row + (uint32_t)(128 + c * 32), r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[c * 32 + i] += __uint_as_float(r[i]);
            }
            tc_fence_before();
        }
        const int qr = q0 + row;
        if (qr < P.Lq) {
            const float inv = 1.0f / l;
            op_t* op = P.out + ((int64_t)qrow0 + qr) * P.ldo + h * kD;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                *reinterpret_cast<uint4*>(op + u * 8) =
                    make_uint4(pack2(acc[8 * u] * inv, acc[8 * u + 1] * inv), pack2(acc[8 * u + 2] * inv, acc[8 * u + 3] * inv),
                               pack2(acc[8 * u + 4] * inv, acc[8 * u + 5] * inv), pack2(acc[8 * u + 6] * inv, acc[8 * u + 7] * inv));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc(tmem_base, 256);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Second kernel shape, for sequences of >= 256 queries and keys: one CTA = 256 query rows (tiles A and B) of one (frame, head),
// one CTA per SM, 320 threads:
//   warps 0-3 / 4-7   softmax warpgroups of tile A / B (thread = query row = TMEM lane)
//   warp 8 (one lane) TMA producer: Q_A, Q_B once, then K and V tiles of 128 keys through 3-stage rings
//   warp 9 (one lane) MMA issuer: S_X = Q_X K^T into TMEM columns [128 X, 128 X + 128); O_X (+)= P_X V into [256 + 64 X, ..+64)
// The output accumulator stays in TMEM: a softmax thread reads its 128 scores once, releases S (the next QK^T starts while
// it computes the exponentials), and rescales O only when the running row maximum has grown by more than 2^8 since the
// reference maximum was fixed (warp-uniform decision; exact: the reference maximum cancels in the final division), so the
// per-tile work of a thread is 128 exp2 + one swizzled bf16 row of P.  While warpgroup A is in its exponentials the tensor
// pipe runs the MMAs of tile B and vice versa.
namespace {
constexpr int kThreads2 = 320;
constexpr int kStages2 = 2;
// smem: Q_A | Q_B | K[2] | V[2] | P_A[2 buffers][2 halves] | P_B[2][2] | barriers.  P is double-buffered: the softmax threads write
// P_X(j) while P_X(j-1) V is still in the tensor pipe (with one buffer 18 % of their samples were the wait for that MMA,
// profiles/r02_attn2_ncu.txt); the K / V rings give up one stage each for it.
constexpr int kOff2Q = 0, kOff2K = 2 * kTileBytes, kOff2V = kOff2K + kStages2 * kTileBytes, kOff2P = kOff2V + kStages2 * kTileBytes,
              kOff2Bar = kOff2P + 8 * kTileBytes;
constexpr size_t kSmem2 = (size_t)kOff2Bar + 256 + 1024;
constexpr float kRescaleLog2 = 8.0f;
}  // namespace

__global__ void __launch_bounds__(kThreads2, 1) attn_tc2_kernel(const __grid_constant__ AttnTcP P) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + kOff2Bar);
    uint64_t* q_full = bars + 0;
    uint64_t* k_full = bars + 1;      // [3]
    uint64_t* k_empty = bars + 4;     // [3]
    uint64_t* v_full = bars + 7;      // [3]
    uint64_t* v_empty = bars + 10;    // [3]
    uint64_t* s_full = bars + 13;     // [2] S_X(j) is in TMEM
    uint64_t* s_free = bars + 15;     // [2] S_X(j) has been read into registers
    uint64_t* p_full = bars + 17;     // [2] P_X(j) is in smem (and O_X rescaled if needed)
    uint64_t* o_done = bars + 19;     // [2 tiles][2 buffers] P_X(j) V_j has been accumulated (barrier [x][j & 1]: one completion per two key tiles)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 23);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int o = blockIdx.y / P.heads, h = blockIdx.y - o * P.heads;
    const int q0 = blockIdx.x * 2 * kTileQ;
    const bool hasB = q0 + kTileQ < P.Lq;
    const int qrow0 = (int)first_row(P.qmap, o), kvrow0 = (int)first_row(P.kvmap, o);

    if (threadIdx.x == 0) {
        prefetch_tmap(&P.tmQ); prefetch_tmap(&P.tmK); prefetch_tmap(&P.tmV);
        mbar_init(q_full, 1);
        for (int i = 0; i < kStages2; ++i) {
            mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1);
            mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1);
        }
        for (int x = 0; x < 2; ++x) {
            mbar_init(&s_full[x], 1); mbar_init(&s_free[x], 4);
            mbar_init(&p_full[x], 4); mbar_init(&o_done[2 * x], 1); mbar_init(&o_done[2 * x + 1], 1);
        }
        fence_mbar_init();
    }
    if (warp == 9) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();
    const int Lk = P.kv_len ? P.kv_len[o / P.kv_len_div] : P.Lk;
    const int ntiles = (Lk + kTileK - 1) / kTileK;

    if (warp == 8) {
        if (lane == 0) {
            // ===== TMA producer =====
            mbar_expect_tx(q_full, hasB ? 2 * kTileBytes : kTileBytes);
            tma_load_2d(base + kOff2Q, &P.tmQ, q_full, h * kD, qrow0 + q0);
            if (hasB) tma_load_2d(base + kOff2Q + kTileBytes, &P.tmQ, q_full, h * kD, qrow0 + q0 + kTileQ);
            for (int j = 0; j < ntiles; ++j) {
                const int s = j % kStages2;
                const uint32_t ph = (uint32_t)(((j / kStages2) & 1) ^ 1);
                mbar_wait(&k_empty[s], ph);
                mbar_expect_tx(&k_full[s], kTileBytes);
                tma_load_2d(base + kOff2K + s * kTileBytes, &P.tmK, &k_full[s], h * kD, kvrow0 + j * kTileK);
                mbar_wait(&v_empty[s], ph);
                mbar_expect_tx(&v_full[s], kTileBytes);
                tma_load_2d(base + kOff2V + s * kTileBytes, &P.tmV, &v_full[s], h * kD, kvrow0 + j * kTileK);
            }
        }
    } else if (warp == 9) {
        if (lane == 0) {
            // ===== MMA issuer =====
            constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128);
            constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64) | (1u << 16);      // B (= V) is MN-major
            const int nx = hasB ? 2 : 1;
            mbar_wait(q_full, 0);
            auto issue_s = [&](int x, int j) {
                const uint64_t dq = umma_desc_sw128(smem_u32(base + kOff2Q + x * kTileBytes));
                const uint64_t dk = umma_desc_sw128(smem_u32(base + kOff2K + (j % kStages2) * kTileBytes));
#pragma unroll
                for (int k = 0; k < kD / 16; ++k)
                    umma_bf16(tmem_base + (uint32_t)(x * 128), dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc_s, (uint32_t)(k != 0));
                umma_commit(&s_full[x]);
            };
            mbar_wait(&k_full[0], 0);
            tc_fence_after();
            for (int x = 0; x < nx; ++x) issue_s(x, 0);
            umma_commit(&k_empty[0]);
            for (int j = 0; j < ntiles; ++j) {
                const uint32_t ph = (uint32_t)(j & 1);
                const bool more = j + 1 < ntiles;
                if (more) { mbar_wait(&k_full[(j + 1) % kStages2], (uint32_t)(((j + 1) / kStages2) & 1)); }
                if (more) {
                    for (int x = 0; x < nx; ++x) {
                        mbar_wait(&s_free[x], ph);            // S_X(j) is in registers: QK^T of the next key tile may overwrite it
                        tc_fence_after();
                        issue_s(x, j + 1);
                    }
                    umma_commit(&k_empty[(j + 1) % kStages2]);
                }
                mbar_wait(&v_full[j % kStages2], (uint32_t)((j / kStages2) & 1));
                for (int x = 0; x < nx; ++x) {
                    mbar_wait(&p_full[x], ph);
                    tc_fence_after();
                    const uint32_t t_o = tmem_base + (uint32_t)(256 + x * 64);
                    const uint32_t pb = smem_u32(base + kOff2P + (x * 2 + (j & 1)) * 2 * kTileBytes);
                    const uint32_t vb = smem_u32(base + kOff2V + (j % kStages2) * kTileBytes);
#pragma unroll
                    for (int k = 0; k < kTileK / 16; ++k) {
                        const uint64_t dp = umma_desc_sw128(pb + (k >> 2) * kTileBytes) + (uint64_t)(2 * (k & 3));
                        const uint64_t dv = umma_desc_sw128(vb + k * 2048);
                        umma_bf16(t_o, dp, dv, idesc_pv, (uint32_t)((j | k) != 0));
                    }
                    umma_commit(&o_done[2 * x + (j & 1)]);
                }
                umma_commit(&v_empty[j % kStages2]);
            }
        }
    } else {
        // ===== softmax warpgroups: x = tile (A / B), thread = query row =====
        const int x = warp >> 2, wq = warp & 3;
        if (x == 0 || hasB) {
            const int row = wq * 32 + lane;
            const uint32_t t_lane = tmem_base + ((uint32_t)(wq * 32) << 16);
            const uint32_t t_s = t_lane + (uint32_t)(x * 128), t_o = t_lane + (uint32_t)(256 + x * 64);
            uint8_t* prow0 = base + kOff2P + x * 4 * kTileBytes + row * 128;       // this row inside either 64-key half of P_X, buffer 0
            float m_ref = -INFINITY, l = 0.f;
            for (int j = 0; j < ntiles; ++j) {
                const uint32_t ph = (uint32_t)(j & 1);
                const int nvalid = min(kTileK, Lk - j * kTileK);
                mbar_wait(&s_full[x], ph);
                tc_fence_after();
                uint32_t r[128];
                tmem_ld32(t_s, r); tmem_ld32(t_s + 32, r + 32); tmem_ld32(t_s + 64, r + 64); tmem_ld32(t_s + 96, r + 96);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_free[x]);
                if (nvalid < kTileK) {
#pragma unroll
                    for (int i = 0; i < 128; ++i)
                        if (i >= nvalid) r[i] = 0xff800000u;             // -inf: exp2 gives exactly 0
                }
                float mxp[8];                                             // 8 independent chains instead of one of 128
#pragma unroll
                for (int i = 0; i < 8; ++i) mxp[i] = __uint_as_float(r[i]);
#pragma unroll
                for (int i = 8; i < 128; ++i) mxp[i & 7] = fmaxf(mxp[i & 7], __uint_as_float(r[i]));
                const float mx = fmaxf(fmaxf(fmaxf(mxp[0], mxp[1]), fmaxf(mxp[2], mxp[3])), fmaxf(fmaxf(mxp[4], mxp[5]), fmaxf(mxp[6], mxp[7])));
                // the reference maximum moves only when some row of the warp outgrew it by 2^8 (first tile: always)
                const bool grow = (mx - m_ref) * P.scale_log2 > kRescaleLog2;
                uint8_t* prow = prow0 + (j & 1) * 2 * kTileBytes;
                if (j >= 2) {
                    mbar_wait(&o_done[2 * x + (j & 1)], (uint32_t)(((j >> 1) - 1) & 1));      // P_X(j-2) V is done: this P buffer is ours
                    tc_fence_after();
                }
                if (__any_sync(0xffffffffu, grow)) {
                    const float m_new = fmaxf(m_ref, mx);
                    const float alpha = ex2((m_ref - m_new) * P.scale_log2);       // first tile: exp2(-inf) = 0, l = 0 anyway
                    m_ref = m_new;
                    l *= alpha;
                    if (j > 0) {
                        mbar_wait(&o_done[2 * x + ((j - 1) & 1)], (uint32_t)(((j - 1) >> 1) & 1));   // P_X(j-1) V is done: O_X is ours
                        tc_fence_after();
#pragma unroll 1
                        for (int c = 0; c < 8; ++c) {                      // rare path: 8 columns at a time keeps the 128 scores in registers
                            uint32_t ov[8];
                            tmem_ld8(t_o + (uint32_t)(c * 8), ov);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 8; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
                            tmem_st8(t_o + (uint32_t)(c * 8), ov);
                        }
                        tmem_st_wait();
                    }
                }
                const float ms = m_ref * P.scale_log2;
                const uint64_t sc2 = f2_pack(P.scale_log2, P.scale_log2), nms2 = f2_pack(-ms, -ms);
                uint64_t rsp[4] = {0ull, 0ull, 0ull, 0ull};              // four chains of (even, odd) partial row sums
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t pk[16];
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        float x0, x1;
                        f2_unpack(f2_fma(f2_pack(__uint_as_float(r[c * 32 + i]), __uint_as_float(r[c * 32 + i + 1])), sc2, nms2), x0, x1);
                        const float p0 = ex2(x0), p1 = ex2(x1);
                        rsp[(i >> 1) & 3] = f2_add(rsp[(i >> 1) & 3], f2_pack(p0, p1));
                        pk[i >> 1] = pack2(p0, p1);
                    }
                    uint8_t* dst = prow + (c >> 1) * kTileBytes;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int chunk = ((c & 1) * 4 + u) ^ (row & 7);
                        *reinterpret_cast<uint4*>(dst + chunk * 16) = make_uint4(pk[4 * u], pk[4 * u + 1], pk[4 * u + 2], pk[4 * u + 3]);
                    }
                }
                {
                    float a, b;
                    f2_unpack(f2_add(f2_add(rsp[0], rsp[1]), f2_add(rsp[2], rsp[3])), a, b);
                    l += a + b;
                }
                tc_fence_before();
                fence_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[x]);
            }
            // epilogue: O_X / l -> bf16
            mbar_wait(&o_done[2 * x + ((ntiles - 1) & 1)], (uint32_t)(((ntiles - 1) >> 1) & 1));
            tc_fence_after();
            const int qr = q0 + x * kTileQ + row;
            const float inv = 1.0f / l;
            op_t* op = P.out + ((int64_t)qrow0 + qr) * P.ldo + h * kD;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t ov[32];
                tmem_ld32(t_o + (uint32_t)(c * 32), ov);
                tmem_ld_wait();
                if (qr < P.Lq) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        float f[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(ov[8 * u + e]) * inv;
                        *reinterpret_cast<uint4*>(op + c * 32 + u * 8) =
                            make_uint4(pack2(f[0], f[1]), pack2(f[2], f[3]), pack2(f[4], f[5]), pack2(f[6], f[7]));
                    }
                }
            }
            tc_fence_before();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------------------------
struct AttnTcLaunchImpl {
    AttnTcP p;
    dim3 grid;
    int version;      // 1: 128 query rows per CTA, two CTAs per SM; 2: 256 query rows per CTA, accumulator in TMEM
};

bool attention_tc_supported(const AttnArgs& a) {
    return a.qmap.step == 1 && a.kvmap.step == 1 && a.Lq >= 64 && a.q_rows > 0 && a.kv_rows > 0 && a.ldq % 8 == 0 && a.ldkv % 8 == 0 &&
           a.ldo % 8 == 0 && ((((uintptr_t)a.q | (uintptr_t)a.k | (uintptr_t)a.v | (uintptr_t)a.out) & 15) == 0) &&
           (int64_t)a.n_outer * a.heads <= 65535;
}

int attention_tc_plan(const AttnArgs& a, AttnTcLaunch* out) {
    MOCA_REQUIRE(a.q && a.k && a.v && a.out, MOCA_E_ARG, "attention_tc: NULL tensor");
    MOCA_REQUIRE(attention_tc_supported(a), MOCA_E_SHAPE, "attention_tc: unsupported geometry (contiguous sequences, Lq >= 64, 16-byte aligned)");
    static_assert(sizeof(AttnTcLaunchImpl) <= sizeof(AttnTcLaunch::storage), "AttnTcLaunch storage too small");
    AttnTcLaunchImpl L;
    memset(&L, 0, sizeof(L));
    const int width = a.heads * kD;
    int rc = make_tmap_bf16(&L.p.tmQ, a.q, a.q_rows, width, a.ldq, 128);
    if (rc) return rc;
    rc = make_tmap_bf16(&L.p.tmK, a.k, a.kv_rows, width, a.ldkv, 128);
    if (rc) return rc;
    rc = make_tmap_bf16(&L.p.tmV, a.v, a.kv_rows, width, a.ldkv, 128);
    if (rc) return rc;
    L.p.out = a.out; L.p.ldo = a.ldo; L.p.qmap = a.qmap; L.p.kvmap = a.kvmap;
    L.p.heads = a.heads; L.p.Lq = a.Lq; L.p.Lk = a.Lk;
    L.p.kv_len = a.kv_len; L.p.kv_len_div = a.kv_len_div > 0 ? a.kv_len_div : 1;
    L.p.scale_log2 = a.scale * 1.4426950408889634f;
    static const int forced = [] { const char* e = getenv("MOCA_ATTN_TC"); return e ? atoi(e) : 0; }();      // 1 / 2 pins the kernel shape
    L.version = forced == 1 || forced == 2 ? forced : (a.Lq >= 256 && a.Lk >= 256 ? 2 : 1);
    L.grid = dim3(cdiv(a.Lq, L.version == 2 ? 2 * kTileQ : kTileQ), a.n_outer * a.heads);
    // per-device attribute, set at plan time (once per plan, not per launch)
    MOCA_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    MOCA_CHECK_CUDA(cudaFuncSetAttribute(attn_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem2));
    memcpy(out->storage, &L, sizeof(L));
    return MOCA_OK;
}

int attention_tc_run(const AttnTcLaunch& l, cudaStream_t st) {
    const AttnTcLaunchImpl* L = reinterpret_cast<const AttnTcLaunchImpl*>(l.storage);
    if (L->version == 2) launch_pdl_tc(attn_tc2_kernel, L->grid, dim3(kThreads2), kSmem2, st, L->p);
    else launch_pdl_tc(attn_tc_kernel, L->grid, dim3(kThreads), kSmem, st, L->p);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

}  // namespace moca
