// See attn.cuh.  Head dim 64, bf16 operands, fp32 softmax/accumulators.  Scores never touch HBM (the reference's
// einsum path materialises a [80, 2560, 2560] fp32 tensor per call, attention.py:92-114).
// Tensor-core path of this first version: warp-level mma.sync m16n8k16 (HMMA) fed by ldmatrix from XOR-swizzled
// shared-memory tiles filled with cp.async; 6 % of the UNet FLOPs live here (SURVEY table 2.2 K6-K8).
#include "attn.cuh"

namespace moca {

namespace {

constexpr int D = 64;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// element offset inside a [rows][64] bf16 tile whose 16-byte chunks are XOR-swizzled by (row & 7)
__device__ __forceinline__ int swz(int row, int col) { return row * D + ((((col >> 3) ^ (row & 7)) << 3) | (col & 7)); }

__device__ __forceinline__ void cp_async16(void* dst, const void* src, bool valid) {
    const uint32_t d = smem_addr(dst);
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32." MOCA_MMA_OP_STR ".f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    op2_t v = floats2op2(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ int64_t first_row(const SeqMap& m, int o) { return (int64_t)(o / m.div) * m.mul_a + (int64_t)(o % m.div) * m.mul_b; }

struct AttnP {
    const op_t* q; int ldq; SeqMap qmap;
    const op_t* k; const op_t* v; int ldkv; SeqMap kvmap;
    op_t* out; int ldo;
    int n_outer, heads, Lq, Lk;
    const int* kv_len; int kv_len_div;
    float scale_log2;
};

// loads ROWS x 64 bf16 (rows r0.. of a sequence) into a swizzled tile; rows >= L are zero-filled
template <int ROWS, int THREADS>
__device__ __forceinline__ void load_tile(op_t* s, const op_t* g, int64_t row0, int step, int ld, int r0, int L,
                                          int tid) {
#pragma unroll
    for (int i = tid; i < ROWS * 8; i += THREADS) {
        const int r = i >> 3, ch = i & 7;
        const bool ok = (r0 + r) < L;
        const op_t* src = g + (row0 + (int64_t)(ok ? r0 + r : 0) * step) * ld + ch * 8;
        cp_async16(s + r * D + ((ch ^ (r & 7)) << 3), src, ok);
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) attn_flash_kernel(const __grid_constant__ AttnP P) {
    __shared__ __align__(128) op_t sQ[64 * D];
    __shared__ __align__(128) op_t sK[2][64 * D];
    __shared__ __align__(128) op_t sV[2][64 * D];
    pdl_wait();
    pdl_launch_dependents();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int o = blockIdx.y / P.heads, h = blockIdx.y - o * P.heads;
    const int q0 = blockIdx.x * 64;
    const int Lk = P.kv_len ? P.kv_len[o / P.kv_len_div] : P.Lk;
    const int64_t qrow0 = first_row(P.qmap, o), kvrow0 = first_row(P.kvmap, o);
    const op_t* qg = P.q + h * D;
    const op_t* kg = P.k + h * D;
    const op_t* vg = P.v + h * D;
    const int ntiles = (Lk + 63) / 64;

    load_tile<64, 128>(sQ, qg, qrow0, P.qmap.step, P.ldq, q0, P.Lq, tid);
    load_tile<64, 128>(sK[0], kg, kvrow0, P.kvmap.step, P.ldkv, 0, Lk, tid);
    load_tile<64, 128>(sV[0], vg, kvrow0, P.kvmap.step, P.ldkv, 0, Lk, tid);
    cp_async_commit();

    uint32_t qf[4][4];
    float oacc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) oacc[i][e] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

    for (int j = 0; j < ntiles; ++j) {
        const int buf = j & 1;
        if (j + 1 < ntiles) {
            load_tile<64, 128>(sK[buf ^ 1], kg, kvrow0, P.kvmap.step, P.ldkv, (j + 1) * 64, Lk, tid);
            load_tile<64, 128>(sV[buf ^ 1], vg, kvrow0, P.kvmap.step, P.ldkv, (j + 1) * 64, Lk, tid);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (j == 0) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const int r = warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
                const int c = kk * 16 + 8 * (lane >> 4);
                ldsm_x4(smem_addr(sQ + swz(r, c)), qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3]);
            }
        }
        float s[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int e = 0; e < 4; ++e) s[i][e] = 0.f;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                const int mi = lane >> 3;
                const int r = np * 16 + (lane & 7) + 8 * (mi >> 1);      // key
                const int c = kk * 16 + 8 * (mi & 1);                    // d
                uint32_t b0, b1, b2, b3;
                ldsm_x4(smem_addr(sK[buf] + swz(r, c)), b0, b1, b2, b3);
                mma_bf16(s[2 * np], qf[kk], b0, b1);
                mma_bf16(s[2 * np + 1], qf[kk], b2, b3);
            }
        }
        // mask + online softmax (rows g and g+8 of this warp's 16 query rows)
        const int kbase = j * 64;
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int key = kbase + i * 8 + 2 * t;
            if (key >= Lk) { s[i][0] = -INFINITY; s[i][2] = -INFINITY; }
            if (key + 1 >= Lk) { s[i][1] = -INFINITY; s[i][3] = -INFINITY; }
            mx0 = fmaxf(mx0, fmaxf(s[i][0], s[i][1]));
            mx1 = fmaxf(mx1, fmaxf(s[i][2], s[i][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
        const float c0 = exp2f((m0 - mn0) * P.scale_log2), c1 = exp2f((m1 - mn1) * P.scale_log2);
        m0 = mn0; m1 = mn1;
        const float ms0 = mn0 * P.scale_log2, ms1 = mn1 * P.scale_log2;
        float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            s[i][0] = exp2f(s[i][0] * P.scale_log2 - ms0); s[i][1] = exp2f(s[i][1] * P.scale_log2 - ms0);
            s[i][2] = exp2f(s[i][2] * P.scale_log2 - ms1); s[i][3] = exp2f(s[i][3] * P.scale_log2 - ms1);
            rs0 += s[i][0] + s[i][1];
            rs1 += s[i][2] + s[i][3];
        }
        l0 = l0 * c0 + rs0;
        l1 = l1 * c1 + rs1;
#pragma unroll
        for (int i = 0; i < 8; ++i) { oacc[i][0] *= c0; oacc[i][1] *= c0; oacc[i][2] *= c1; oacc[i][3] *= c1; }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            uint32_t pa[4];
            pa[0] = pack2(s[2 * kk][0], s[2 * kk][1]);
            pa[1] = pack2(s[2 * kk][2], s[2 * kk][3]);
            pa[2] = pack2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            pa[3] = pack2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
            for (int dp = 0; dp < 4; ++dp) {
                const int mi = lane >> 3;
                const int r = kk * 16 + (lane & 7) + 8 * (mi & 1);       // key
                const int c = dp * 16 + 8 * (mi >> 1);                   // d
                uint32_t b0, b1, b2, b3;
                ldsm_x4_t(smem_addr(sV[buf] + swz(r, c)), b0, b1, b2, b3);
                mma_bf16(oacc[2 * dp], pa, b0, b1);
                mma_bf16(oacc[2 * dp + 1], pa, b2, b3);
            }
        }
        __syncthreads();
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
    const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
    op_t* og = P.out + h * D + 2 * t;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (r0 < P.Lq)
            *reinterpret_cast<uint32_t*>(og + (qrow0 + (int64_t)r0 * P.qmap.step) * P.ldo + i * 8) = pack2(oacc[i][0] * i0, oacc[i][1] * i0);
        if (r1 < P.Lq)
            *reinterpret_cast<uint32_t*>(og + (qrow0 + (int64_t)r1 * P.qmap.step) * P.ldo + i * 8) = pack2(oacc[i][2] * i1, oacc[i][3] * i1);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// One warp per (outer, head) sequence of L <= 16 elements (temporal attention: sequence = the T frames of one
// spatial position, rows `step` = H*W apart in the channels-last token matrix — no rearrange copy is ever made).
__global__ void __launch_bounds__(128) attn_short_kernel(const __grid_constant__ AttnP P) {
    __shared__ __align__(128) op_t sm[4][3][16 * D];
    pdl_wait();
    pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int item = blockIdx.x * 4 + warp;
    if (item >= P.n_outer * P.heads) return;
    const int o = item / P.heads, h = item - o * P.heads;
    const int L = P.Lq;
    const int64_t qrow0 = first_row(P.qmap, o), kvrow0 = first_row(P.kvmap, o);
    op_t* sQ = sm[warp][0];
    op_t* sK = sm[warp][1];
    op_t* sV = sm[warp][2];
    load_tile<16, 32>(sQ, P.q + h * D, qrow0, P.qmap.step, P.ldq, 0, L, lane);
    load_tile<16, 32>(sK, P.k + h * D, kvrow0, P.kvmap.step, P.ldkv, 0, L, lane);
    load_tile<16, 32>(sV, P.v + h * D, kvrow0, P.kvmap.step, P.ldkv, 0, L, lane);
    cp_async_commit();
    cp_async_wait<0>();
    __syncwarp();
    float s[2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) s[i][e] = 0.f;
    const int mi = lane >> 3;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
        uint32_t a[4], b0, b1, b2, b3;
        ldsm_x4(smem_addr(sQ + swz((lane & 7) + 8 * (mi & 1), kk * 16 + 8 * (mi >> 1))), a[0], a[1], a[2], a[3]);
        ldsm_x4(smem_addr(sK + swz((lane & 7) + 8 * (mi >> 1), kk * 16 + 8 * (mi & 1))), b0, b1, b2, b3);
        mma_bf16(s[0], a, b0, b1);
        mma_bf16(s[1], a, b2, b3);
    }
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int key = i * 8 + 2 * t;
        if (key >= L) { s[i][0] = -INFINITY; s[i][2] = -INFINITY; }
        if (key + 1 >= L) { s[i][1] = -INFINITY; s[i][3] = -INFINITY; }
        mx0 = fmaxf(mx0, fmaxf(s[i][0], s[i][1]));
        mx1 = fmaxf(mx1, fmaxf(s[i][2], s[i][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        s[i][0] = exp2f((s[i][0] - mx0) * P.scale_log2); s[i][1] = exp2f((s[i][1] - mx0) * P.scale_log2);
        s[i][2] = exp2f((s[i][2] - mx1) * P.scale_log2); s[i][3] = exp2f((s[i][3] - mx1) * P.scale_log2);
        l0 += s[i][0] + s[i][1];
        l1 += s[i][2] + s[i][3];
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    uint32_t pa[4] = {pack2(s[0][0], s[0][1]), pack2(s[0][2], s[0][3]), pack2(s[1][0], s[1][1]), pack2(s[1][2], s[1][3])};
    float oacc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) oacc[i][e] = 0.f;
#pragma unroll
    for (int dp = 0; dp < 4; ++dp) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(smem_addr(sV + swz((lane & 7) + 8 * (mi & 1), dp * 16 + 8 * (mi >> 1))), b0, b1, b2, b3);
        mma_bf16(oacc[2 * dp], pa, b0, b1);
        mma_bf16(oacc[2 * dp + 1], pa, b2, b3);
    }
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
    op_t* og = P.out + h * D + 2 * t;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (g < L) *reinterpret_cast<uint32_t*>(og + (qrow0 + (int64_t)g * P.qmap.step) * P.ldo + i * 8) = pack2(oacc[i][0] * i0, oacc[i][1] * i0);
        if (g + 8 < L)
            *reinterpret_cast<uint32_t*>(og + (qrow0 + (int64_t)(g + 8) * P.qmap.step) * P.ldo + i * 8) = pack2(oacc[i][2] * i1, oacc[i][3] * i1);
    }
}

static int fill(const AttnArgs& a, AttnP& P) {
    MOCA_REQUIRE(a.q && a.k && a.v && a.out, MOCA_E_ARG, "attention: NULL tensor");
    MOCA_REQUIRE(a.n_outer > 0 && a.heads > 0 && a.Lq > 0 && a.Lk > 0, MOCA_E_SHAPE, "attention: bad sizes outer %d heads %d Lq %d Lk %d",
                 a.n_outer, a.heads, a.Lq, a.Lk);
    MOCA_REQUIRE(a.ldq % 8 == 0 && a.ldkv % 8 == 0 && a.ldo % 2 == 0, MOCA_E_ALIGN, "attention: row pitches must be multiples of 8 bf16");
    MOCA_REQUIRE((((uintptr_t)a.q | (uintptr_t)a.k | (uintptr_t)a.v) & 15) == 0 && ((uintptr_t)a.out & 3) == 0, MOCA_E_ALIGN,
                 "attention: pointers must be 16-byte aligned");
    P.q = a.q; P.ldq = a.ldq; P.qmap = a.qmap; P.k = a.k; P.v = a.v; P.ldkv = a.ldkv; P.kvmap = a.kvmap;
    P.out = a.out; P.ldo = a.ldo; P.n_outer = a.n_outer; P.heads = a.heads; P.Lq = a.Lq; P.Lk = a.Lk;
    P.kv_len = a.kv_len; P.kv_len_div = a.kv_len_div > 0 ? a.kv_len_div : 1;
    P.scale_log2 = a.scale * 1.4426950408889634f;
    return MOCA_OK;
}

int attention_flash(const AttnArgs& a, cudaStream_t st) {
    AttnP P;
    int rc = fill(a, P);
    if (rc) return rc;
    MOCA_REQUIRE((int64_t)a.n_outer * a.heads <= 65535, MOCA_E_SHAPE, "attention_flash: outer*heads = %lld exceeds grid.y",
                 (long long)a.n_outer * a.heads);
    dim3 grid(cdiv(a.Lq, 64), a.n_outer * a.heads);
    launch_pdl(attn_flash_kernel, grid, dim3(128), 0, st, P);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

int attention_short(const AttnArgs& a, cudaStream_t st) {
    AttnP P;
    int rc = fill(a, P);
    if (rc) return rc;
    MOCA_REQUIRE(a.Lq == a.Lk && a.Lq <= 16 && a.kv_len == nullptr, MOCA_E_SHAPE, "attention_short: needs Lq == Lk <= 16 (got %d, %d)", a.Lq, a.Lk);
    const int64_t items = (int64_t)a.n_outer * a.heads;
    launch_pdl(attn_short_kernel, dim3(cdiv(items, 4)), dim3(128), 0, st, P);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

}  // namespace moca
