// See norm.cuh.  All kernels here are bandwidth-bound: 16-byte vector accesses, channels-last (tokens x C) layout.
#include "norm.cuh"

namespace moca {

// ---------------------------------------------------------------------------------------------------------------
__global__ void pack_weight_kernel(const float* __restrict__ src, int N, int K, int taps, int Kp, __nv_bfloat16* __restrict__ dst,
                                   int64_t pitch, int row0, int geglu_bn) {
    const int64_t total = (int64_t)N * taps * Kp;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % Kp);
        const int tap = (int)((i / Kp) % taps);
        const int n = (int)(i / ((int64_t)Kp * taps));
        float v = 0.0f;
        if (k < K) v = src[((int64_t)n * K + k) * taps + tap];
        int r = n;
        if (geglu_bn > 0) {
            // rows [0, N/2) are values, [N/2, N) gates (x, gate = proj(x).chunk(2), attention.py:382)
            const int half = N / 2, hb = geglu_bn / 2;
            const int j = n < half ? n : n - half;
            r = (j / hb) * geglu_bn + (n < half ? 0 : hb) + (j % hb);
        }
        dst[(int64_t)(row0 + r) * pitch + (int64_t)tap * Kp + k] = __float2bfloat16_rn(v);
    }
}

int pack_weight(const float* src, int N, int K, int taps, int Kp, __nv_bfloat16* dst, int64_t dst_pitch, int row0, int geglu_bn,
                cudaStream_t st) {
    const int64_t total = (int64_t)N * taps * Kp;
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    pack_weight_kernel<<<blocks, 256, 0, st>>>(src, N, K, taps, Kp, dst, dst_pitch, row0, geglu_bn);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// GroupNorm statistics
// Block = TPR x RP threads: thread (x, y) owns the float4 channel vectors x, x+TPR, .. of rows y, y+RP, .. of its 32-row
// chunk, so its channels (and groups) are fixed and partial sums stay in registers.
constexpr int kStatSlots = 3;      // float4 slots per thread: supports C <= 256*3*4 = 3072
constexpr int kStatChunk = 32;     // rows per block; fixed so that statistics do not depend on the batch size

__global__ void __launch_bounds__(256) gn_stats_kernel(const float* __restrict__ x1, int C1, const float* __restrict__ x2,
                                                       int C2, int rows_per_stat, int Cg, int tpr, int rp,
                                                       double* __restrict__ acc) {
    // per-(row lane, channel) partial sums; a thread owns its channels so plain stores suffice
    __shared__ float s_sum[256 * kStatSlots * 4];
    __shared__ float s_sq[256 * kStatSlots * 4];
    const int stat = blockIdx.y;
    const int r0 = blockIdx.x * kStatChunk;
    const int r1 = min(rows_per_stat, r0 + kStatChunk);
    const int C = C1 + C2;
    const int nvec = C >> 2;
    const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
    float s[kStatSlots][4], ss[kStatSlots][4];
#pragma unroll
    for (int u = 0; u < kStatSlots; ++u)
#pragma unroll
        for (int e = 0; e < 4; ++e) s[u][e] = ss[u][e] = 0.0f;
    if (ty < rp) {
#pragma unroll 4
        for (int r = r0 + ty; r < r1; r += rp) {
            const size_t row = (size_t)stat * rows_per_stat + r;
#pragma unroll
            for (int u = 0; u < kStatSlots; ++u) {
                const int v = tx + u * tpr;
                if (v < nvec) {
                    const int c = v << 2;
                    const float4 t = c < C1 ? __ldg(reinterpret_cast<const float4*>(x1 + row * C1 + c))
                                            : __ldg(reinterpret_cast<const float4*>(x2 + row * C2 + (c - C1)));
                    s[u][0] += t.x; ss[u][0] += t.x * t.x;
                    s[u][1] += t.y; ss[u][1] += t.y * t.y;
                    s[u][2] += t.z; ss[u][2] += t.z * t.z;
                    s[u][3] += t.w; ss[u][3] += t.w * t.w;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kStatSlots; ++u) {
            const int v = tx + u * tpr;
            if (v < nvec) {
                *reinterpret_cast<float4*>(&s_sum[ty * C + (v << 2)]) = make_float4(s[u][0], s[u][1], s[u][2], s[u][3]);
                *reinterpret_cast<float4*>(&s_sq[ty * C + (v << 2)]) = make_float4(ss[u][0], ss[u][1], ss[u][2], ss[u][3]);
            }
        }
    }
    __syncthreads();
    // 64 threads: (group, sum|sumsq) -> fixed-order double sum over the group's channels and row lanes, one atomic each
    if (threadIdx.x < 64) {
        const int g = threadIdx.x >> 1;
        const float* src = (threadIdx.x & 1) ? s_sq : s_sum;
        double a = 0.0;
        for (int y = 0; y < rp; ++y)
            for (int c = g * Cg; c < (g + 1) * Cg; ++c) a += (double)src[y * C + c];
        atomicAdd(&acc[(size_t)stat * 64 + threadIdx.x], a);
    }
}

static void stat_geometry(int nvec, int& slots, int& tpr, int& rp) {
    slots = (nvec + 255) / 256;
    tpr = (((nvec + slots - 1) / slots) + 31) / 32 * 32;
    if (tpr > 256) tpr = 256;
    rp = 256 / tpr;
    if (rp < 1) rp = 1;
}

int gn_stats(const float* x1, int C1, const float* x2, int C2, int64_t rows, int rows_per_stat, int groups, double* acc,
             cudaStream_t st) {
    const int C = C1 + C2;
    MOCA_REQUIRE(groups == 32 && C % 32 == 0 && C % 4 == 0 && C1 % 4 == 0 && C <= 256 * kStatSlots * 4, MOCA_E_SHAPE,
                 "gn_stats: unsupported channels C1=%d C2=%d groups=%d", C1, C2, groups);
    MOCA_REQUIRE(rows % rows_per_stat == 0, MOCA_E_SHAPE, "gn_stats: rows %lld not a multiple of %d", (long long)rows, rows_per_stat);
    const int nstat = (int)(rows / rows_per_stat);
    int slots, tpr, rp;
    stat_geometry(C / 4, slots, tpr, rp);
    dim3 grid(cdiv(rows_per_stat, kStatChunk), nstat);
    gn_stats_kernel<<<grid, 256, 0, st>>>(x1, C1, x2, C2, rows_per_stat, C / 32, tpr, rp, acc);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// GroupNorm apply (+SiLU) with layout scatter.  Block = TPR x RP threads as above; each block walks kApplyRows destination
// rows.  A thread's channels are fixed: gamma / beta / group ids live in registers; mean / rstd of the (at most two)
// statistics groups a block touches are derived once per block into shared memory.
struct ApplyP {
    const float* x1; int C1;
    const float* x2; int C2;
    int rows_per_stat; int Cg;
    const double* acc; const float* gamma; const float* beta;
    float eps; int silu; int mode;
    int NF, H, W;        // source geometry
    int T, HW;
    int dst_rows;        // rows of the destination layout
    int nstat;
    int tpr, rp;
    __nv_bfloat16* dst; __nv_bfloat16* raw;
};

constexpr int kApplyRows = 16;

__device__ __forceinline__ float silu_f(float v) { return __fdividef(v, 1.0f + __expf(-v)); }

// destination row -> source row (or -1 for a zero border / padding row)
__device__ __forceinline__ int apply_src_row(const ApplyP& P, int drow) {
    if (P.mode == DST_LINEAR) return drow;
    if (P.mode == DST_TIME_PAD) {
        const int per = P.HW * (P.T + 2);
        const int b = drow / per, r = drow - b * per;
        const int tp = r / P.HW, hw = r - tp * P.HW;
        return (tp < 1 || tp > P.T) ? -1 : (b * P.T + (tp - 1)) * P.HW + hw;
    }
    const int up = P.mode == DST_SPATIAL_PAD_UP2 ? 1 : 0;
    const int Hd = P.H << up, Wd = P.W << up, Wp = Wd + 2, Hp = Hd + 2;
    const int f = drow / (Wp * Hp), r = drow - f * (Wp * Hp);
    const int hp = r / Wp, wp = r - hp * Wp;
    if (hp < 1 || hp > Hd || wp < 1 || wp > Wd) return -1;
    return (f * P.H + ((hp - 1) >> up)) * P.W + ((wp - 1) >> up);
}

__global__ void __launch_bounds__(256) gn_apply_kernel(const __grid_constant__ ApplyP P) {
    __shared__ float2 s_mr[kApplyRows + 1][32];   // (mean, rstd) of statistics groups s_first, s_first + 1, ..
    __shared__ int s_first;
    const int C = P.C1 + P.C2;
    const int nvec = C >> 2;
    const int tx = threadIdx.x % P.tpr, ty = threadIdx.x / P.tpr;
    const int d0 = blockIdx.x * kApplyRows;
    const int d1 = min(P.dst_rows, d0 + kApplyRows);
    if (P.rows_per_stat > 0) {
        if (threadIdx.x == 0) {
            int first = -1;
            for (int d = d0; d < d1 && first < 0; ++d) {
                const int sr = apply_src_row(P, d);
                if (sr >= 0) first = sr / P.rows_per_stat;
            }
            s_first = first;
        }
        __syncthreads();
        const int first = s_first;
        if (first >= 0) {
            // a block's kApplyRows destination rows map to at most kApplyRows consecutive statistics groups
            const int nwhich = min(kApplyRows + 1, (kApplyRows + P.rows_per_stat - 1) / P.rows_per_stat + 1);
            for (int i = threadIdx.x; i < nwhich * 32; i += blockDim.x) {
                const int which = i >> 5, g = i & 31;
                const int stat = first + which;
                if (stat < P.nstat) {
                    const double n = (double)P.rows_per_stat * P.Cg;
                    const double sum = P.acc[(size_t)stat * 64 + 2 * g], sq = P.acc[(size_t)stat * 64 + 2 * g + 1];
                    const double m = sum / n;
                    double var = sq / n - m * m;
                    var = var < 0.0 ? 0.0 : var;
                    s_mr[which][g] = make_float2((float)m, rsqrtf((float)var + P.eps));
                }
            }
        }
        __syncthreads();
    }
    if (ty >= P.rp) return;
    float4 ga[kStatSlots], be[kStatSlots];
    int grp[kStatSlots][4];
#pragma unroll
    for (int u = 0; u < kStatSlots; ++u) {
        const int v = tx + u * P.tpr;
        if (v < nvec && P.rows_per_stat > 0) {
            ga[u] = *reinterpret_cast<const float4*>(P.gamma + (v << 2));
            be[u] = *reinterpret_cast<const float4*>(P.beta + (v << 2));
#pragma unroll
            for (int e = 0; e < 4; ++e) grp[u][e] = ((v << 2) + e) / P.Cg;
        }
    }
    const int first = P.rows_per_stat > 0 ? s_first : 0;
    for (int d = d0 + ty; d < d1; d += P.rp) {
        const int sr = apply_src_row(P, d);
        const int which = (sr >= 0 && P.rows_per_stat > 0) ? sr / P.rows_per_stat - first : 0;
#pragma unroll
        for (int u = 0; u < kStatSlots; ++u) {
            const int v = tx + u * P.tpr;
            if (v >= nvec) continue;
            const int c = v << 2;
            uint2 o = make_uint2(0u, 0u);
            if (sr >= 0) {
                const float4 t = c < P.C1 ? __ldg(reinterpret_cast<const float4*>(P.x1 + (size_t)sr * P.C1 + c))
                                          : __ldg(reinterpret_cast<const float4*>(P.x2 + (size_t)sr * P.C2 + (c - P.C1)));
                float x[4] = {t.x, t.y, t.z, t.w};
                if (P.raw) {
                    __nv_bfloat162 a = __floats2bfloat162_rn(x[0], x[1]), b = __floats2bfloat162_rn(x[2], x[3]);
                    *reinterpret_cast<uint2*>(P.raw + (size_t)sr * C + c) = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
                }
                if (P.rows_per_stat > 0) {
                    const float g4[4] = {ga[u].x, ga[u].y, ga[u].z, ga[u].w};
                    const float b4[4] = {be[u].x, be[u].y, be[u].z, be[u].w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 mr = s_mr[which][grp[u][e]];
                        float y = (x[e] - mr.x) * mr.y * g4[e] + b4[e];
                        if (P.silu) y = silu_f(y);
                        x[e] = y;
                    }
                } else if (P.silu) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) x[e] = silu_f(x[e]);
                }
                __nv_bfloat162 a = __floats2bfloat162_rn(x[0], x[1]), b = __floats2bfloat162_rn(x[2], x[3]);
                o = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
            }
            *reinterpret_cast<uint2*>(P.dst + (size_t)d * C + c) = o;
        }
    }
}

int gn_apply(const ApplyArgs& a, cudaStream_t st) {
    const int C = a.C1 + a.C2;
    MOCA_REQUIRE(C % 4 == 0 && a.C1 % 4 == 0 && C <= 256 * kStatSlots * 4, MOCA_E_SHAPE, "gn_apply: channels C1=%d C2=%d unsupported", a.C1, a.C2);
    MOCA_REQUIRE(a.rows_per_stat == 0 || (a.groups == 32 && C % 32 == 0 && a.acc && a.gamma && a.beta), MOCA_E_ARG,
                 "gn_apply: bad normalisation arguments");
    MOCA_REQUIRE(a.rows_per_stat == 0 || a.rows % a.rows_per_stat == 0, MOCA_E_SHAPE, "gn_apply: bad statistics span");
    ApplyP P;
    P.x1 = a.x1; P.C1 = a.C1; P.x2 = a.x2; P.C2 = a.C2;
    P.rows_per_stat = a.rows_per_stat; P.Cg = C / 32 > 0 ? C / 32 : 1;
    P.acc = a.acc; P.gamma = a.gamma; P.beta = a.beta; P.eps = a.eps; P.silu = a.silu; P.mode = a.dst_mode;
    P.NF = a.NF; P.H = a.H; P.W = a.W; P.T = a.T; P.HW = a.HW;
    P.dst = a.dst; P.raw = a.raw;
    int64_t dst_rows = 0;
    switch (a.dst_mode) {
        case DST_LINEAR: dst_rows = a.rows; break;
        case DST_SPATIAL_PAD:
            MOCA_REQUIRE((int64_t)a.NF * a.H * a.W == a.rows, MOCA_E_SHAPE, "gn_apply: spatial geometry does not match rows");
            dst_rows = (int64_t)a.NF * (a.H + 2) * (a.W + 2); break;
        case DST_SPATIAL_PAD_UP2:
            MOCA_REQUIRE((int64_t)a.NF * a.H * a.W == a.rows, MOCA_E_SHAPE, "gn_apply: spatial geometry does not match rows");
            dst_rows = (int64_t)a.NF * (2 * a.H + 2) * (2 * a.W + 2); break;
        case DST_TIME_PAD:
            MOCA_REQUIRE((int64_t)a.B * a.T * a.HW == a.rows, MOCA_E_SHAPE, "gn_apply: time geometry does not match rows");
            dst_rows = (int64_t)a.B * (a.T + 2) * a.HW; break;
        default: MOCA_REQUIRE(false, MOCA_E_ARG, "gn_apply: unknown dst mode %d", a.dst_mode);
    }
    MOCA_REQUIRE(dst_rows < (1ll << 31) && a.rows < (1ll << 31), MOCA_E_SHAPE, "gn_apply: too many rows");
    P.dst_rows = (int)dst_rows;
    P.nstat = a.rows_per_stat > 0 ? (int)(a.rows / a.rows_per_stat) : 0;
    int slots;
    stat_geometry(C / 4, slots, P.tpr, P.rp);
    gn_apply_kernel<<<cdiv(dst_rows, kApplyRows), 256, 0, st>>>(P);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

// ---------------------------------------------------------------------------------------------------------------
constexpr int kLnSlots = 10;   // C <= 32*4*10 = 1280

__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, int64_t rows, int C,
                                                        __nv_bfloat16* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int nvec = C >> 2;
    const float4* xr = reinterpret_cast<const float4*>(x + row * C);
    float4 v[kLnSlots];
    float s = 0.f;
#pragma unroll
    for (int u = 0; u < kLnSlots; ++u) {
        const int i = lane + 32 * u;
        if (i < nvec) {
            v[u] = xr[i];
            s += (v[u].x + v[u].y) + (v[u].z + v[u].w);
        }
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / (float)C;
    float q = 0.f;
#pragma unroll
    for (int u = 0; u < kLnSlots; ++u) {
        const int i = lane + 32 * u;
        if (i < nvec) {
            const float a = v[u].x - mean, b = v[u].y - mean, c = v[u].z - mean, d = v[u].w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
    }
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q / (float)C + 1e-5f);
    uint2* orow = reinterpret_cast<uint2*>(out + row * C);
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    const float4* b4 = reinterpret_cast<const float4*>(beta);
#pragma unroll
    for (int u = 0; u < kLnSlots; ++u) {
        const int i = lane + 32 * u;
        if (i < nvec) {
            const float4 g = g4[i], b = b4[i];
            __nv_bfloat162 lo = __floats2bfloat162_rn((v[u].x - mean) * rstd * g.x + b.x, (v[u].y - mean) * rstd * g.y + b.y);
            __nv_bfloat162 hi = __floats2bfloat162_rn((v[u].z - mean) * rstd * g.z + b.z, (v[u].w - mean) * rstd * g.w + b.w);
            orow[i] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
        }
    }
}

int layernorm_bf16(const float* x, const float* gamma, const float* beta, int64_t rows, int C, __nv_bfloat16* out, cudaStream_t st) {
    MOCA_REQUIRE(C % 4 == 0 && C <= 128 * kLnSlots, MOCA_E_SHAPE, "layernorm: C=%d unsupported (multiple of 4, <= %d)", C, 128 * kLnSlots);
    layernorm_kernel<<<cdiv(rows, 8), 256, 0, st>>>(x, gamma, beta, rows, C, out);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

// ---------------------------------------------------------------------------------------------------------------
__global__ void timestep_embedding_kernel(const int* __restrict__ t, int n, int dim, __nv_bfloat16* __restrict__ out) {
    const int half = dim / 2;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * half) return;
    const int j = i % half, r = i / half;
    const float f = expf(-9.210340371976184f * (float)j / (float)half);      // -log(10000) * j / half
    const float a = (float)t[r] * f;
    out[(int64_t)r * dim + j] = __float2bfloat16_rn(cosf(a));
    out[(int64_t)r * dim + half + j] = __float2bfloat16_rn(sinf(a));
}

int timestep_embedding_bf16(const int* t_dev, int n, int dim, __nv_bfloat16* out, cudaStream_t st) {
    MOCA_REQUIRE(dim % 2 == 0, MOCA_E_SHAPE, "timestep_embedding: odd dim %d", dim);
    timestep_embedding_kernel<<<cdiv((int64_t)n * dim / 2, 256), 256, 0, st>>>(t_dev, n, dim, out);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

__global__ void silu_add_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t n, int do_silu,
                                __nv_bfloat16* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float v = a[i];
        if (b) v += b[i];
        if (do_silu) v = v / (1.0f + expf(-v));
        out[i] = __float2bfloat16_rn(v);
    }
}

int silu_add_bf16(const float* a, const float* b, int64_t n, __nv_bfloat16* out, cudaStream_t st) {
    const int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    silu_add_kernel<<<blocks, 256, 0, st>>>(a, b, n, 1, out);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

int f32_to_bf16(const float* a, int64_t n, __nv_bfloat16* out, cudaStream_t st) {
    const int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    silu_add_kernel<<<blocks, 256, 0, st>>>(a, nullptr, n, 0, out);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

// ---------------------------------------------------------------------------------------------------------------
__global__ void ncthw_to_pad_kernel(const float* __restrict__ x, int B, int C, int T, int H, int W, int Cp,
                                    __nv_bfloat16* __restrict__ dst) {
    const int Hp = H + 2, Wp = W + 2;
    const int64_t total = (int64_t)B * T * Hp * Wp * Cp;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % Cp);
        const int wp = (int)((i / Cp) % Wp);
        const int hp = (int)((i / ((int64_t)Cp * Wp)) % Hp);
        const int f = (int)(i / ((int64_t)Cp * Wp * Hp));
        float v = 0.f;
        if (c < C && hp >= 1 && hp <= H && wp >= 1 && wp <= W) {
            const int b = f / T, t = f % T;
            v = x[((((int64_t)b * C + c) * T + t) * H + (hp - 1)) * W + (wp - 1)];
        }
        dst[i] = __float2bfloat16_rn(v);
    }
}

int ncthw_to_pad(const float* x, int B, int C, int T, int H, int W, int Cp, __nv_bfloat16* dst, cudaStream_t st) {
    const int64_t total = (int64_t)B * T * (H + 2) * (W + 2) * Cp;
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    ncthw_to_pad_kernel<<<blocks, 256, 0, st>>>(x, B, C, T, H, W, Cp, dst);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

__global__ void tokens_to_ncthw_kernel(const float* __restrict__ tok, int ld, int B, int C, int T, int H, int W, float* __restrict__ y) {
    const int64_t total = (int64_t)B * C * T * H * W;
    const int HW = H * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int hw = (int)(i % HW);
        const int t = (int)((i / HW) % T);
        const int c = (int)((i / ((int64_t)HW * T)) % C);
        const int b = (int)(i / ((int64_t)HW * T * C));
        y[i] = tok[(((int64_t)b * T + t) * HW + hw) * ld + c];
    }
}

int tokens_to_ncthw(const float* tok, int ld, int B, int C, int T, int H, int W, float* y, cudaStream_t st) {
    const int64_t total = (int64_t)B * C * T * H * W;
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    tokens_to_ncthw_kernel<<<blocks, 256, 0, st>>>(tok, ld, B, C, T, H, W, y);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

int nhwc_to_pad(const float* x, int NF, int H, int W, int C, __nv_bfloat16* dst, cudaStream_t st) {
    ApplyArgs a;
    a.x1 = x; a.C1 = C; a.rows = (int64_t)NF * H * W; a.rows_per_stat = 0; a.dst_mode = DST_SPATIAL_PAD;
    a.NF = NF; a.H = H; a.W = W; a.dst = dst;
    return gn_apply(a, st);
}

}  // namespace moca
