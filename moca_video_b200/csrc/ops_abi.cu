// Primitive entry points used by the unit parity tests (tests/test_gpu_ops.py): they run the same kernels the UNet plan
// uses (tcgen05 GEMM core, shifted-row implicit-GEMM convolutions, fused attention) on caller tensors.
// These helpers allocate and free scratch per call and synchronise: test surface, not the hot path.
#include "attn.cuh"
#include "gemm_tc.cuh"
#include "norm.cuh"

using namespace moca;

#include <vector>

namespace {
template <typename T>
static T* dmalloc(size_t n) {
    void* p = nullptr;
    if (cudaMalloc(&p, n * sizeof(T) + 16) != cudaSuccess) return nullptr;
    return (T*)p;
}
static int round_up64(int v) { return (v + 63) / 64 * 64; }
}  // namespace

extern "C" int32_t moca_gemm_op16(moca_ctx* ctx, const void* A, const void* Wt, const float* bias, int32_t M, int32_t N, int32_t K,
                                  void* D, int32_t out_op16, void* stream) {
    MOCA_REQUIRE(ctx && A && Wt && D, MOCA_E_ARG, "moca_gemm_op16: NULL argument");
    MOCA_REQUIRE(K % 64 == 0, MOCA_E_SHAPE, "moca_gemm_op16: K=%d must be a multiple of 64", K);
    GemmDesc d;
    d.A = A; d.a_rows = M; d.a_cols = K; d.a_pitch = K; d.Wt = Wt; d.N = N; d.ntaps = 1; d.Kp = K;
    d.mode = GEMM_LINEAR; d.M = M; d.epi = out_op16 ? EPI_BF16 : EPI_F32; d.out = D; d.ldo = N; d.bias = bias;
    GemmLaunch L;
    int rc = gemm_plan(ctx->c, d, &L);
    if (rc) return rc;
    return gemm_run(L, (cudaStream_t)stream);
}

// GEGLU-fused FeedForward first layer (attention.py:376-383): out bf16 [M, N/2] from fp32 weight [N, K], bias [N]
extern "C" int32_t moca_geglu_f32(moca_ctx* ctx, const void* A_op16, const float* w, const float* bias, int32_t M, int32_t N, int32_t K,
                                  void* out_op16, void* stream) {
    MOCA_REQUIRE(ctx && A_op16 && w && bias && out_op16, MOCA_E_ARG, "moca_geglu_f32: NULL argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int Kp = round_up64(K);
    op_t* wp = dmalloc<op_t>((size_t)N * Kp);
    float* bp = dmalloc<float>(N);
    MOCA_REQUIRE(wp && bp, MOCA_E_CUDA, "moca_geglu_f32: allocation failed");
    cudaMemsetAsync(wp, 0, (size_t)N * Kp * 2, st);
    int rc = pack_weight(w, N, K, 1, Kp, wp, Kp, 0, 256, st);
    // the bias takes the same row permutation as the weight (value rows / gate rows per 256-row tile)
    std::vector<float> hb(N), hp(N);
    cudaStreamSynchronize(st);
    cudaMemcpy(hb.data(), bias, N * sizeof(float), cudaMemcpyDeviceToHost);
    for (int i = 0; i < N; ++i) {
        const int half = N / 2, hbk = 128;
        const int j = i < half ? i : i - half;
        hp[(j / hbk) * 256 + (i < half ? 0 : hbk) + (j % hbk)] = hb[i];
    }
    cudaMemcpy(bp, hp.data(), N * sizeof(float), cudaMemcpyHostToDevice);
    GemmDesc d;
    d.A = A_op16; d.a_rows = M; d.a_cols = K; d.a_pitch = K; d.Wt = wp; d.N = N; d.ntaps = 1; d.Kp = Kp;
    d.mode = GEMM_LINEAR; d.M = M; d.epi = EPI_GEGLU; d.out = out_op16; d.ldo = N / 2; d.bias = bp;
    GemmLaunch L;
    if (!rc) rc = gemm_plan(ctx->c, d, &L);
    if (!rc) rc = gemm_run(L, st);
    cudaStreamSynchronize(st);
    cudaFree(wp); cudaFree(bp);
    return rc;
}

// conv3x3 (padding 1, stride 1 or 2) on channels-last fp32 tokens [NF, H, W, Cin] with a torch-layout weight [Cout, Cin, 3, 3]
extern "C" int32_t moca_conv3x3_f32(moca_ctx* ctx, const float* x_nhwc, const float* w_oihw, const float* bias, int32_t NF, int32_t H,
                                    int32_t W, int32_t Cin, int32_t Cout, int32_t stride, float* y_nhwc, void* stream) {
    MOCA_REQUIRE(ctx && x_nhwc && w_oihw && y_nhwc, MOCA_E_ARG, "moca_conv3x3_f32: NULL argument");
    MOCA_REQUIRE(Cin % 8 == 0 && (stride == 1 || stride == 2), MOCA_E_SHAPE, "moca_conv3x3_f32: Cin %% 8 and stride in {1,2}");
    cudaStream_t st = (cudaStream_t)stream;
    const int Kp = round_up64(Cin), Hp = H + 2, Wp = W + 2;
    op_t* wp = dmalloc<op_t>((size_t)Cout * 9 * Kp);
    op_t* pad = dmalloc<op_t>((size_t)NF * Hp * Wp * Cin);
    MOCA_REQUIRE(wp && pad, MOCA_E_CUDA, "moca_conv3x3_f32: allocation failed");
    int rc = pack_weight(w_oihw, Cout, Cin, 9, Kp, wp, (int64_t)9 * Kp, 0, 0, st);
    GemmDesc d;
    d.a_cols = Cin; d.a_pitch = Cin; d.Wt = wp; d.N = Cout; d.ntaps = 9; d.Kp = Kp;
    d.g_H = H; d.g_W = W; d.stride = stride;
    d.A = pad;
    if (conv4d_supported(NF, H, W, stride)) {
        // unpadded tokens through the 4-D tensor map (the path the UNet plan takes for its own geometries)
        ApplyArgs a;
        a.x1 = x_nhwc; a.C1 = Cin; a.rows = (int64_t)NF * H * W; a.rows_per_stat = 0; a.dst_mode = DST_LINEAR; a.NF = NF; a.H = H; a.W = W; a.dst = pad;
        if (!rc) rc = gn_apply(a, st);
        d.a_rows = (int64_t)NF * H * W; d.mode = GEMM_SPATIAL_4D; d.M = NF * (H / stride) * (W / stride);
    } else {
        if (!rc) rc = nhwc_to_pad(x_nhwc, NF, H, W, Cin, pad, st);
        d.a_rows = (int64_t)NF * Hp * Wp;
        for (int ky = 0; ky < 3; ++ky)
            for (int kx = 0; kx < 3; ++kx) d.tap_shift[ky * 3 + kx] = (ky - 1) * Wp + (kx - 1);
        d.mode = GEMM_SPATIAL_PAD; d.g_rows = Hp * Wp; d.g_w = Wp;
        d.M = (int)d.a_rows;
    }
    d.epi = EPI_F32; d.out = y_nhwc; d.ldo = Cout; d.bias = bias; d.rows_per_frame = (H / stride) * (W / stride);
    GemmLaunch L;
    if (!rc) rc = gemm_plan(ctx->c, d, &L);
    if (!rc) rc = gemm_run(L, st);
    cudaStreamSynchronize(st);
    cudaFree(wp); cudaFree(pad);
    return rc;
}

// Conv3d (3,1,1), padding (1,0,0) on tokens [B, T, HW, C] fp32 with torch-layout weight [Cout, Cin, 3, 1, 1]
extern "C" int32_t moca_tconv_f32(moca_ctx* ctx, const float* x, const float* w, const float* bias, int32_t B, int32_t T, int32_t HW,
                                  int32_t Cin, int32_t Cout, float* y, void* stream) {
    MOCA_REQUIRE(ctx && x && w && y, MOCA_E_ARG, "moca_tconv_f32: NULL argument");
    MOCA_REQUIRE(Cin % 8 == 0, MOCA_E_SHAPE, "moca_tconv_f32: Cin %% 8");
    cudaStream_t st = (cudaStream_t)stream;
    const int Kp = round_up64(Cin);
    op_t* wp = dmalloc<op_t>((size_t)Cout * 3 * Kp);
    op_t* pad = dmalloc<op_t>((size_t)B * (T + 2) * HW * Cin);
    MOCA_REQUIRE(wp && pad, MOCA_E_CUDA, "moca_tconv_f32: allocation failed");
    int rc = pack_weight(w, Cout, Cin, 3, Kp, wp, (int64_t)3 * Kp, 0, 0, st);
    ApplyArgs a;
    a.x1 = x; a.C1 = Cin; a.rows = (int64_t)B * T * HW; a.rows_per_stat = 0; a.dst_mode = DST_TIME_PAD; a.B = B; a.T = T; a.HW = HW; a.dst = pad;
    if (!rc) rc = gn_apply(a, st);
    GemmDesc d;
    d.A = pad; d.a_rows = (int64_t)B * (T + 2) * HW; d.a_cols = Cin; d.a_pitch = Cin; d.Wt = wp; d.N = Cout; d.ntaps = 3; d.Kp = Kp;
    for (int dt = 0; dt < 3; ++dt) d.tap_shift[dt] = (dt - 1) * HW;
    d.mode = GEMM_TIME_PAD; d.g_rows = (T + 2) * HW; d.g_w = HW; d.g_H = T;
    d.M = (int)d.a_rows; d.epi = EPI_F32; d.out = y; d.ldo = Cout; d.bias = bias;
    GemmLaunch L;
    if (!rc) rc = gemm_plan(ctx->c, d, &L);
    if (!rc) rc = gemm_run(L, st);
    cudaStreamSynchronize(st);
    cudaFree(wp); cudaFree(pad);
    return rc;
}

// GroupNorm(32) (+SiLU) on tokens [rows, C] fp32 -> bf16 [rows, C]; rows_per_stat rows share statistics
extern "C" int32_t moca_groupnorm_op16(moca_ctx* ctx, const float* x, const float* gamma, const float* beta, int64_t rows, int32_t C,
                                       int32_t rows_per_stat, float eps, int32_t silu, void* out_op16, void* stream) {
    MOCA_REQUIRE(ctx && x && gamma && beta && out_op16, MOCA_E_ARG, "moca_groupnorm_op16: NULL argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int nstat = (int)(rows / rows_per_stat);
    double* acc = dmalloc<double>((size_t)nstat * 64);
    float2* table = dmalloc<float2>((size_t)nstat * 32);
    unsigned* counter = dmalloc<unsigned>((size_t)nstat);
    MOCA_REQUIRE(acc && table && counter, MOCA_E_CUDA, "moca_groupnorm_op16: allocation failed");
    cudaMemsetAsync(acc, 0, (size_t)nstat * 64 * sizeof(double), st);
    cudaMemsetAsync(counter, 0, (size_t)nstat * sizeof(unsigned), st);
    int rc = gn_stats(x, C, nullptr, 0, rows, rows_per_stat, 32, acc, eps, table, counter, st);
    ApplyArgs a;
    a.x1 = x; a.C1 = C; a.rows = rows; a.rows_per_stat = rows_per_stat; a.table = table; a.gamma = gamma; a.beta = beta; a.eps = eps;
    a.silu = silu; a.dst_mode = DST_LINEAR; a.dst = (op_t*)out_op16;
    if (!rc) rc = gn_apply(a, st);
    cudaStreamSynchronize(st);
    cudaFree(acc); cudaFree(table); cudaFree(counter);
    return rc;
}

extern "C" int32_t moca_layernorm_op16(moca_ctx* ctx, const float* x, const float* gamma, const float* beta, int64_t rows, int32_t C,
                                       void* out_op16, void* stream) {
    MOCA_REQUIRE(ctx && x && gamma && beta && out_op16, MOCA_E_ARG, "moca_layernorm_op16: NULL argument");
    return layernorm_op16(x, gamma, beta, rows, C, (op_t*)out_op16, (cudaStream_t)stream);
}

// softmax(q k^T / 8) v, head dim 64.  q [batch*Lq, heads*64] (pitch q_stride), k/v [batch*Lk, heads*64] (pitch kv_stride).
// temporal != 0: sequences are interleaved (element i of sequence o is row (o / inner)*L*inner + (o % inner) + i*inner), Lq == Lk <= 16.
extern "C" int32_t moca_attention_op16(moca_ctx* ctx, const void* q, const void* k, const void* v, int32_t batch, int32_t heads,
                                       int32_t Lq, int32_t Lk, int32_t Lk_valid, int32_t q_stride, int32_t kv_stride, void* out,
                                       int32_t o_stride, int32_t temporal_inner, void* stream) {
    MOCA_REQUIRE(ctx && q && k && v && out, MOCA_E_ARG, "moca_attention_op16: NULL argument");
    AttnArgs a;
    a.q = (const op_t*)q; a.k = (const op_t*)k; a.v = (const op_t*)v; a.out = (op_t*)out;
    a.ldq = q_stride; a.ldkv = kv_stride; a.ldo = o_stride; a.n_outer = batch; a.heads = heads; a.Lq = Lq; a.Lk = Lk_valid;
    a.scale = 0.125f;
    if (temporal_inner > 0) {
        a.qmap.div = temporal_inner; a.qmap.mul_a = Lq * temporal_inner; a.qmap.mul_b = 1; a.qmap.step = temporal_inner;
        a.kvmap = a.qmap;
        return attention_short(a, (cudaStream_t)stream);
    }
    a.qmap.div = 1; a.qmap.mul_a = Lq; a.qmap.step = 1;
    a.kvmap.div = 1; a.kvmap.mul_a = Lk; a.kvmap.step = 1;
    a.q_rows = (int64_t)batch * Lq; a.kv_rows = (int64_t)batch * Lk;
    if (attention_tc_supported(a)) {
        AttnTcLaunch l;
        int rc = attention_tc_plan(a, &l);
        if (rc) return rc;
        return attention_tc_run(l, (cudaStream_t)stream);
    }
    return attention_flash(a, (cudaStream_t)stream);
}
