// See unet.cuh.  Part 2: per-shape execution plan of UNetModel.forward (openaimodel3d.py:534-578) and its modules.
//
// Activation layout: channels-last token matrix [B*T*H*W, C], token order (b, t, h, w) at every level.  The residual
// stream and every GroupNorm/LayerNorm input stay fp32; every GEMM operand is a bf16 copy produced by the norm that
// precedes it (fused normalise + SiLU + layout scatter), so no rearrange/permute/contiguous copy of the reference
// (openaimodel3d.py:42-45,230-233,552; attention.py:264-277,333-372) exists here: temporal modules address the same
// matrix with a row stride of H*W.
#include "unet.cuh"

namespace moca {

typedef std::function<int(cudaStream_t)> Op;

struct Tap {
    std::string name;
    float* buf = nullptr;
    int64_t rows = 0;
    int C = 0;
};

struct Plan {
    int B = 0, T = 0, H = 0, W = 0, Lmax = 0;
    std::vector<Op> ops;
    std::vector<std::string> labels;      // one per op (profiling: moca_unet_profile)
    std::vector<double> op_flops;
    int launches = 0;
    uint64_t last_use = 0;                // LRU stamp of the plan cache
    double flops = 0;
    std::vector<void*> allocs;
    float* x_in = nullptr;
    float* ctx_in = nullptr;
    int* d_ts = nullptr;
    int* d_fps = nullptr;
    int* d_ctxlen = nullptr;
    float* out_tok = nullptr;
    float* emb_in = nullptr;              // layer plans: [NF, 4 mc] time (+ fps) embedding input of a ResBlock
    int out_C = 0;                        // layer plans: channels of out_tok
    double* stats = nullptr;
    size_t stats_bytes = 0;
    float2* ab_table = nullptr;       // (scale, shift) per (statistics group, channel) of the GroupNorm being applied (shared)
    cudaGraphExec_t gexec = nullptr;
    std::vector<Tap> taps;
    bool with_taps = false;
    size_t workspace_bytes = 0;
    ~Plan() {
        if (gexec) cudaGraphExecDestroy(gexec);
        for (void* p : allocs) cudaFree(p);
    }
};

struct Act {
    float* p = nullptr;
    int C = 0, H = 0, W = 0;
    StatSrc st;               // GroupNorm statistics material left by the GEMM epilogue that wrote p
};

// x [B, C, T, H, W] fp32 -> 3x3 patches as 16-bit rows [B*T*H*W, 64]: column c * 9 + ky * 3 + kx = x[b, c, t, h + ky - 1, w + kx - 1]
// (zero outside the image and for columns >= 9 C): the stem convolution (openaimodel3d.py:375-379) becomes one k-block of a Linear
__global__ void ncthw_patches_kernel(const float* __restrict__ x, int B, int C, int T, int H, int W, op_t* __restrict__ dst) {
    const int64_t total = (int64_t)B * T * H * W * 64;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i & 63);
        const int64_t r = i >> 6;
        const int w = (int)(r % W), h = (int)((r / W) % H), f = (int)(r / ((int64_t)W * H));
        float v = 0.f;
        if (k < 9 * C) {
            const int c = k / 9, tap = k - 9 * c;
            const int hh = h + tap / 3 - 1, ww = w + tap % 3 - 1;
            if (hh >= 0 && hh < H && ww >= 0 && ww < W) v = x[((((int64_t)(f / T) * C + c) * T + (f % T)) * H + hh) * W + ww];
        }
        dst[i] = float2op(v);
    }
}

enum ScratchId { S_PAD = 0, S_RAW, S_LIN, S_QKV, S_ATT, S_FFH, S_F32A, S_F32B, S_Y, S_COUNT };

struct Builder {
    moca_unet* net;
    Plan* plan;
    bool dry;
    int rc = MOCA_OK;
    size_t need[S_COUNT] = {0};
    void* sptr[S_COUNT] = {nullptr};
    size_t cs_doubles = 0;    // bump allocator over plan->stats (zeroed at the start of every forward)
    int B, T, H0, W0, NF, Lmax;
    std::string scope;

    void fail(int code) { if (rc == MOCA_OK) rc = code; }
    void raw_push(Op op, const std::string& label, double flops) {
        plan->ops.push_back(std::move(op));
        plan->labels.push_back(scope + " " + label);
        plan->op_flops.push_back(flops);
    }

    void* palloc(size_t bytes) {
        if (dry) { plan->workspace_bytes += bytes; return nullptr; }
        void* p = nullptr;
        if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) { set_error("plan: cudaMalloc of %zu bytes failed", bytes); fail(MOCA_E_CUDA); return nullptr; }
        plan->allocs.push_back(p);
        return p;
    }
    template <typename Tp>
    Tp* scratch(int id, size_t elems) {
        const size_t bytes = elems * sizeof(Tp);
        if (dry) { if (bytes > need[id]) need[id] = bytes; return nullptr; }
        return (Tp*)sptr[id];
    }
    void push(Op op, int nlaunch = 1, const std::string& label = "op", double flops = 0) {
        plan->launches += nlaunch;
        if (!dry) raw_push(std::move(op), label, flops);
    }
    // Statistics buffer of a tensor with C channels: {sum, sumsq} per (frame, channel), accumulated by the producing GEMM's epilogue
    // with fp64 atomics.  Always per frame, also when the consumer normalises over the T frames of a sample (the 5-D GroupNorms
    // of TemporalConvBlock / TemporalTransformer; the apply kernel adds the T entries): per-sample entries would put
    // every atomic of the GEMM on B*C addresses, and same-address fp64 atomics serialise at ~40 ns each in L2 (measured:
    // the 81920 x 320 proj_out GEMM went from 56 to 107 us).
    // Large frames (the VAE decoder's 160 x 256 and 320 x 512 levels: thousands of M tiles on one or two frames) split a frame
    // into entries of <= 2048 rows for the same reason; gn_finalize adds the entries of a group.
    StatSrc stat_alloc(int C, int HW) {
        StatSrc r;
        r.C = C;
        r.rows = HW;
        if (HW > 4096)
            for (int e = 2048; e >= 32; e -= 32)
                if (HW % e == 0) { r.rows = e; break; }
        const size_t off = cs_doubles;
        cs_doubles += (size_t)NF * (HW / r.rows) * C * 2;
        r.cs = dry ? (const double*)16 : plan->stats + off;
        return r;
    }

    // ---- GEMM flavours -------------------------------------------------------------------------------------
    void gemm(GemmDesc d, double useful_flops) {
        plan->launches += 1;
        plan->flops += useful_flops;
        if (dry || rc) return;
        GemmLaunch L;
        int r = gemm_plan(net->ctx->c, d, &L);
        if (r) { fail(r); return; }
        char buf[192];
        snprintf(buf, sizeof(buf), "gemm M=%d N=%d K=%d taps=%d BN=%d tiles=%d epi=%d mode=%d", d.M, d.N, d.a_cols, d.ntaps, L.BN,
                 L.p.num_m_tiles * L.p.num_n_tiles, d.epi, d.mode);
        raw_push([L](cudaStream_t st) { return gemm_run(L, st); }, buf, useful_flops);
    }
    void linear(const op_t* A, int64_t rows, int K, const PackedW& w, int epi, void* out, int ldo, const float* residual,
                int n_override = 0, const StatSrc* ss = nullptr, int rpf = 0, const op_t* residual16 = nullptr) {
        GemmDesc d;
        d.A = A; d.a_rows = rows; d.a_cols = K; d.a_pitch = K;
        d.Wt = w.w; d.N = n_override ? n_override : w.N; d.ntaps = 1; d.Kp = w.Kp;
        d.mode = GEMM_LINEAR; d.M = (int)rows; d.epi = epi; d.out = out; d.ldo = ldo; d.bias = w.b; d.residual = residual; d.residual16 = residual16;
        if (ss) { d.chan_stats = (double*)ss->cs; d.stat_rows_per_frame = ss->rows; }
        d.stat_sample_rows = T * rpf;
        gemm(d, 2.0 * rows * d.N * K);
    }
    // 3x3 conv of the 16-bit operand `src`: unpadded tokens [NF, H, W, Cin] through the 4-D tensor map when the geometry allows
    // (conv_dst / conv4d tell the producer of src which layout to write), else the zero-bordered [NF, H+2, W+2, Cin] buffer
    // (at 5x8 the unpadded matrix has 10 M tiles: fewer CTAs than SMs and slower than the 18 padded ones, 83 vs 69 us measured;
    // the stride-2 convolutions always win by not computing the dropped rows)
    bool conv4d(int H, int W, int stride) const {
        return conv4d_supported(NF, H, W, stride) && (stride == 2 || (int64_t)NF * H * W >= 4096);
    }
    int conv_dst(int H, int W, int stride) const { return conv4d(H, W, stride) ? DST_LINEAR : DST_SPATIAL_PAD; }
    void conv3x3(const op_t* src, int H, int W, int Cin, const PackedW& w, int stride, void* out, const float* rowbias,
                 int rb_ld, const float* residual, const StatSrc* ss, int out16 = 0, int pad_lo = 1) {
        GemmDesc d;
        d.pad_lo = pad_lo;
        if (pad_lo != 1 && !conv4d(H, W, stride)) {
            set_error("plan: a convolution with padding (0,1,0,1) needs the 4-D tensor-map geometry (NF=%d H=%d W=%d stride=%d)", NF, H, W, stride);
            fail(MOCA_E_SHAPE);
            return;
        }
        const int Ho = H / stride, Wo = W / stride;
        d.a_cols = Cin; d.a_pitch = Cin;
        d.Wt = w.w; d.N = w.N; d.ntaps = 9; d.Kp = w.Kp;
        d.g_H = H; d.g_W = W; d.stride = stride;
        if (conv4d(H, W, stride)) {
            d.A = src; d.a_rows = (int64_t)NF * H * W;
            d.mode = GEMM_SPATIAL_4D; d.M = NF * Ho * Wo;
            d.stat_sample_rows = T * Ho * Wo;
        } else {
            const int Hp = H + 2, Wp = W + 2;
            d.A = src; d.a_rows = (int64_t)NF * Hp * Wp;
            for (int ky = 0; ky < 3; ++ky)
                for (int kx = 0; kx < 3; ++kx) d.tap_shift[ky * 3 + kx] = (ky - 1) * Wp + (kx - 1);
            d.mode = GEMM_SPATIAL_PAD; d.g_rows = Hp * Wp; d.g_w = Wp;
            d.M = (int)d.a_rows;
            d.stat_sample_rows = T * Hp * Wp;
        }
        d.epi = EPI_F32; d.out = out; d.ldo = w.N; d.bias = w.b; d.out16 = out16;
        d.rowbias = rowbias; d.rb_ld = rb_ld; d.rows_per_frame = Ho * Wo; d.residual = residual;
        if (ss) { d.chan_stats = (double*)ss->cs; d.stat_rows_per_frame = ss->rows; }
        gemm(d, 2.0 * NF * Ho * Wo * w.N * 9.0 * w.K);
    }
    void tconv(const op_t* padT, int HW, int C, const PackedW& w, void* out, const float* residual, const StatSrc* ss, int out16 = 0) {
        GemmDesc d;
        d.A = padT; d.a_rows = (int64_t)B * (T + 2) * HW; d.a_cols = C; d.a_pitch = C;
        d.Wt = w.w; d.N = w.N; d.ntaps = 3; d.Kp = w.Kp;
        for (int dt = 0; dt < 3; ++dt) d.tap_shift[dt] = (dt - 1) * HW;
        d.mode = GEMM_TIME_PAD; d.g_rows = (T + 2) * HW; d.g_w = HW; d.g_H = T;
        d.M = (int)d.a_rows; d.epi = EPI_F32; d.out = out; d.ldo = w.N; d.bias = w.b; d.residual = residual; d.out16 = out16;
        if (ss) { d.chan_stats = (double*)ss->cs; d.stat_rows_per_frame = ss->rows; }
        d.stat_sample_rows = (T + 2) * HW;
        gemm(d, 2.0 * NF * HW * w.N * 3.0 * C);
    }

    // ---- normalisation flavours ----------------------------------------------------------------------------
    // GroupNorm(32) (+ SiLU) of x1 (fp32, or 16-bit with in16) [+ virtual channel concat with x2] over rows_per_stat rows per group,
    // statistics from the producers' epilogues (finalised inside the apply kernel), scattered into the consumer's operand layout.
    void gnorm(const void* x1, const StatSrc& s1, int C1, const float* x2, const StatSrc& s2in, int C2, int64_t rows, int rows_per_stat,
               const NormW& w, float eps, int silu, int dst_mode, int H, int W, op_t* dst, op_t* raw, int in16 = 0) {
        StatSrc s2 = s2in;
        if (!x2) s2 = StatSrc();
        if (!dry && (s1.C != C1 || s2.C != C2 || !s1.cs || (x2 && !s2.cs))) {
            set_error("plan: statistics source does not match the GroupNorm input (%d/%d vs %d/%d)", s1.C, s2.C, C1, C2); fail(MOCA_E_STATE);
        }
        if (!dry && (s1.rows <= 0 || rows_per_stat % s1.rows != 0 || (x2 && s2.rows != s1.rows))) {
            set_error("plan: statistics of %d rows per entry cannot feed a GroupNorm over %d rows", s1.rows, rows_per_stat); fail(MOCA_E_STATE);
        }
        ApplyArgs a;
        a.x1 = (const float*)x1; a.C1 = C1; a.x2 = x2; a.C2 = C2; a.rows = rows; a.rows_per_stat = rows_per_stat;
        a.gamma = w.g; a.beta = w.b; a.eps = eps; a.silu = silu; a.dst_mode = dst_mode; a.in16 = in16;
        a.NF = NF; a.H = H; a.W = W; a.B = B; a.T = T; a.HW = H * W; a.dst = dst; a.raw = raw;
        // (scale, shift) per (statistics group, channel): one buffer for all GroupNorms, finalize -> apply are adjacent in the stream
        float2* ab = dry ? nullptr : plan->ab_table;
        a.ab = ab;
        const int nstat = (int)(rows / rows_per_stat);
        const float* gam = w.g; const float* bet = w.b;
        const std::string dims = " rows=" + std::to_string(rows) + " C=" + std::to_string(C1 + C2) + " dst=" + std::to_string(dst_mode) +
                                 (in16 ? " in16" : "");
        push([=](cudaStream_t st) { return gn_finalize(s1, s2, nstat, rows_per_stat, eps, gam, bet, ab, st); }, 1, "gn_finalize" + dims);
        push([=](cudaStream_t st) { return gn_apply(a, st); }, 1, "gn_apply" + dims);
    }
    void convert(const float* x, int C, int64_t rows, int dst_mode, int H, int W, op_t* dst) {
        ApplyArgs a;
        a.x1 = x; a.C1 = C; a.rows = rows; a.rows_per_stat = 0; a.dst_mode = dst_mode;
        a.NF = NF; a.H = H; a.W = W; a.B = B; a.T = T; a.HW = H * W; a.dst = dst;
        push([=](cudaStream_t st) { return gn_apply(a, st); }, 1, "convert rows=" + std::to_string(rows) + " C=" + std::to_string(C) + " dst=" + std::to_string(dst_mode));
    }
    void lnorm(const op_t* x, const NormW& w, int64_t rows, int C, op_t* out) {
        push([=](cudaStream_t st) { return layernorm_op16_in16(x, w.g, w.b, rows, C, out, st); }, 1, "layernorm rows=" + std::to_string(rows) + " C=" + std::to_string(C));
    }
    void tap(const std::string& name, const Act& a) {
        if (!plan->with_taps) return;
        const int64_t rows = (int64_t)NF * a.H * a.W;
        Tap t;
        t.name = name; t.rows = rows; t.C = a.C;
        t.buf = (float*)palloc((size_t)rows * a.C * sizeof(float));
        if (dry) return;
        plan->taps.push_back(t);
        const float* src = a.p;
        float* dst = t.buf;
        const size_t bytes = (size_t)rows * a.C * sizeof(float);
        raw_push([=](cudaStream_t st) {
            return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st) == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, "tap", 0);
    }

    // rowbias [NF, rb_ld]: this block's emb_layers output; kv [B*Lmax, kv_ld]: this block's cross-attention keys | values
    Act res_block(const LayerW& L, const Act& x, const Act* skip, const float* rowbias, int rb_ld);
    void transformer(const LayerW& L, Act& x, const op_t* kv, int kv_ld);
    void build();
    void build_layer(const LayerW& L);
    Act vae_res(const VaeResW& r, const Act& x, const std::string& name);
    void vae_attn(Act& x);
    void build_vae(float inv_scale);
    void build_vae_encoder();
    void build_text();
};

// ResBlock._forward (openaimodel3d.py:208-234) followed by its TemporalConvBlock (:269-276)
Act Builder::res_block(const LayerW& L, const Act& x, const Act* skip, const float* rowbias, int rb_ld) {
    const ResW& r = L.res;
    scope = L.name;
    const int H = x.H, W = x.W, HW = H * W;
    const int64_t rows = (int64_t)NF * HW;
    const int C2 = skip ? skip->C : 0;
    const int Cin = x.C + C2;
    if (Cin != r.cin) { set_error("plan: %s expects %d input channels, got %d", L.name.c_str(), r.cin, Cin); fail(MOCA_E_SHAPE); }
    if (skip && !r.has_skip) { set_error("plan: %s concatenates a skip tensor but has an identity skip_connection", L.name.c_str()); fail(MOCA_E_SHAPE); }
    const size_t pad_elems = (size_t)NF * (H + 2) * (W + 2);
    op_t* pad = scratch<op_t>(S_PAD, pad_elems * (Cin > r.cout ? Cin : r.cout));
    op_t* raw = r.has_skip ? scratch<op_t>(S_RAW, (size_t)rows * Cin) : nullptr;
    // tensors consumed only by the next GroupNorm are kept in 16 bits (their statistics come from the fp32 accumulators' epilogue)
    op_t* t1 = scratch<op_t>(S_F32A, (size_t)rows * r.cout);
    op_t* t2 = scratch<op_t>(S_F32B, (size_t)rows * r.cout);
    Act out;
    out.C = r.cout; out.H = H; out.W = W;
    out.p = (float*)palloc((size_t)rows * r.cout * sizeof(float));
    // in_layers: GN(fp32, eps 1e-5) -> SiLU -> conv3x3 ; + emb_layers(emb)[:, :, None, None]
    gnorm(x.p, x.st, x.C, skip ? skip->p : nullptr, skip ? skip->st : StatSrc(), C2, rows, HW, r.gn1, 1e-5f, 1, conv_dst(H, W, 1), H, W, pad, raw);
    const StatSrc st_t1 = stat_alloc(r.cout, HW);
    conv3x3(pad, H, W, Cin, r.conv1, 1, t1, rowbias, rb_ld, nullptr, &st_t1, 1);
    // out_layers: GN -> SiLU -> (dropout) -> conv3x3 ; + skip_connection(x)
    gnorm(t1, st_t1, r.cout, nullptr, StatSrc(), 0, rows, HW, r.gn2, 1e-5f, 1, conv_dst(H, W, 1), H, W, pad, nullptr, 1);
    StatSrc st_cur = stat_alloc(r.cout, HW);
    if (r.has_skip) {
        linear(raw, rows, Cin, r.skip, EPI_F32, out.p, r.cout, nullptr);
        conv3x3(pad, H, W, r.cout, r.conv2, 1, out.p, nullptr, 0, out.p, &st_cur);
    } else {
        conv3x3(pad, H, W, r.cout, r.conv2, 1, out.p, nullptr, 0, x.p, &st_cur);
    }
    // TemporalConvBlock: 4 x [GN over (C/32, T, H, W) -> SiLU -> Conv3d (3,1,1)] + identity
    const size_t padt_elems = (size_t)B * (T + 2) * HW * r.cout;
    op_t* padT = scratch<op_t>(S_PAD, padt_elems);
    const void* src = out.p;
    void* dsts[4] = {t1, t2, t1, out.p};
    for (int i = 0; i < 4; ++i) {
        gnorm(src, st_cur, r.cout, nullptr, StatSrc(), 0, rows, T * HW, r.tgn[i], 1e-5f, 1, DST_TIME_PAD, H, W, padT, nullptr, i > 0);
        st_cur = stat_alloc(r.cout, HW);
        tconv(padT, HW, r.cout, r.tconv[i], dsts[i], i == 3 ? out.p : nullptr, &st_cur, i < 3);
        src = dsts[i];
    }
    out.st = st_cur;
    return out;
}


// SpatialTransformer.forward (attention.py:262-278) / TemporalTransformer.forward (:331-373) with one
// BasicTransformerBlock (:216-220): x += attn1(LN1 x); x += attn2(LN2 x, ctx); x += ff(LN3 x).  Runs in place on x.
void Builder::transformer(const LayerW& L, Act& x, const op_t* kv, int kv_ld) {
    const TfW& t = L.tf;
    scope = L.name;
    const int H = x.H, W = x.W, HW = H * W;
    const int64_t rows = (int64_t)NF * HW;
    const int C = t.C, I = t.inner;
    if (x.C != C) { set_error("plan: %s expects %d channels, got %d", L.name.c_str(), C, x.C); fail(MOCA_E_SHAPE); }
    op_t* lin = scratch<op_t>(S_LIN, (size_t)rows * (C > I ? C : I));
    op_t* qkv = scratch<op_t>(S_QKV, (size_t)rows * 3 * I);
    op_t* att = scratch<op_t>(S_ATT, (size_t)rows * I);
    op_t* ffh = scratch<op_t>(S_FFH, (size_t)rows * 4 * I);
    // inner residual stream of the block (proj_in output + attn1 + attn2 + ff): 16-bit, updated in place by the 16-bit epilogues
    op_t* y = scratch<op_t>(S_Y, (size_t)rows * I);
    // norm (GroupNorm eps 1e-6; spatial: per frame, temporal: 5-D input so statistics span all T frames)
    gnorm(x.p, x.st, C, nullptr, StatSrc(), 0, rows, t.spatial ? HW : T * HW, t.gn, 1e-6f, 0, DST_LINEAR, H, W, lin, nullptr);
    linear(lin, rows, C, t.proj_in, EPI_BF16, y, I, nullptr);

    AttnArgs self;
    self.q = qkv; self.k = qkv + I; self.v = qkv + 2 * I; self.ldq = 3 * I; self.ldkv = 3 * I;
    self.out = att; self.ldo = I; self.heads = t.heads; self.scale = 0.125f;
    if (t.spatial) {
        self.n_outer = NF; self.Lq = HW; self.Lk = HW;
        self.qmap.div = 1; self.qmap.mul_a = HW; self.qmap.mul_b = 0; self.qmap.step = 1;
    } else {
        self.n_outer = B * HW; self.Lq = T; self.Lk = T;
        self.qmap.div = HW; self.qmap.mul_a = T * HW; self.qmap.mul_b = 1; self.qmap.step = HW;
    }
    self.kvmap = self.qmap;
    self.q_rows = rows; self.kv_rows = rows;
    const double attn_flops = 4.0 * (double)self.n_outer * t.heads * self.Lq * self.Lk * 64.0;
    const bool spatial = t.spatial;
    // spatial sequences of >= 64 tokens run on the tcgen05 kernel (tensor maps built here, once per plan)
    const bool self_tc = spatial && attention_tc_supported(self);
    AttnTcLaunch self_l;
    if (self_tc && !dry && !rc) { int r = attention_tc_plan(self, &self_l); if (r) fail(r); }
    auto run_self = [=](cudaStream_t st) {
        return self_tc ? attention_tc_run(self_l, st) : spatial ? attention_flash(self, st) : attention_short(self, st);
    };

    // attn1: self-attention
    lnorm(y, t.ln1, rows, I, lin);
    linear(lin, rows, I, t.qkv1, EPI_BF16, qkv, 3 * I, nullptr);
    push(run_self, 1, "attn_self", attn_flops); plan->flops += attn_flops;
    linear(att, rows, I, t.out1, EPI_BF16, y, I, nullptr, 0, nullptr, 0, y);
    // attn2: cross-attention on the text context (spatial) or a second self-attention (temporal_selfatt_only, attention.py:313-314)
    lnorm(y, t.ln2, rows, I, lin);
    if (t.spatial) {
        linear(lin, rows, I, t.q2, EPI_BF16, qkv, I, nullptr);
        AttnArgs ca;
        ca.q = qkv; ca.ldq = I; ca.qmap = self.qmap;
        ca.k = kv; ca.v = kv + I; ca.ldkv = kv_ld;
        ca.kvmap.div = T; ca.kvmap.mul_a = Lmax; ca.kvmap.mul_b = 0; ca.kvmap.step = 1;     // frame o uses the context of sample o / T
        ca.out = att; ca.ldo = I; ca.n_outer = NF; ca.heads = t.heads; ca.Lq = HW; ca.Lk = Lmax;
        ca.kv_len = plan->d_ctxlen; ca.kv_len_div = T; ca.scale = 0.125f;
        ca.q_rows = rows; ca.kv_rows = (int64_t)B * Lmax;
        const bool ca_tc = attention_tc_supported(ca);
        AttnTcLaunch ca_l;
        if (ca_tc && !dry && !rc) { int r = attention_tc_plan(ca, &ca_l); if (r) fail(r); }
        push([=](cudaStream_t st) { return ca_tc ? attention_tc_run(ca_l, st) : attention_flash(ca, st); }, 1, "attn_cross", 4.0 * (double)NF * t.heads * HW * Lmax * 64.0);
        plan->flops += 4.0 * (double)NF * t.heads * HW * Lmax * 64.0;
    } else {
        linear(lin, rows, I, t.qkv2, EPI_BF16, qkv, 3 * I, nullptr);
        push(run_self, 1, "attn_self", attn_flops); plan->flops += attn_flops;
    }
    linear(att, rows, I, t.out2, EPI_BF16, y, I, nullptr, 0, nullptr, 0, y);
    // feed-forward: Linear(C -> 8C) with the GEGLU gate fused into the GEMM epilogue, Linear(4C -> C) + residual
    lnorm(y, t.ln3, rows, I, lin);
    linear(lin, rows, I, t.ff1, EPI_GEGLU, ffh, 4 * I, nullptr);
    linear(ffh, rows, 4 * I, t.ff2, EPI_BF16, y, I, nullptr, 0, nullptr, 0, y);
    // proj_out + x_in (reads the 16-bit stream directly)
    x.st = stat_alloc(C, HW);
    linear(y, rows, I, t.proj_out, EPI_F32, x.p, C, x.p, 0, &x.st, HW);
}


struct IntPack {
    int ts[128];
    int fps[128];
    int ctxlen[16];
};

__global__ void set_ints_kernel(const __grid_constant__ IntPack P, int nf, int nb, int* __restrict__ ts, int* __restrict__ fps,
                                int* __restrict__ ctxlen) {
    const int i = threadIdx.x;
    if (i < nf) { ts[i] = P.ts[i]; fps[i] = P.fps[i]; }
    if (i < nb) ctxlen[i] = P.ctxlen[i];
}

// UNetModel.forward (openaimodel3d.py:534-578)
void Builder::build() {
    moca_unet* n = net;
    const int mc = n->mc, te = 4 * mc;
    const int64_t rows0 = (int64_t)NF * H0 * W0;
    // ---- fixed input staging (the graph reads these; the caller's pointers are copied in before every launch)
    plan->x_in = (float*)palloc((size_t)B * n->in_ch * T * H0 * W0 * sizeof(float));
    plan->ctx_in = (float*)palloc((size_t)B * Lmax * n->ctx_dim * sizeof(float));
    plan->d_ts = (int*)palloc(128 * sizeof(int));
    plan->d_fps = (int*)palloc(128 * sizeof(int));
    plan->d_ctxlen = (int*)palloc(16 * sizeof(int));
    plan->out_tok = (float*)palloc((size_t)rows0 * n->out_ch * sizeof(float));
    op_t* te_bf = (op_t*)palloc((size_t)NF * mc * 2);
    op_t* fe_bf = (op_t*)palloc((size_t)NF * mc * 2);
    float* e1 = (float*)palloc((size_t)NF * te * 4);
    op_t* e1_bf = (op_t*)palloc((size_t)NF * te * 2);
    float* temb = (float*)palloc((size_t)NF * te * 4);
    float* femb = (float*)palloc((size_t)NF * te * 4);
    op_t* semb = (op_t*)palloc((size_t)NF * te * 2);
    float* embout = (float*)palloc((size_t)NF * n->n_emb * 4);
    op_t* ctx_bf = (op_t*)palloc((size_t)B * Lmax * n->ctx_dim * 2);
    op_t* kv = (op_t*)palloc((size_t)B * Lmax * (n->n_kv > 0 ? n->n_kv : 16) * 2);
    scope = "embed";
    if (!dry) {
        double* stats = plan->stats;
        const size_t sb = plan->stats_bytes;
        raw_push([=](cudaStream_t st) { return cudaMemsetAsync(stats, 0, sb, st) == cudaSuccess ? MOCA_OK : MOCA_E_CUDA; }, "memset stats", 0);
    }
    plan->launches += 1;
    // ---- time / fps embeddings (openaimodel3d.py:536-543) and all 22 emb_layers in one GEMM (:166-172, 221-226)
    {
        const int* d_ts = plan->d_ts; const int* d_fps = plan->d_fps; const int nf = NF;
        push([=](cudaStream_t st) { return timestep_embedding_bf16(d_ts, nf, mc, te_bf, st); });
        push([=](cudaStream_t st) { return timestep_embedding_bf16(d_fps, nf, mc, fe_bf, st); });
        const int64_t ne = (int64_t)NF * te;
        linear(te_bf, NF, mc, n->time0, EPI_F32, e1, te, nullptr);
        push([=](cudaStream_t st) { return silu_add_bf16(e1, nullptr, ne, e1_bf, st); });
        linear(e1_bf, NF, te, n->time2, EPI_F32, temb, te, nullptr);
        linear(fe_bf, NF, mc, n->fps0, EPI_F32, e1, te, nullptr);
        push([=](cudaStream_t st) { return silu_add_bf16(e1, nullptr, ne, e1_bf, st); });
        linear(e1_bf, NF, te, n->fps2, EPI_F32, femb, te, nullptr);
        push([=](cudaStream_t st) { return silu_add_bf16(temb, femb, ne, semb, st); });       // SiLU(emb), emb = t_emb + fps_emb
        linear(semb, NF, te, n->emb_all, EPI_F32, embout, n->n_emb, nullptr);
    }
    // ---- cross-attention keys / values of every SpatialTransformer in one GEMM (context is the same for all frames)
    if (n->n_kv > 0) {
        const float* ctx_in = plan->ctx_in;
        const int64_t nc = (int64_t)B * Lmax * n->ctx_dim;
        push([=](cudaStream_t st) { return f32_to_bf16(ctx_in, nc, ctx_bf, st); });
        linear(ctx_bf, (int64_t)B * Lmax, n->ctx_dim, n->kv_all, EPI_BF16, kv, n->n_kv, nullptr, n->n_kv);
    }
    // ---- input conv
    Act cur;
    cur.C = mc; cur.H = H0; cur.W = W0;
    {
        op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (H0 + 2) * (W0 + 2) * 64);
        const float* x_in = plan->x_in;
        const int b = B, c = n->in_ch, t = T, h = H0, w = W0;
        const bool c4 = conv4d(H0, W0, 1);
        const PackedW& cw = n->inp[0].layers[0].conv;
        cur.p = (float*)palloc((size_t)rows0 * mc * 4);
        cur.st = stat_alloc(mc, H0 * W0);
        if (cw.taps == 1) {
            // patches gathered by the layout conversion: 81920 x 320 x 64 instead of nine taps of 60 zero channels each (85 -> 30 us)
            push([=](cudaStream_t st) {
                const int64_t total = (int64_t)b * t * h * w * 64;
                ncthw_patches_kernel<<<(int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16), 256, 0, st>>>(x_in, b, c, t, h, w, pad);
                return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
            }, 1, "stem patches");
            GemmDesc d;
            d.A = pad; d.a_rows = rows0; d.a_cols = 64; d.a_pitch = 64;
            d.Wt = cw.w; d.N = cw.N; d.ntaps = 1; d.Kp = cw.Kp;
            d.mode = GEMM_LINEAR; d.M = (int)rows0; d.epi = EPI_F32; d.out = cur.p; d.ldo = cw.N; d.bias = cw.b;
            d.chan_stats = (double*)cur.st.cs; d.stat_rows_per_frame = cur.st.rows; d.stat_sample_rows = T * H0 * W0;
            gemm(d, 2.0 * rows0 * cw.N * 9.0 * c);
        } else {
            push([=](cudaStream_t st) { return c4 ? ncthw_to_tokens(x_in, b, c, t, h, w, 64, pad, st) : ncthw_to_pad(x_in, b, c, t, h, w, 64, pad, st); });
            conv3x3(pad, H0, W0, 64, cw, 1, cur.p, nullptr, 0, nullptr, &cur.st);
        }
        tap("input_blocks.0.0", cur);
    }
    auto run_layers = [&](const BlockW& blk, const Act* skip) {
        for (size_t li = 0; li < blk.layers.size(); ++li) {
            const LayerW& L = blk.layers[li];
            if (L.kind == "conv_in") continue;
            if (L.kind == "res") {
                cur = res_block(L, cur, li == 0 ? skip : nullptr, embout + L.res.emb_off, n->n_emb);
            } else if (L.kind == "st" || L.kind == "tt") {
                transformer(L, cur, kv ? kv + L.tf.kv_off : nullptr, n->n_kv);
            } else if (L.kind == "down") {
                scope = L.name;
                const int64_t rows = (int64_t)NF * cur.H * cur.W;
                op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (cur.H + 2) * (cur.W + 2) * cur.C);
                convert(cur.p, cur.C, rows, conv_dst(cur.H, cur.W, 2), cur.H, cur.W, pad);
                Act o; o.C = L.cout; o.H = cur.H / 2; o.W = cur.W / 2;
                o.p = (float*)palloc((size_t)NF * o.H * o.W * o.C * 4);
                o.st = stat_alloc(o.C, o.H * o.W);
                conv3x3(pad, cur.H, cur.W, cur.C, L.conv, 2, o.p, nullptr, 0, nullptr, &o.st);
                cur = o;
            } else if (L.kind == "up") {
                scope = L.name;
                const int64_t rows = (int64_t)NF * cur.H * cur.W;
                const int H2 = cur.H * 2, W2 = cur.W * 2;
                op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (H2 + 2) * (W2 + 2) * cur.C);
                convert(cur.p, cur.C, rows, conv4d(H2, W2, 1) ? DST_UP2 : DST_SPATIAL_PAD_UP2, cur.H, cur.W, pad);
                Act o; o.C = L.cout; o.H = H2; o.W = W2;
                o.p = (float*)palloc((size_t)NF * H2 * W2 * o.C * 4);
                o.st = stat_alloc(o.C, H2 * W2);
                conv3x3(pad, H2, W2, cur.C, L.conv, 1, o.p, nullptr, 0, nullptr, &o.st);
                cur = o;
            }
            tap(L.name, cur);
        }
    };
    std::vector<Act> hs;
    for (size_t i = 0; i < n->inp.size(); ++i) {
        if (i > 0) run_layers(n->inp[i], nullptr);
        if (i == 0) run_layers(n->init_attn, nullptr);        // openaimodel3d.py:559-560
        hs.push_back(cur);
    }
    run_layers(n->mid, nullptr);
    for (size_t i = 0; i < n->outb.size(); ++i) {
        Act skip = hs.back();
        hs.pop_back();
        if (skip.H != cur.H || skip.W != cur.W) { set_error("plan: skip tensor geometry mismatch at output block %zu", i); fail(MOCA_E_SHAPE); }
        run_layers(n->outb[i], &skip);                        // h = cat([h, hs.pop()], 1) is virtual: two-source GroupNorm
    }
    // ---- out: GN -> SiLU -> conv3x3 (openaimodel3d.py:528-532, 574)
    scope = "out";
    {
        const int64_t rows = rows0;
        op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (H0 + 2) * (W0 + 2) * cur.C);
        gnorm(cur.p, cur.st, cur.C, nullptr, StatSrc(), 0, rows, H0 * W0, n->gn_out, 1e-5f, 1, conv_dst(H0, W0, 1), H0, W0, pad, nullptr);
        conv3x3(pad, H0, W0, cur.C, n->conv_out, 1, plan->out_tok, nullptr, 0, nullptr, nullptr);
    }
}


// ---- one module of the UNet on its own (the per-module entry points of the boundary) --------------------------
// {sum, sumsq} doubles per (frame, channel) of a token matrix that no GEMM epilogue produced (the module's input)
__global__ void token_stats_kernel(const float* __restrict__ x, int C, int HW, double* __restrict__ cs) {
    __shared__ double rs[8][32], rq[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = blockIdx.y, c = blockIdx.x * 32 + lane;
    double s = 0.0, q = 0.0;
    if (c < C)
        for (int r = warp; r < HW; r += 8) {
            const double v = (double)x[((size_t)f * HW + r) * C + c];
            s += v; q += v * v;
        }
    rs[warp][lane] = s; rq[warp][lane] = q;
    __syncthreads();
    if (warp == 0 && c < C) {
        for (int w = 1; w < 8; ++w) { s += rs[w][lane]; q += rq[w][lane]; }
        cs[((size_t)f * C + c) * 2] = s;
        cs[((size_t)f * C + c) * 2 + 1] = q;
    }
}

__global__ void ncthw_to_tokens_f32_kernel(const float* __restrict__ x, int B, int C, int T, int H, int W, float* __restrict__ tok) {
    const int64_t total = (int64_t)B * C * T * H * W;
    const int HW = H * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const int64_t r = i / C;                       // token row (b, t, hw)
        const int hw = (int)(r % HW), t = (int)((r / HW) % T), b = (int)(r / ((int64_t)HW * T));
        tok[i] = x[(((int64_t)b * C + c) * T + t) * HW + hw];
    }
}

// ResBlock(+TemporalConvBlock) / SpatialTransformer / TemporalTransformer `L` applied to plan->x_in ([B, C, T, H, W] fp32);
// emb_in [NF, 4 mc] is the (time + fps) embedding a ResBlock receives, ctx_in the text context of a SpatialTransformer.
void Builder::build_layer(const LayerW& L) {
    moca_unet* n = net;
    const int mc = n->mc, te = 4 * mc;
    const bool is_res = L.kind == "res", is_st = L.kind == "st";
    const int Cin = is_res ? L.res.cin : L.tf.C;
    const int64_t rows = (int64_t)NF * H0 * W0;
    plan->x_in = (float*)palloc((size_t)rows * Cin * sizeof(float));
    plan->emb_in = (float*)palloc((size_t)NF * te * sizeof(float));
    plan->ctx_in = (float*)palloc((size_t)B * Lmax * n->ctx_dim * sizeof(float));
    plan->d_ctxlen = (int*)palloc(16 * sizeof(int));
    scope = L.name;
    if (!dry) {
        double* stats = plan->stats;
        const size_t sb = plan->stats_bytes;
        raw_push([=](cudaStream_t st) { return cudaMemsetAsync(stats, 0, sb, st) == cudaSuccess ? MOCA_OK : MOCA_E_CUDA; }, "memset stats", 0);
    }
    plan->launches += 1;
    Act cur;
    cur.C = Cin; cur.H = H0; cur.W = W0;
    cur.p = (float*)palloc((size_t)rows * Cin * sizeof(float));
    cur.st = stat_alloc(Cin, H0 * W0);
    {
        const float* x_in = plan->x_in;
        float* tok = cur.p;
        double* cs = (double*)cur.st.cs;
        const int b = B, c = Cin, t = T, h = H0, w = W0, nf = NF;
        push([=](cudaStream_t st) {
            const int64_t total = (int64_t)b * c * t * h * w;
            const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
            ncthw_to_tokens_f32_kernel<<<blocks, 256, 0, st>>>(x_in, b, c, t, h, w, tok);
            token_stats_kernel<<<dim3(cdiv(c, 32), nf), 256, 0, st>>>(tok, c, h * w, cs);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 2, "stage input");
    }
    if (is_res) {
        // emb_layers = SiLU -> Linear (openaimodel3d.py:166-172): this block's rows of the stacked projection
        const ResW& r = L.res;
        op_t* semb = (op_t*)palloc((size_t)NF * te * sizeof(op_t));
        float* embout = (float*)palloc((size_t)NF * r.cout * sizeof(float));
        const float* emb_in = plan->emb_in;
        const int64_t ne = (int64_t)NF * te;
        push([=](cudaStream_t st) { return silu_add_bf16(emb_in, nullptr, ne, semb, st); });
        PackedW slice = n->emb_all;
        slice.w = n->emb_all.w + (size_t)r.emb_off * n->emb_all.Kp; slice.b = n->emb_all.b + r.emb_off; slice.N = r.cout;
        linear(semb, NF, te, slice, EPI_F32, embout, r.cout, nullptr);
        cur = res_block(L, cur, nullptr, embout, r.cout);
    } else {
        const TfW& t = L.tf;
        op_t* kv = nullptr;
        if (is_st) {
            // this block's to_k | to_v rows of the stacked context projection (attention.py:81-82)
            const int I = t.inner;
            const float* ctx_in = plan->ctx_in;
            const int64_t nc = (int64_t)B * Lmax * n->ctx_dim;
            op_t* ctx16 = (op_t*)palloc((size_t)nc * sizeof(op_t));
            kv = (op_t*)palloc((size_t)B * Lmax * 2 * I * sizeof(op_t));
            push([=](cudaStream_t st) { return f32_to_bf16(ctx_in, nc, ctx16, st); });
            PackedW slice = n->kv_all;
            slice.w = n->kv_all.w + (size_t)t.kv_off * n->kv_all.Kp; slice.b = nullptr; slice.N = 2 * I;
            linear(ctx16, (int64_t)B * Lmax, n->ctx_dim, slice, EPI_BF16, kv, 2 * I, nullptr, 2 * I);
        }
        transformer(L, cur, kv, is_st ? 2 * t.inner : 0);
    }
    plan->out_tok = cur.p;
    plan->out_C = cur.C;
}

// ---- SD-VAE decoder (row N1): decode_first_stage_2DAE (ddpm3d.py:555-562) = post_quant_conv + Decoder.forward
// (autoencoder.py:104-107, ae_modules.py:536-578), one frame = one "frame" of the token matrix (NF = B, T = 1) ------------------
__global__ void vae_stage_kernel(const float* __restrict__ z, int B, int C, int HW, float inv_scale, const float* __restrict__ w,
                                 const float* __restrict__ b, float* __restrict__ tok) {
    // tok[(f, hw)][c] = post_quant_conv(z / scale_factor); columns C .. 63 zero (conv_in's K is padded to one 64-wide k-block)
    const int64_t total = (int64_t)B * HW * 64;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i & 63);
        const int64_t r = i >> 6;
        const int hw = (int)(r % HW), f = (int)(r / HW);
        float v = 0.f;
        if (c < C) {
            if (w) {
                v = b[c];
                for (int k = 0; k < C; ++k) v = fmaf(w[c * C + k], z[((int64_t)f * C + k) * HW + hw] * inv_scale, v);
            } else {
                v = z[((int64_t)f * C + c) * HW + hw] * inv_scale;
            }
        }
        tok[i] = v;
    }
}

// moments[f][c][hw] = quant_conv(tok)[c] (autoencoder.py:100): 1x1 conv of the encoder's conv_out tokens [B*HW, C], written NCHW
__global__ void vae_moments_kernel(const float* __restrict__ tok, int B, int C, int HW, const float* __restrict__ w,
                                   const float* __restrict__ b, float* __restrict__ out) {
    const int64_t total = (int64_t)B * C * HW;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int hw = (int)(i % HW);
        const int c = (int)((i / HW) % C);
        const int f = (int)(i / ((int64_t)HW * C));
        const float* t = tok + ((int64_t)f * HW + hw) * C;
        float v = b[c];
        for (int k = 0; k < C; ++k) v = fmaf(w[c * C + k], t[k], v);
        out[i] = v;
    }
}

// softmax(scale * S) over the rows of an fp32 matrix -> 16-bit P (AttnBlock, ae_modules.py:64-66); one block per row
__global__ void __launch_bounds__(256) softmax_rows_kernel(const float* __restrict__ S, int cols, int ld, float scale, op_t* __restrict__ Pm) {
    __shared__ float red[8];
    const float* row = S + (size_t)blockIdx.x * ld;
    op_t* out = Pm + (size_t)blockIdx.x * ld;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float mx = -INFINITY;
    for (int c = threadIdx.x; c < cols; c += 256) mx = fmaxf(mx, row[c]);
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    for (int w = 1; w < 8; ++w) mx = fmaxf(mx, red[w]);
    __syncthreads();
    float sum = 0.f;
    for (int c = threadIdx.x; c < cols; c += 256) sum += __expf((row[c] - mx) * scale);
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = 0.f;
    for (int w = 0; w < 8; ++w) sum += red[w];
    const float inv = 1.0f / sum;
    for (int c = threadIdx.x; c < cols; c += 256) out[c] = float2op(__expf((row[c] - mx) * scale) * inv);
}

// ResnetBlock.forward with temb = None (ae_modules.py:189-210)
Act Builder::vae_res(const VaeResW& r, const Act& x, const std::string& name) {
    scope = name;
    const int H = x.H, W = x.W, HW = H * W;
    const int64_t rows = (int64_t)NF * HW;
    if (x.C != r.cin) { set_error("plan: %s expects %d input channels, got %d", name.c_str(), r.cin, x.C); fail(MOCA_E_SHAPE); }
    const size_t pad_elems = (size_t)NF * (H + 2) * (W + 2);
    op_t* pad = scratch<op_t>(S_PAD, pad_elems * (r.cin > r.cout ? r.cin : r.cout));
    op_t* raw = r.has_nin ? scratch<op_t>(S_RAW, (size_t)rows * r.cin) : nullptr;
    op_t* t1 = scratch<op_t>(S_F32A, (size_t)rows * r.cout);
    Act out;
    out.C = r.cout; out.H = H; out.W = W;
    out.p = (float*)palloc((size_t)rows * r.cout * sizeof(float));
    gnorm(x.p, x.st, x.C, nullptr, StatSrc(), 0, rows, HW, r.n1, 1e-6f, 1, conv_dst(H, W, 1), H, W, pad, raw);
    const StatSrc st1 = stat_alloc(r.cout, HW);
    conv3x3(pad, H, W, r.cin, r.conv1, 1, t1, nullptr, 0, nullptr, &st1, 1);
    gnorm(t1, st1, r.cout, nullptr, StatSrc(), 0, rows, HW, r.n2, 1e-6f, 1, conv_dst(H, W, 1), H, W, pad, nullptr, 1);
    out.st = stat_alloc(r.cout, HW);
    if (r.has_nin) {
        linear(raw, rows, r.cin, r.nin, EPI_F32, out.p, r.cout, nullptr);
        conv3x3(pad, H, W, r.cout, r.conv2, 1, out.p, nullptr, 0, out.p, &out.st);
    } else {
        conv3x3(pad, H, W, r.cout, r.conv2, 1, out.p, nullptr, 0, x.p, &out.st);
    }
    return out;
}

// AttnBlock.forward (ae_modules.py:52-78): one head of C channels over the H*W tokens of a frame.  S = q k^T and O = P v run on the
// GEMM core with the keys / the transposed values in the weight operand's place; v's bias is added after the product (the rows of P
// sum to one); in place on x.
void Builder::vae_attn(Act& x) {
    const VaeW& v = net->vae;
    scope = v.encoder ? "encoder.mid.attn_1" : "decoder.mid.attn_1";
    const int C = x.C, HW = x.H * x.W;
    const int64_t rows = (int64_t)NF * HW;
    if (HW % 64 != 0 || C % 64 != 0) { set_error("plan: the VAE attention needs H*W (%d) and C (%d) to be multiples of 64", HW, C); fail(MOCA_E_SHAPE); }
    op_t* xn = scratch<op_t>(S_LIN, (size_t)rows * C);
    op_t* q = scratch<op_t>(S_QKV, (size_t)rows * C * 2);
    op_t* k = dry ? nullptr : q + (size_t)rows * C;
    op_t* vT = scratch<op_t>(S_ATT, (size_t)C * HW);
    float* Sm = scratch<float>(S_Y, (size_t)HW * HW);
    op_t* Pm = scratch<op_t>(S_FFH, (size_t)HW * HW);
    op_t* att = scratch<op_t>(S_RAW, (size_t)rows * C);
    gnorm(x.p, x.st, C, nullptr, StatSrc(), 0, rows, HW, v.attn_norm, 1e-6f, 0, DST_LINEAR, x.H, x.W, xn, nullptr);
    linear(xn, rows, C, v.aq, EPI_BF16, q, C, nullptr);
    linear(xn, rows, C, v.ak, EPI_BF16, k, C, nullptr);
    const float scale = 1.0f / sqrtf((float)C);
    for (int f = 0; f < NF; ++f) {
        const op_t* xf = dry ? nullptr : xn + (size_t)f * HW * C;
        const op_t* qf = dry ? nullptr : q + (size_t)f * HW * C;
        const op_t* kf = dry ? nullptr : k + (size_t)f * HW * C;
        PackedW wx; wx.w = (op_t*)xf; wx.N = HW; wx.K = C; wx.Kp = C;               // v^T = W_v x^T (no bias here)
        linear(v.av.w, C, C, wx, EPI_BF16, vT, HW, nullptr);
        PackedW wk; wk.w = (op_t*)kf; wk.N = HW; wk.K = C; wk.Kp = C;
        linear(qf, HW, C, wk, EPI_F32, Sm, HW, nullptr);
        push([=](cudaStream_t st) {
            softmax_rows_kernel<<<HW, 256, 0, st>>>(Sm, HW, HW, scale, Pm);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 1, "vae softmax");
        PackedW wv; wv.w = vT; wv.b = v.av.b; wv.N = C; wv.K = HW; wv.Kp = HW;
        linear(Pm, HW, HW, wv, EPI_BF16, dry ? nullptr : att + (size_t)f * HW * C, C, nullptr);
    }
    x.st = stat_alloc(C, HW);
    linear(att, rows, C, v.aproj, EPI_F32, x.p, C, x.p, 0, &x.st, HW);
}

void Builder::build_vae(float inv_scale) {
    moca_unet* n = net;
    const VaeW& v = n->vae;
    const int L = (int)v.mult.size();
    const int64_t rows0 = (int64_t)NF * H0 * W0;
    plan->x_in = (float*)palloc((size_t)rows0 * v.zc * sizeof(float));
    scope = "vae";
    if (!dry) {
        double* stats = plan->stats;
        const size_t sb = plan->stats_bytes;
        raw_push([=](cudaStream_t st) { return cudaMemsetAsync(stats, 0, sb, st) == cudaSuccess ? MOCA_OK : MOCA_E_CUDA; }, "memset stats", 0);
    }
    plan->launches += 1;
    Act h;
    h.H = H0; h.W = W0;
    {
        float* tok = (float*)palloc((size_t)rows0 * 64 * sizeof(float));
        op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (H0 + 2) * (W0 + 2) * 64);
        const float* z = plan->x_in;
        const float* pw = v.pq_w; const float* pb = v.pq_b;
        const int b = NF, c = v.zc, hw = H0 * W0;
        push([=](cudaStream_t st) {
            const int64_t total = (int64_t)b * hw * 64;
            const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
            vae_stage_kernel<<<blocks, 256, 0, st>>>(z, b, c, hw, inv_scale, pw, pb, tok);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 1, "vae stage");
        convert(tok, 64, rows0, conv_dst(H0, W0, 1), H0, W0, pad);
        h.C = v.ch * v.mult[L - 1];
        h.p = (float*)palloc((size_t)rows0 * h.C * sizeof(float));
        h.st = stat_alloc(h.C, H0 * W0);
        conv3x3(pad, H0, W0, 64, v.conv_in, 1, h.p, nullptr, 0, nullptr, &h.st);
    }
    h = vae_res(v.mid1, h, "decoder.mid.block_1");
    vae_attn(h);
    h = vae_res(v.mid2, h, "decoder.mid.block_2");
    for (int lvl = L - 1; lvl >= 0; --lvl) {
        for (size_t j = 0; j < v.up[lvl].size(); ++j)
            h = vae_res(v.up[lvl][j], h, "decoder.up." + std::to_string(lvl) + ".block." + std::to_string(j));
        if (lvl != 0) {
            scope = "decoder.up." + std::to_string(lvl) + ".upsample";
            const int64_t rows = (int64_t)NF * h.H * h.W;
            const int H2 = h.H * 2, W2 = h.W * 2;
            op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (H2 + 2) * (W2 + 2) * h.C);
            convert(h.p, h.C, rows, conv4d(H2, W2, 1) ? DST_UP2 : DST_SPATIAL_PAD_UP2, h.H, h.W, pad);
            Act o; o.C = h.C; o.H = H2; o.W = W2;
            o.p = (float*)palloc((size_t)NF * H2 * W2 * o.C * sizeof(float));
            o.st = stat_alloc(o.C, H2 * W2);
            conv3x3(pad, H2, W2, h.C, v.upconv[lvl], 1, o.p, nullptr, 0, nullptr, &o.st);
            h = o;
        }
    }
    scope = "decoder.out";
    {
        const int64_t rows = (int64_t)NF * h.H * h.W;
        op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (h.H + 2) * (h.W + 2) * h.C);
        gnorm(h.p, h.st, h.C, nullptr, StatSrc(), 0, rows, h.H * h.W, v.norm_out, 1e-6f, 1, conv_dst(h.H, h.W, 1), h.H, h.W, pad, nullptr);
        plan->out_tok = (float*)palloc((size_t)rows * v.out_ch * sizeof(float));
        plan->out_C = v.out_ch;
        conv3x3(pad, h.H, h.W, h.C, v.conv_out, 1, plan->out_tok, nullptr, 0, nullptr, nullptr);
    }
}

// Encoder.forward (ae_modules.py:430-463); out_tok = conv_out tokens [B*h*w, mom_ch], quant_conv is applied by moca_vae_encode
void Builder::build_vae_encoder() {
    moca_unet* n = net;
    const VaeW& v = n->vae;
    const int L = (int)v.mult.size();
    const int64_t rows0 = (int64_t)NF * H0 * W0;
    plan->x_in = (float*)palloc((size_t)rows0 * v.in_ch * sizeof(float));
    scope = "vae";
    if (!dry) {
        double* stats = plan->stats;
        const size_t sb = plan->stats_bytes;
        raw_push([=](cudaStream_t st) { return cudaMemsetAsync(stats, 0, sb, st) == cudaSuccess ? MOCA_OK : MOCA_E_CUDA; }, "memset stats", 0);
    }
    plan->launches += 1;
    Act h;
    h.H = H0; h.W = W0;
    {
        float* tok = (float*)palloc((size_t)rows0 * 64 * sizeof(float));
        op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (H0 + 2) * (W0 + 2) * 64);
        const float* x = plan->x_in;
        const int b = NF, c = v.in_ch, hw = H0 * W0;
        push([=](cudaStream_t st) {
            const int64_t total = (int64_t)b * hw * 64;
            const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
            vae_stage_kernel<<<blocks, 256, 0, st>>>(x, b, c, hw, 1.0f, nullptr, nullptr, tok);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 1, "vae stage");
        convert(tok, 64, rows0, conv_dst(H0, W0, 1), H0, W0, pad);
        h.C = v.ch;
        h.p = (float*)palloc((size_t)rows0 * h.C * sizeof(float));
        h.st = stat_alloc(h.C, H0 * W0);
        conv3x3(pad, H0, W0, 64, v.conv_in, 1, h.p, nullptr, 0, nullptr, &h.st);
    }
    for (int lvl = 0; lvl < L; ++lvl) {
        for (size_t j = 0; j < v.down[lvl].size(); ++j)
            h = vae_res(v.down[lvl][j], h, "encoder.down." + std::to_string(lvl) + ".block." + std::to_string(j));
        if (lvl != L - 1) {
            // Downsample (ae_modules.py:80-100): F.pad (0,1,0,1) + 3x3 stride-2 conv without padding = the stride-2 tensor-map conv
            // with its window starting AT the output pixel (no leading zero row / column); right / bottom zeros come from the TMA fill
            scope = "encoder.down." + std::to_string(lvl) + ".downsample";
            if ((h.H | h.W) & 1) { set_error("plan: odd image size %d x %d at the encoder's downsample %d", h.H, h.W, lvl); fail(MOCA_E_SHAPE); return; }
            const int64_t rows = (int64_t)NF * h.H * h.W;
            op_t* lin = scratch<op_t>(S_PAD, (size_t)rows * h.C);
            convert(h.p, h.C, rows, DST_LINEAR, h.H, h.W, lin);
            Act o; o.C = h.C; o.H = h.H / 2; o.W = h.W / 2;
            o.p = (float*)palloc((size_t)NF * o.H * o.W * o.C * sizeof(float));
            o.st = stat_alloc(o.C, o.H * o.W);
            conv3x3(lin, h.H, h.W, h.C, v.downconv[lvl], 2, o.p, nullptr, 0, nullptr, &o.st, 0, 0);
            h = o;
        }
    }
    h = vae_res(v.mid1, h, "encoder.mid.block_1");
    vae_attn(h);
    h = vae_res(v.mid2, h, "encoder.mid.block_2");
    scope = "encoder.out";
    {
        const int64_t rows = (int64_t)NF * h.H * h.W;
        op_t* pad = scratch<op_t>(S_PAD, (size_t)NF * (h.H + 2) * (h.W + 2) * h.C);
        gnorm(h.p, h.st, h.C, nullptr, StatSrc(), 0, rows, h.H * h.W, v.norm_out, 1e-6f, 1, conv_dst(h.H, h.W, 1), h.H, h.W, pad, nullptr);
        plan->out_tok = (float*)palloc((size_t)rows * v.mom_ch * sizeof(float));
        plan->out_C = v.mom_ch;
        conv3x3(pad, h.H, h.W, h.C, v.conv_out, 1, plan->out_tok, nullptr, 0, nullptr, nullptr);
    }
}

// ---- OpenCLIP text tower (row N4): FrozenOpenCLIPEmbedder.encode_with_transformer (condition.py:212-230) -----------------------
// x[r] = token_embedding[tokens[r]] + positional_embedding[r mod L] (condition.py:213-214)
__global__ void text_embed_kernel(const int* __restrict__ tokens, const float* __restrict__ tok_emb, const float* __restrict__ pos_emb,
                                  int rows, int L, int W, int vocab, float* __restrict__ x) {
    const int r = blockIdx.x;
    if (r >= rows) return;
    int t = tokens[r];
    t = t < 0 ? 0 : (t >= vocab ? vocab - 1 : t);              // ids are validated on the host; never read outside the table
    const float4* e = reinterpret_cast<const float4*>(tok_emb + (size_t)t * W);
    const float4* p = reinterpret_cast<const float4*>(pos_emb + (size_t)(r % L) * W);
    float4* o = reinterpret_cast<float4*>(x + (size_t)r * W);
    for (int i = threadIdx.x; i < W / 4; i += blockDim.x) {
        const float4 a = e[i], b = p[i];
        o[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    }
}

// nn.MultiheadAttention with the causal additive mask of the text tower (open_clip build_attention_mask: -inf above the diagonal),
// head dim 64, L <= 128 keys: blocks of a (sample, head) hold its K / V in shared memory, one warp per query row.
__global__ void __launch_bounds__(128) text_attn_kernel(const op_t* __restrict__ qkv, int L, int W, int heads, op_t* __restrict__ out) {
    extern __shared__ float sm[];
    float* Ks = sm;
    float* Vs = sm + (size_t)L * 65;
    float* Qs = Vs + (size_t)L * 65;                           // [4 warps][64]
    const int b = blockIdx.x / heads, h = blockIdx.x % heads;
    const op_t* base = qkv + (size_t)b * L * 3 * W + h * 64;
    for (int i = threadIdx.x; i < L * 64; i += blockDim.x) {
        const int j = i >> 6, d = i & 63;
        Ks[j * 65 + d] = op2float(base[(size_t)j * 3 * W + W + d]);
        Vs[j * 65 + d] = op2float(base[(size_t)j * 3 * W + 2 * W + d]);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* qs = Qs + warp * 64;
    // gridDim.y blocks share a (sample, head): query rows interleaved over them (K / V are re-read, 20 KB from L2)
    for (int qi = warp + 4 * blockIdx.y; qi < L; qi += 4 * gridDim.y) {
        qs[lane] = op2float(base[(size_t)qi * 3 * W + lane]) * 0.125f;
        qs[lane + 32] = op2float(base[(size_t)qi * 3 * W + lane + 32]) * 0.125f;
        __syncwarp();
        float sc[4], mx = -INFINITY;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int j = lane + 32 * c;
            float a = -INFINITY;
            if (j <= qi) {
                a = 0.f;
                for (int d = 0; d < 64; ++d) a = fmaf(qs[d], Ks[j * 65 + d], a);
            }
            sc[c] = a;
            mx = fmaxf(mx, a);
        }
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            sc[c] = (lane + 32 * c <= qi) ? __expf(sc[c] - mx) : 0.f;
            sum += sc[c];
        }
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (32 * c > qi) break;                              // warp-uniform
            const int jn = min(32, qi + 1 - 32 * c);
            for (int jj = 0; jj < jn; ++jj) {
                const float pj = __shfl_sync(0xffffffffu, sc[c], jj);
                const float* vr = Vs + (32 * c + jj) * 65;
                a0 = fmaf(pj, vr[lane], a0);
                a1 = fmaf(pj, vr[lane + 32], a1);
            }
        }
        const float inv = 1.0f / sum;
        op_t* o = out + ((size_t)b * L + qi) * W + h * 64;
        o[lane] = float2op(a0 * inv);
        o[lane + 32] = float2op(a1 * inv);
        __syncwarp();
    }
}

// nn.GELU() (exact erf form; open_clip's ViT-H-14 text config has quick_gelu = false): fp32 -> 16-bit GEMM operand
__global__ void gelu_op16_kernel(const float* __restrict__ x, int64_t n, op_t* __restrict__ y) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float g = x[i];
        y[i] = float2op(0.5f * g * (1.0f + erff(g * 0.70710678118654752f)));
    }
}

// nn.LayerNorm(W), eps 1e-5, fp32 out (ln_final, condition.py:218): one warp per row, two passes over registers-in-cache
__global__ void __launch_bounds__(128) layernorm_f32_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, int rows, int W, float* __restrict__ y) {
    const int r = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= rows) return;
    const float* xr = x + (size_t)r * W;
    float s = 0.f;
    for (int i = lane; i < W; i += 32) s += xr[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / (float)W;
    float q = 0.f;
    for (int i = lane; i < W; i += 32) { const float d = xr[i] - mean; q = fmaf(d, d, q); }
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q / (float)W + 1e-5f);
    for (int i = lane; i < W; i += 32) y[(size_t)r * W + i] = (xr[i] - mean) * rstd * gamma[i] + beta[i];
}

void Builder::build_text() {
    const TextW& t = net->text;
    const int L = t.ctx_len, W = t.width;
    const int rows = B * L;
    plan->d_ts = (int*)palloc((size_t)rows * sizeof(int));                 // token ids [B, L]
    float* x = (float*)palloc((size_t)rows * W * sizeof(float));           // fp32 residual stream
    plan->out_tok = (float*)palloc((size_t)rows * W * sizeof(float));
    plan->out_C = W;
    op_t* ln = scratch<op_t>(S_LIN, (size_t)rows * W);
    op_t* qkv = scratch<op_t>(S_QKV, (size_t)rows * 3 * W);
    op_t* att = scratch<op_t>(S_ATT, (size_t)rows * W);
    float* fc = scratch<float>(S_F32A, (size_t)rows * 4 * W);
    op_t* act = scratch<op_t>(S_FFH, (size_t)rows * 4 * W);
    scope = "text";
    {
        const int* tok = plan->d_ts;
        const float* te = t.tok_emb; const float* pe = t.pos_emb;
        const int vocab = t.vocab;
        push([=](cudaStream_t st) {
            text_embed_kernel<<<rows, 128, 0, st>>>(tok, te, pe, rows, L, W, vocab, x);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 1, "embed");
    }
    const int heads = t.heads, nb = B;
    const size_t attn_smem = ((size_t)2 * L * 65 + 4 * 64) * sizeof(float);
    for (int i = 0; i < t.used; ++i) {
        const TextLayerW& b = t.blocks[i];
        scope = "model.transformer.resblocks." + std::to_string(i);
        // x = x + attn(ln_1(x)); x = x + mlp(ln_2(x))   (open_clip ResidualAttentionBlock.forward)
        push([=](cudaStream_t st) { return layernorm_op16(x, b.ln1.g, b.ln1.b, rows, W, ln, st); }, 1, "layernorm");
        linear(ln, rows, W, b.in_proj, EPI_BF16, qkv, 3 * W, nullptr);
        push([=](cudaStream_t st) {
            text_attn_kernel<<<dim3(nb * heads, 4), 128, attn_smem, st>>>(qkv, L, W, heads, att);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 1, "causal attention");
        linear(att, rows, W, b.out_proj, EPI_F32, x, W, x);
        push([=](cudaStream_t st) { return layernorm_op16(x, b.ln2.g, b.ln2.b, rows, W, ln, st); }, 1, "layernorm");
        linear(ln, rows, W, b.c_fc, EPI_F32, fc, 4 * W, nullptr);
        push([=](cudaStream_t st) {
            const int64_t n = (int64_t)rows * 4 * W;
            gelu_op16_kernel<<<(int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8), 256, 0, st>>>(fc, n, act);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 1, "gelu");
        linear(act, rows, 4 * W, b.c_proj, EPI_F32, x, W, x);
    }
    scope = "model.ln_final";
    {
        const float* g = t.ln_final.g; const float* bb = t.ln_final.b;
        float* y = plan->out_tok;
        push([=](cudaStream_t st) {
            layernorm_f32_kernel<<<(rows + 3) / 4, 128, 0, st>>>(x, g, bb, rows, W, y);
            return cudaGetLastError() == cudaSuccess ? MOCA_OK : MOCA_E_CUDA;
        }, 1, "layernorm f32");
    }
}

}  // namespace moca

using namespace moca;

// bounded plan cache: a plan owns its workspace (1.5 GB at full size) and CUDA graph, so varying prompt lengths, batch sizes or
// per-module calls must not accumulate them.  Least recently used plans go first; the one in use stays.
static void evict_plans(moca_unet* n) {
    static const size_t max_plans = [] { const char* e = getenv("MOCA_MAX_PLANS"); return (size_t)(e && atoi(e) > 0 ? atoi(e) : 4); }();
    while (n->plans.size() >= max_plans) {
        auto victim = n->plans.end();
        for (auto it2 = n->plans.begin(); it2 != n->plans.end(); ++it2)
            if (it2->second.get() != n->last && (victim == n->plans.end() || it2->second->last_use < victim->second->last_use)) victim = it2;
        if (victim == n->plans.end()) break;
        cudaDeviceSynchronize();                            // its graph / kernels may still be running
        n->plans.erase(victim);
    }
}

// dry pass (sizes) -> allocation -> real pass; `build` runs the same Builder method both times
template <typename F>
static int make_plan(moca_unet* n, const std::vector<int>& key, int B, int T, int H, int W, int Lmax, bool taps, F build, Plan** out) {
    auto plan = std::make_shared<Plan>();
    plan->B = B; plan->T = T; plan->H = H; plan->W = W; plan->Lmax = Lmax; plan->with_taps = taps;
    Builder dry;
    dry.net = n; dry.plan = plan.get(); dry.dry = true; dry.B = B; dry.T = T; dry.H0 = H; dry.W0 = W; dry.NF = B * T; dry.Lmax = Lmax;
    build(dry);
    if (dry.rc) return dry.rc;
    Builder real = dry;
    real.dry = false; real.cs_doubles = 0;
    plan->launches = 0; plan->flops = 0;
    plan->stats_bytes = dry.cs_doubles * sizeof(double);
    plan->stats = (double*)real.palloc(plan->stats_bytes);
    plan->ab_table = (float2*)real.palloc((size_t)B * T * 3072 * sizeof(float2));      // gn_apply supports C <= 3072
    for (int i = 0; i < S_COUNT; ++i) {
        real.sptr[i] = real.palloc(dry.need[i] + 256);
        plan->workspace_bytes += dry.need[i];
    }
    if (real.rc) return real.rc;
    build(real);
    if (real.rc) return real.rc;
    evict_plans(n);
    plan->last_use = ++n->use_clock;
    n->plans[key] = plan;
    *out = plan.get();
    return MOCA_OK;
}

static int get_plan(moca_unet* n, int B, int T, int H, int W, int Lmax, Plan** out) {
    std::vector<int> key{B, T, H, W, Lmax, n->taps ? 1 : 0};
    auto it = n->plans.find(key);
    if (it != n->plans.end()) { it->second->last_use = ++n->use_clock; *out = it->second.get(); return MOCA_OK; }
    int down = 1;
    for (size_t i = 1; i < n->mult.size(); ++i) down *= 2;
    MOCA_REQUIRE(H % down == 0 && W % down == 0, MOCA_E_SHAPE, "moca_unet_forward: H=%d W=%d must be multiples of %d", H, W, down);
    MOCA_REQUIRE(B * T <= 128 && B <= 16 && T <= 16, MOCA_E_SHAPE, "moca_unet_forward: B=%d T=%d unsupported (B*T <= 128, T <= 16)", B, T);
    return make_plan(n, key, B, T, H, W, Lmax, n->taps, [](Builder& b) { b.build(); }, out);
}

// launch a plan's ops: as one CUDA graph (captured on first use) unless taps are collected or MOCA_NO_GRAPH is set
static int run_plan(Plan* plan, cudaStream_t st, const char* who) {
    const bool use_graph = !plan->with_taps && getenv("MOCA_NO_GRAPH") == nullptr;
    if (!use_graph) {
        for (auto& op : plan->ops) {
            int rc = op(st);
            if (rc) return rc;
        }
        return MOCA_OK;
    }
    if (!plan->gexec) {
        // capture on a private stream (the caller's may be the legacy default stream, which cannot be captured)
        cudaGraph_t graph = nullptr;
        cudaStream_t cs = nullptr;
        MOCA_CHECK_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
        MOCA_CHECK_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
        int orc = MOCA_OK;
        for (auto& op : plan->ops) {
            orc = op(cs);
            if (orc) break;
        }
        cudaError_t ce = cudaStreamEndCapture(cs, &graph);
        cudaStreamDestroy(cs);
        if (orc) { if (graph) cudaGraphDestroy(graph); return orc; }
        MOCA_REQUIRE(ce == cudaSuccess, MOCA_E_CUDA, "%s: stream capture failed: %s", who, cudaGetErrorString(ce));
        ce = cudaGraphInstantiate(&plan->gexec, graph, 0);
        cudaGraphDestroy(graph);
        MOCA_REQUIRE(ce == cudaSuccess, MOCA_E_CUDA, "%s: graph instantiate failed: %s", who, cudaGetErrorString(ce));
    }
    MOCA_CHECK_CUDA(cudaGraphLaunch(plan->gexec, st));
    return MOCA_OK;
}

// AutoencoderKL.decode of B latent frames (autoencoder.py:104-107) with the 1 / scale_factor of decode_first_stage_2DAE
// (ddpm3d.py:555-562) folded into the staging kernel: z [B, zc, H, W] fp32 -> image [B, out_ch, H * 2^(L-1), W * 2^(L-1)] fp32
extern "C" int32_t moca_vae_decode(moca_unet* n, const float* z, int32_t B, int32_t H, int32_t W, float scale_factor, float* image,
                                   void* stream) {
    MOCA_REQUIRE(n && z && image, MOCA_E_ARG, "moca_vae_decode: NULL argument");
    MOCA_REQUIRE(n->vae.present && !n->vae.encoder, MOCA_E_STATE, "moca_vae_decode: handle was not made by moca_vae_create");
    MOCA_REQUIRE(n->finalized, MOCA_E_STATE, "moca_vae_decode: call moca_unet_finalize after loading the weights");
    MOCA_REQUIRE(B > 0 && B <= 16 && H > 0 && W > 0 && scale_factor != 0.f, MOCA_E_SHAPE, "moca_vae_decode: bad shape B=%d H=%d W=%d", B, H, W);
    const int up = 1 << ((int)n->vae.mult.size() - 1);
    MOCA_REQUIRE((int64_t)B * (H * up + 2) * (W * up + 2) < (1ll << 31) / 64, MOCA_E_SHAPE, "moca_vae_decode: B=%d frames of %d x %d are too many rows", B, H * up, W * up);
    cudaStream_t st = (cudaStream_t)stream;
    int sbits;
    memcpy(&sbits, &scale_factor, sizeof(int));
    std::vector<int> key{B, 1, H, W, -1, sbits};
    Plan* plan = nullptr;
    auto it = n->plans.find(key);
    if (it != n->plans.end()) { it->second->last_use = ++n->use_clock; plan = it->second.get(); }
    else {
        const float inv = 1.0f / scale_factor;
        int rc = make_plan(n, key, B, 1, H, W, 0, false, [inv](Builder& b) { b.build_vae(inv); }, &plan);
        if (rc) return rc;
    }
    n->last = plan;
    MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->x_in, z, (size_t)B * n->vae.zc * H * W * sizeof(float), cudaMemcpyDeviceToDevice, st));
    int rc = run_plan(plan, st, "moca_vae_decode");
    if (rc) return rc;
    return tokens_to_ncthw(plan->out_tok, n->vae.out_ch, B, n->vae.out_ch, 1, H * up, W * up, image, st);
}

// FrozenOpenCLIPEmbedder.encode_with_transformer (condition.py:212-219): tokens DEVICE int32 [B, context_length] (open_clip.tokenize
// output; the BPE tokenizer itself is host string work and stays with the caller) -> z [B, context_length, width] fp32
extern "C" int32_t moca_text_encode(moca_unet* n, const int32_t* tokens, int32_t B, float* z, void* stream) {
    MOCA_REQUIRE(n && tokens && z, MOCA_E_ARG, "moca_text_encode: NULL argument");
    MOCA_REQUIRE(n->text.present, MOCA_E_STATE, "moca_text_encode: handle was not made by moca_text_create");
    MOCA_REQUIRE(n->finalized, MOCA_E_STATE, "moca_text_encode: call moca_unet_finalize after loading the weights");
    MOCA_REQUIRE(B > 0 && B <= 64, MOCA_E_SHAPE, "moca_text_encode: B=%d outside 1..64", B);
    cudaStream_t st = (cudaStream_t)stream;
    const int L = n->text.ctx_len, W = n->text.width;
    std::vector<int> key{B, 1, L, 1, -3, 0};
    Plan* plan = nullptr;
    auto it = n->plans.find(key);
    if (it != n->plans.end()) { it->second->last_use = ++n->use_clock; plan = it->second.get(); }
    else {
        int rc = make_plan(n, key, B, 1, L, 1, 0, false, [](Builder& b) { b.build_text(); }, &plan);
        if (rc) return rc;
        MOCA_CHECK_CUDA(cudaFuncSetAttribute(text_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024));
    }
    n->last = plan;
    MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->d_ts, tokens, (size_t)B * L * sizeof(int), cudaMemcpyDeviceToDevice, st));
    int rc = run_plan(plan, st, "moca_text_encode");
    if (rc) return rc;
    MOCA_CHECK_CUDA(cudaMemcpyAsync(z, plan->out_tok, (size_t)B * L * W * sizeof(float), cudaMemcpyDeviceToDevice, st));
    return MOCA_OK;
}

// AutoencoderKL.encode up to the posterior's parameters (autoencoder.py:97-102): x [B, in_ch, H, W] fp32 image in [-1, 1] ->
// moments [B, 2 zc, H / 2^(L-1), W / 2^(L-1)] fp32 (mean | logvar, what DiagonalGaussianDistribution is built from)
extern "C" int32_t moca_vae_encode(moca_unet* n, const float* x, int32_t B, int32_t H, int32_t W, float* moments, void* stream) {
    MOCA_REQUIRE(n && x && moments, MOCA_E_ARG, "moca_vae_encode: NULL argument");
    MOCA_REQUIRE(n->vae.present && n->vae.encoder, MOCA_E_STATE, "moca_vae_encode: handle was not made by moca_vae_encoder_create");
    MOCA_REQUIRE(n->finalized, MOCA_E_STATE, "moca_vae_encode: call moca_unet_finalize after loading the weights");
    const int down = 1 << ((int)n->vae.mult.size() - 1);
    MOCA_REQUIRE(B > 0 && B <= 16 && H > 0 && W > 0 && H % down == 0 && W % down == 0, MOCA_E_SHAPE,
                 "moca_vae_encode: bad shape B=%d H=%d W=%d (multiples of %d)", B, H, W, down);
    MOCA_REQUIRE((int64_t)B * (H + 2) * (W + 2) < (1ll << 31) / 64, MOCA_E_SHAPE, "moca_vae_encode: B=%d frames of %d x %d are too many rows", B, H, W);
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<int> key{B, 1, H, W, -2, 0};
    Plan* plan = nullptr;
    auto it = n->plans.find(key);
    if (it != n->plans.end()) { it->second->last_use = ++n->use_clock; plan = it->second.get(); }
    else {
        int rc = make_plan(n, key, B, 1, H, W, 0, false, [](Builder& b) { b.build_vae_encoder(); }, &plan);
        if (rc) return rc;
    }
    n->last = plan;
    MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->x_in, x, (size_t)B * n->vae.in_ch * H * W * sizeof(float), cudaMemcpyDeviceToDevice, st));
    int rc = run_plan(plan, st, "moca_vae_encode");
    if (rc) return rc;
    const int hw = (H / down) * (W / down), C = n->vae.mom_ch;
    const int64_t total = (int64_t)B * C * hw;
    vae_moments_kernel<<<(int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16), 256, 0, st>>>(plan->out_tok, B, C, hw, n->vae.q_w,
                                                                                                             n->vae.q_b, moments);
    MOCA_LAUNCH_CHECK();
    return MOCA_OK;
}

extern "C" int32_t moca_unet_forward(moca_unet* n, const float* x, const int32_t* timesteps, const float* context,
                                     const int32_t* ctx_len, int32_t Lmax, const int32_t* fps, int32_t B, int32_t T, int32_t H,
                                     int32_t W, float* y, void* stream) {
    MOCA_REQUIRE(n && x && timesteps && context && ctx_len && fps && y, MOCA_E_ARG, "moca_unet_forward: NULL argument");
    MOCA_REQUIRE(!n->vae.present && !n->text.present, MOCA_E_STATE, "moca_unet_forward: this handle is a VAE or a text tower (moca_vae_decode / moca_vae_encode / moca_text_encode)");
    MOCA_REQUIRE(n->finalized, MOCA_E_STATE, "moca_unet_forward: call moca_unet_finalize after loading the weights");
    MOCA_REQUIRE(B > 0 && T > 0 && H > 0 && W > 0 && Lmax > 0, MOCA_E_SHAPE, "moca_unet_forward: bad shape");
    // fps: int32[B], one entry per sample (the Python mirror validates the length: a shorter array would be read past its end)
    MOCA_REQUIRE((Lmax * n->ctx_dim) % 8 == 0, MOCA_E_SHAPE, "moca_unet_forward: context size");
    for (int b = 0; b < B; ++b)
        MOCA_REQUIRE(ctx_len[b] >= 1 && ctx_len[b] <= Lmax, MOCA_E_ARG, "moca_unet_forward: ctx_len[%d]=%d outside 1..%d", b, ctx_len[b], Lmax);
    cudaStream_t st = (cudaStream_t)stream;
    Plan* plan = nullptr;
    int rc = get_plan(n, B, T, H, W, Lmax, &plan);
    if (rc) return rc;
    n->last = plan;
    // inputs -> fixed staging (kernel parameters carry the small integer tables, so no pinned staging is needed)
    IntPack ip;
    memset(&ip, 0, sizeof(ip));
    for (int i = 0; i < B * T; ++i) { ip.ts[i] = timesteps[i]; ip.fps[i] = fps[i / T]; }
    for (int b = 0; b < B; ++b) ip.ctxlen[b] = ctx_len[b];
    set_ints_kernel<<<1, 128, 0, st>>>(ip, B * T, B, plan->d_ts, plan->d_fps, plan->d_ctxlen);
    MOCA_LAUNCH_CHECK();
    MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->x_in, x, (size_t)B * n->in_ch * T * H * W * sizeof(float), cudaMemcpyDeviceToDevice, st));
    MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->ctx_in, context, (size_t)B * Lmax * n->ctx_dim * sizeof(float), cudaMemcpyDeviceToDevice, st));
    rc = run_plan(plan, st, "moca_unet_forward");
    if (rc) return rc;
    return tokens_to_ncthw(plan->out_tok, n->out_ch, B, n->out_ch, T, H, W, y, st);
}

// ---- per-module entry points (ResBlock.forward openaimodel3d.py:208-234 incl. its TemporalConvBlock :269-276,
// SpatialTransformer.forward attention.py:262-278, TemporalTransformer.forward :331-373) ---------------------------------------
static const LayerW* find_layer(moca_unet* n, const char* name) {
    auto scan = [&](const BlockW& b) -> const LayerW* {
        for (const LayerW& L : b.layers)
            if (L.name == name) return &L;
        return nullptr;
    };
    for (const BlockW& b : n->inp) if (const LayerW* L = scan(b)) return L;
    if (const LayerW* L = scan(n->init_attn)) return L;
    if (const LayerW* L = scan(n->mid)) return L;
    for (const BlockW& b : n->outb) if (const LayerW* L = scan(b)) return L;
    return nullptr;
}

extern "C" int32_t moca_unet_layer_forward(moca_unet* n, const char* layer, const float* x, const float* emb, const float* context,
                                           const int32_t* ctx_len, int32_t Lmax, int32_t B, int32_t T, int32_t H, int32_t W, float* y,
                                           void* stream) {
    MOCA_REQUIRE(n && layer && x && y, MOCA_E_ARG, "moca_unet_layer_forward: NULL argument");
    MOCA_REQUIRE(n->finalized, MOCA_E_STATE, "moca_unet_layer_forward: call moca_unet_finalize after loading the weights");
    const LayerW* L = find_layer(n, layer);
    MOCA_REQUIRE(L && (L->kind == "res" || L->kind == "st" || L->kind == "tt"), MOCA_E_ARG,
                 "moca_unet_layer_forward: '%s' is not a ResBlock / SpatialTransformer / TemporalTransformer of this network", layer);
    MOCA_REQUIRE(B > 0 && T > 0 && H > 0 && W > 0 && B * T <= 128 && B <= 16 && T <= 16, MOCA_E_SHAPE, "moca_unet_layer_forward: bad shape");
    const bool is_res = L->kind == "res", is_st = L->kind == "st";
    MOCA_REQUIRE(!is_res || emb, MOCA_E_ARG, "moca_unet_layer_forward: a ResBlock needs the embedding [B*T, %d]", 4 * n->mc);
    MOCA_REQUIRE(!is_st || (context && ctx_len && Lmax > 0 && (Lmax * n->ctx_dim) % 8 == 0), MOCA_E_ARG,
                 "moca_unet_layer_forward: a SpatialTransformer needs the context and its lengths");
    if (!is_st) Lmax = 8;
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<int> key{B, T, H, W, Lmax, 2};
    for (const char* c = layer; *c; ++c) key.push_back((int)*c);
    Plan* plan = nullptr;
    auto it = n->plans.find(key);
    if (it != n->plans.end()) {
        it->second->last_use = ++n->use_clock;
        plan = it->second.get();
    } else {
        auto np = std::make_shared<Plan>();
        np->B = B; np->T = T; np->H = H; np->W = W; np->Lmax = Lmax;
        Builder dry;
        dry.net = n; dry.plan = np.get(); dry.dry = true; dry.B = B; dry.T = T; dry.H0 = H; dry.W0 = W; dry.NF = B * T; dry.Lmax = Lmax;
        dry.build_layer(*L);
        if (dry.rc) return dry.rc;
        Builder real = dry;
        real.dry = false; real.cs_doubles = 0;
        np->launches = 0; np->flops = 0;
        np->stats_bytes = dry.cs_doubles * sizeof(double);
        np->stats = (double*)real.palloc(np->stats_bytes);
        np->ab_table = (float2*)real.palloc((size_t)B * T * 3072 * sizeof(float2));
        for (int i = 0; i < S_COUNT; ++i) real.sptr[i] = real.palloc(dry.need[i] + 256);
        if (real.rc) return real.rc;
        real.build_layer(*L);
        if (real.rc) return real.rc;
        evict_plans(n);
        np->last_use = ++n->use_clock;
        n->plans[key] = np;
        plan = np.get();
    }
    const int Cin = is_res ? L->res.cin : L->tf.C;
    MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->x_in, x, (size_t)B * Cin * T * H * W * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (is_res) MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->emb_in, emb, (size_t)B * T * 4 * n->mc * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (is_st) {
        MOCA_CHECK_CUDA(cudaMemcpyAsync(plan->ctx_in, context, (size_t)B * Lmax * n->ctx_dim * sizeof(float), cudaMemcpyDeviceToDevice, st));
        IntPack ip;
        memset(&ip, 0, sizeof(ip));
        for (int b = 0; b < B; ++b) {
            MOCA_REQUIRE(ctx_len[b] >= 1 && ctx_len[b] <= Lmax, MOCA_E_ARG, "moca_unet_layer_forward: ctx_len[%d]=%d outside 1..%d", b, ctx_len[b], Lmax);
            ip.ctxlen[b] = ctx_len[b];
        }
        int* scratch_ints = plan->d_ctxlen;        // set_ints_kernel also writes 2 x nf ints: none here (nf = 0)
        set_ints_kernel<<<1, 128, 0, st>>>(ip, 0, B, scratch_ints, scratch_ints, plan->d_ctxlen);
        MOCA_LAUNCH_CHECK();
    }
    for (auto& op : plan->ops) {
        int rc = op(st);
        if (rc) return rc;
    }
    return tokens_to_ncthw(plan->out_tok, plan->out_C, B, plan->out_C, T, H, W, y, st);
}

extern "C" int32_t moca_unet_layer_channels(moca_unet* n, const char* layer, int32_t* c_in, int32_t* c_out) {
    MOCA_REQUIRE(n && layer && c_in && c_out, MOCA_E_ARG, "moca_unet_layer_channels: NULL argument");
    const LayerW* L = find_layer(n, layer);
    MOCA_REQUIRE(L, MOCA_E_ARG, "moca_unet_layer_channels: no layer '%s'", layer);
    if (L->kind == "res") { *c_in = L->res.cin; *c_out = L->res.cout; }
    else if (L->kind == "st" || L->kind == "tt") { *c_in = L->tf.C; *c_out = L->tf.C; }
    else { *c_in = L->cin; *c_out = L->cout; }
    return MOCA_OK;
}


extern "C" int32_t moca_unet_destroy(moca_unet* n) {
    if (!n) return MOCA_OK;
    n->plans.clear();
    for (void* p : n->allocs) cudaFree(p);
    delete n;
    return MOCA_OK;
}

extern "C" int32_t moca_unet_set_taps(moca_unet* n, int32_t enable) {
    MOCA_REQUIRE(n, MOCA_E_ARG, "moca_unet_set_taps: NULL");
    n->taps = enable != 0;
    return MOCA_OK;
}
extern "C" int32_t moca_unet_num_taps(moca_unet* n) { return (n && n->last) ? (int32_t)n->last->taps.size() : 0; }
extern "C" int32_t moca_unet_tap_info(moca_unet* n, int32_t i, char* name, int32_t name_cap, int64_t* rows, int64_t* channels) {
    MOCA_REQUIRE(n && n->last && i >= 0 && i < (int)n->last->taps.size() && name && rows && channels, MOCA_E_ARG, "moca_unet_tap_info: bad argument");
    const Tap& t = n->last->taps[i];
    snprintf(name, name_cap, "%s", t.name.c_str());
    *rows = t.rows;
    *channels = t.C;
    return MOCA_OK;
}
extern "C" int32_t moca_unet_tap_read(moca_unet* n, int32_t i, float* dst, void* stream) {
    MOCA_REQUIRE(n && n->last && i >= 0 && i < (int)n->last->taps.size() && dst, MOCA_E_ARG, "moca_unet_tap_read: bad argument");
    const Tap& t = n->last->taps[i];
    MOCA_CHECK_CUDA(cudaMemcpyAsync(dst, t.buf, (size_t)t.rows * t.C * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return MOCA_OK;
}
extern "C" int32_t moca_unet_launches(moca_unet* n) { return (n && n->last) ? n->last->launches + 4 : 0; }
extern "C" double moca_unet_flops(moca_unet* n) { return (n && n->last) ? n->last->flops : 0.0; }
extern "C" int64_t moca_unet_workspace_bytes(moca_unet* n) { return (n && n->last) ? (int64_t)n->last->workspace_bytes : 0; }
extern "C" int64_t moca_unet_weight_bytes(moca_unet* n) { return n ? (int64_t)n->weight_bytes : 0; }

// Per-op device timing of the last planned forward (inputs must have been staged by a moca_unet_forward call).
// Writes CSV "idx,us,gflop,label" lines into buf; returns the number of bytes needed.
extern "C" int64_t moca_unet_profile(moca_unet* n, int32_t iters, char* buf, int64_t cap, void* stream) {
    if (!n || !n->last || iters < 1) return -1;
    Plan* plan = n->last;
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::string out;
    for (size_t i = 0; i < plan->ops.size(); ++i) {
        plan->ops[i](st);                                 // warm
        cudaEventRecord(e0, st);
        for (int k = 0; k < iters; ++k) plan->ops[i](st);
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        char line[512];
        snprintf(line, sizeof(line), "%zu,%.3f,%.3f,%s\n", i, ms * 1e3 / iters, plan->op_flops[i] / 1e9, plan->labels[i].c_str());
        out += line;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (buf && cap > 0) {
        snprintf(buf, (size_t)cap, "%s", out.c_str());
    }
    return (int64_t)out.size() + 1;
}
