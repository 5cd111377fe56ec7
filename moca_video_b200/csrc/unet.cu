// See unet.cuh.  Part 1: architecture enumeration and weight packing (moca_unet_create / _load / _finalize).
#include "unet.cuh"

using namespace moca;

namespace moca {

static int round_up(int v, int m) { return (v + m - 1) / m * m; }

static void* dalloc(moca_unet* n, size_t bytes) {
    void* p = nullptr;
    if (n->dry) {                       // architecture listing only (CPU tests): never dereferenced
        n->weight_bytes += bytes;
        return (void*)(uintptr_t)0x1000;
    }
    if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) return nullptr;
    cudaMemset(p, 0, bytes ? bytes : 16);
    n->allocs.push_back(p);
    n->weight_bytes += bytes;
    return p;
}

static float* reg_f32(moca_unet* n, const std::string& name, int64_t numel, float* into = nullptr, int geglu_bn = 0) {
    float* p = into ? into : (float*)dalloc(n, numel * sizeof(float));
    LoadSlot s;
    s.kind = LoadSlot::COPY_F32;
    s.dst_f32 = p;
    s.numel = numel;
    s.geglu_bn = geglu_bn;
    n->slots[name] = s;
    return p;
}

static NormW reg_norm(moca_unet* n, const std::string& prefix, int C) {
    NormW w;
    w.g = reg_f32(n, prefix + ".weight", C);
    w.b = reg_f32(n, prefix + ".bias", C);
    return w;
}

static PackedW alloc_packed(moca_unet* n, int N, int K, int taps, bool bias) {
    PackedW w;
    w.N = N; w.K = K; w.taps = taps; w.Kp = round_up(K, 64);
    w.w = (op_t*)dalloc(n, (size_t)N * taps * w.Kp * sizeof(op_t));
    if (bias) w.b = (float*)dalloc(n, (size_t)N * sizeof(float));
    return w;
}

// registers state_dict tensor `name` ([rowsN, K, taps]) to be packed into rows [row0, row0+rowsN) of pw
static void reg_pack(moca_unet* n, const std::string& name, const PackedW& pw, int row0, int rowsN, int geglu_bn = 0) {
    LoadSlot s;
    s.kind = LoadSlot::PACK;
    s.dst = pw.w;
    s.N = rowsN; s.K = pw.K; s.taps = pw.taps; s.Kp = pw.Kp; s.row0 = row0; s.geglu_bn = geglu_bn;
    s.pitch = (int64_t)pw.taps * pw.Kp;
    s.numel = (int64_t)rowsN * pw.K * pw.taps;
    n->slots[name] = s;
}

static PackedW reg_linear(moca_unet* n, const std::string& prefix, int N, int K, int taps, bool bias, int geglu_bn = 0) {
    PackedW w = alloc_packed(n, N, K, taps, bias);
    reg_pack(n, prefix + ".weight", w, 0, N, geglu_bn);
    if (bias) reg_f32(n, prefix + ".bias", N, w.b, geglu_bn);
    return w;
}

static void build_res(moca_unet* n, LayerW& L, int te) {
    ResW& r = L.res;
    const std::string& p = L.name;
    r.cin = L.cin; r.cout = L.cout;
    r.gn1 = reg_norm(n, p + ".in_layers.0", r.cin);
    r.conv1 = reg_linear(n, p + ".in_layers.2", r.cout, r.cin, 9, true);
    r.emb_off = n->n_emb;
    n->n_emb += r.cout;
    r.gn2 = reg_norm(n, p + ".out_layers.0", r.cout);
    r.conv2 = reg_linear(n, p + ".out_layers.3", r.cout, r.cout, 9, true);
    r.has_skip = r.cin != r.cout;
    if (r.has_skip) r.skip = reg_linear(n, p + ".skip_connection", r.cout, r.cin, 1, true);
    for (int i = 0; i < 4; ++i) {
        const std::string q = p + ".temopral_conv.conv" + std::to_string(i + 1);
        r.tgn[i] = reg_norm(n, q + ".0", r.cout);
        r.tconv[i] = reg_linear(n, q + (i == 0 ? ".2" : ".3"), r.cout, r.cout, 3, true);
    }
    (void)te;
}

static void build_tf(moca_unet* n, LayerW& L, bool spatial, int heads, int inner, bool linear) {
    TfW& t = L.tf;
    const std::string& p = L.name;
    t.spatial = spatial; t.linear = linear; t.C = L.cin; t.inner = inner; t.heads = heads;
    t.gn = reg_norm(n, p + ".norm", t.C);
    t.proj_in = reg_linear(n, p + ".proj_in", inner, t.C, 1, true);
    t.proj_out = reg_linear(n, p + ".proj_out", t.C, inner, 1, true);
    const std::string b = p + ".transformer_blocks.0";
    t.ln1 = reg_norm(n, b + ".norm1", inner);
    t.ln2 = reg_norm(n, b + ".norm2", inner);
    t.ln3 = reg_norm(n, b + ".norm3", inner);
    t.qkv1 = alloc_packed(n, 3 * inner, inner, 1, false);
    reg_pack(n, b + ".attn1.to_q.weight", t.qkv1, 0, inner);
    reg_pack(n, b + ".attn1.to_k.weight", t.qkv1, inner, inner);
    reg_pack(n, b + ".attn1.to_v.weight", t.qkv1, 2 * inner, inner);
    t.out1 = reg_linear(n, b + ".attn1.to_out.0", inner, inner, 1, true);
    if (spatial) {
        t.q2 = reg_linear(n, b + ".attn2.to_q", inner, inner, 1, false);
        t.kv_off = n->n_kv;
        n->n_kv += 2 * inner;          // k rows then v rows, registered once kv_all is allocated
    } else {
        t.qkv2 = alloc_packed(n, 3 * inner, inner, 1, false);
        reg_pack(n, b + ".attn2.to_q.weight", t.qkv2, 0, inner);
        reg_pack(n, b + ".attn2.to_k.weight", t.qkv2, inner, inner);
        reg_pack(n, b + ".attn2.to_v.weight", t.qkv2, 2 * inner, inner);
    }
    t.out2 = reg_linear(n, b + ".attn2.to_out.0", inner, inner, 1, true);
    t.ff1 = reg_linear(n, b + ".ff.net.0.proj", 8 * inner, inner, 1, true, 256);
    t.ff2 = reg_linear(n, b + ".ff.net.2", inner, 4 * inner, 1, true);
}

static bool has(const std::vector<int>& v, int x) {
    for (int e : v)
        if (e == x) return true;
    return false;
}

// Enumerates blocks exactly like UNetModel.__init__ (openaimodel3d.py:373-526); same walk as moca_video_b200/arch.py.
static void build_arch(moca_unet* n) {
    const int mc = n->mc, dh = n->dh, te = 4 * mc;
    n->time0 = reg_linear(n, "time_embed.0", te, mc, 1, true);
    n->time2 = reg_linear(n, "time_embed.2", te, te, 1, true);
    n->fps0 = reg_linear(n, "fps_embedding.0", te, mc, 1, true);
    n->fps2 = reg_linear(n, "fps_embedding.2", te, te, 1, true);
    auto mk = [](const char* kind, const std::string& name, int cin, int cout) {
        LayerW L; L.kind = kind; L.name = name; L.cin = cin; L.cout = cout; return L;
    };
    {
        BlockW b;
        LayerW L = mk("conv_in", "input_blocks.0.0", n->in_ch, mc);
        // the stem's [mc, in_ch, 3, 3] weight read as a Linear over k = c * 9 + tap when the 9 * in_ch patch fits one 64-wide k-block
        // (unet_plan.cu gathers the patches while it converts the input layout), else as a 9-tap convolution
        if (n->in_ch * 9 <= 64) L.conv = reg_linear(n, L.name, mc, n->in_ch * 9, 1, true);
        else L.conv = reg_linear(n, L.name, mc, n->in_ch, 9, true);
        b.layers.push_back(L);
        n->inp.push_back(b);
    }
    {
        LayerW L = mk("tt", "init_attn.0", mc, mc);
        build_tf(n, L, false, n->init_heads, n->init_heads * dh, false);
        n->init_attn.layers.push_back(L);
    }
    std::vector<int> chans{mc};
    int ch = mc, ds = 1, bi = 1;
    for (size_t level = 0; level < n->mult.size(); ++level) {
        const int mult = n->mult[level];
        for (int r = 0; r < n->nres; ++r) {
            BlockW b;
            const std::string pre = "input_blocks." + std::to_string(bi);
            LayerW L = mk("res", pre + ".0", ch, mult * mc);
            build_res(n, L, te);
            b.layers.push_back(L);
            ch = mult * mc;
            if (has(n->attn_res, ds)) {
                LayerW S = mk("st", pre + ".1", ch, ch);
                build_tf(n, S, true, ch / dh, ch, true);
                b.layers.push_back(S);
                LayerW T = mk("tt", pre + ".2", ch, ch);
                build_tf(n, T, false, ch / dh, ch, true);
                b.layers.push_back(T);
            }
            n->inp.push_back(b);
            chans.push_back(ch);
            ++bi;
        }
        if (level != n->mult.size() - 1) {
            BlockW b;
            LayerW L = mk("down", "input_blocks." + std::to_string(bi) + ".0", ch, ch);
            L.conv = reg_linear(n, L.name + ".op", ch, ch, 9, true);
            b.layers.push_back(L);
            n->inp.push_back(b);
            chans.push_back(ch);
            ++bi;
            ds *= 2;
        }
    }
    {
        LayerW A = mk("res", "middle_block.0", ch, ch); build_res(n, A, te); n->mid.layers.push_back(A);
        LayerW S = mk("st", "middle_block.1", ch, ch); build_tf(n, S, true, ch / dh, ch, true); n->mid.layers.push_back(S);
        LayerW T = mk("tt", "middle_block.2", ch, ch); build_tf(n, T, false, ch / dh, ch, true); n->mid.layers.push_back(T);
        LayerW B2 = mk("res", "middle_block.3", ch, ch); build_res(n, B2, te); n->mid.layers.push_back(B2);
    }
    int oi = 0;
    for (int level = (int)n->mult.size() - 1; level >= 0; --level) {
        const int mult = n->mult[level];
        for (int i = 0; i <= n->nres; ++i) {
            const int ich = chans.back();
            chans.pop_back();
            BlockW b;
            b.skip_cin = ich;
            const std::string pre = "output_blocks." + std::to_string(oi);
            LayerW L = mk("res", pre + ".0", ch + ich, mc * mult);
            build_res(n, L, te);
            b.layers.push_back(L);
            ch = mc * mult;
            int li = 1;
            if (has(n->attn_res, ds)) {
                LayerW S = mk("st", pre + ".1", ch, ch);
                build_tf(n, S, true, ch / dh, ch, true);
                b.layers.push_back(S);
                LayerW T = mk("tt", pre + ".2", ch, ch);
                build_tf(n, T, false, ch / dh, ch, true);
                b.layers.push_back(T);
                li = 3;
            }
            if (level && i == n->nres) {
                LayerW U = mk("up", pre + "." + std::to_string(li), ch, ch);
                U.conv = reg_linear(n, U.name + ".conv", ch, ch, 9, true);
                b.layers.push_back(U);
                ds /= 2;
            }
            n->outb.push_back(b);
            ++oi;
        }
    }
    n->gn_out = reg_norm(n, "out.0", mc);
    n->conv_out = reg_linear(n, "out.2", n->out_ch, mc, 9, true);
    // stacked matrices: all ResBlock emb_layers (one GEMM per step) and all cross-attention to_k/to_v (one GEMM per step)
    n->emb_all = alloc_packed(n, n->n_emb, te, 1, true);
    n->kv_all = alloc_packed(n, n->n_kv > 0 ? n->n_kv : 16, n->ctx_dim, 1, false);
    auto visit = [&](BlockW& b) {
        for (LayerW& L : b.layers) {
            if (L.kind == "res") {
                reg_pack(n, L.name + ".emb_layers.1.weight", n->emb_all, L.res.emb_off, L.res.cout);
                reg_f32(n, L.name + ".emb_layers.1.bias", L.res.cout, n->emb_all.b + L.res.emb_off);
            } else if (L.kind == "st") {
                const std::string a = L.name + ".transformer_blocks.0.attn2";
                reg_pack(n, a + ".to_k.weight", n->kv_all, L.tf.kv_off, L.tf.inner);
                reg_pack(n, a + ".to_v.weight", n->kv_all, L.tf.kv_off + L.tf.inner, L.tf.inner);
            }
        }
    };
    for (auto& b : n->inp) visit(b);
    visit(n->mid);
    for (auto& b : n->outb) visit(b);
}

__global__ void copy_bias_kernel(const float* __restrict__ src, float* __restrict__ dst, int N, int geglu_bn) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int r = i;
    if (geglu_bn > 0) {
        const int half = N / 2, hb = geglu_bn / 2;
        const int j = i < half ? i : i - half;
        r = (j / hb) * geglu_bn + (i < half ? 0 : hb) + (j % hb);
    }
    dst[r] = src[i];
}

// One launch for a whole state_dict (moca_unet_load_many): job j covers work elements [start[j], start[j+1]) of the
// concatenated work list; a block takes 1024-element chunks and finds its job by bisection.
struct LoadJob {
    const float* src;
    float* dst_f32;
    op_t* dst;
    long long pitch, start;
    int kind, N, K, taps, Kp, row0, geglu_bn, pad;
};

__device__ __forceinline__ int geglu_row(int n, int N, int geglu_bn) {
    if (geglu_bn <= 0) return n;
    const int half = N / 2, hb = geglu_bn / 2;
    const int j = n < half ? n : n - half;
    return (j / hb) * geglu_bn + (n < half ? 0 : hb) + (j % hb);
}

__global__ void __launch_bounds__(256) load_many_kernel(const LoadJob* __restrict__ jobs, int njobs, long long total) {
    __shared__ int s_job;
    const long long nchunks = (total + 1023) / 1024;
    for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const long long w0 = c * 1024;
        __syncthreads();
        if (threadIdx.x == 0) {
            int lo = 0, hi = njobs - 1;                 // last job with start <= w0
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (jobs[mid].start <= w0) lo = mid; else hi = mid - 1;
            }
            s_job = lo;
        }
        __syncthreads();
        int j = s_job;
        for (int e = threadIdx.x; e < 1024; e += 256) {
            const long long w = w0 + e;
            if (w >= total) break;
            while (j + 1 < njobs && jobs[j + 1].start <= w) ++j;
            const LoadJob J = jobs[j];
            const long long i = w - J.start;
            if (J.kind == 0) {
                J.dst_f32[geglu_row((int)i, J.N, J.geglu_bn)] = J.src[i];
            } else {
                const int k = (int)(i % J.Kp);
                const int tap = (int)((i / J.Kp) % J.taps);
                const int n = (int)(i / ((long long)J.Kp * J.taps));
                const float v = k < J.K ? J.src[((long long)n * J.K + k) * J.taps + tap] : 0.0f;
                J.dst[(long long)(J.row0 + geglu_row(n, J.N, J.geglu_bn)) * J.pitch + (long long)tap * J.Kp + k] = float2op(v);
            }
        }
    }
}

// AutoencoderKL decode path: post_quant_conv (autoencoder.py:104-107) + Decoder (ae_modules.py:466-578), names as in the
// first_stage_model state_dict.
static VaeResW build_vae_res(moca_unet* n, const std::string& p, int cin, int cout) {
    VaeResW r;
    r.cin = cin; r.cout = cout;
    r.n1 = reg_norm(n, p + ".norm1", cin);
    r.conv1 = reg_linear(n, p + ".conv1", cout, cin, 9, true);
    r.n2 = reg_norm(n, p + ".norm2", cout);
    r.conv2 = reg_linear(n, p + ".conv2", cout, cout, 9, true);
    r.has_nin = cin != cout;
    if (r.has_nin) r.nin = reg_linear(n, p + ".nin_shortcut", cout, cin, 1, true);
    return r;
}

static void build_vae_arch(moca_unet* n) {
    VaeW& v = n->vae;
    const int L = (int)v.mult.size();
    int block_in = v.ch * v.mult[L - 1];
    v.pq_w = reg_f32(n, "post_quant_conv.weight", (int64_t)v.zc * v.zc);
    v.pq_b = reg_f32(n, "post_quant_conv.bias", v.zc);
    v.conv_in = reg_linear(n, "decoder.conv_in", block_in, v.zc, 9, true);
    v.mid1 = build_vae_res(n, "decoder.mid.block_1", block_in, block_in);
    v.attn_norm = reg_norm(n, "decoder.mid.attn_1.norm", block_in);
    v.aq = reg_linear(n, "decoder.mid.attn_1.q", block_in, block_in, 1, true);
    v.ak = reg_linear(n, "decoder.mid.attn_1.k", block_in, block_in, 1, true);
    v.av = reg_linear(n, "decoder.mid.attn_1.v", block_in, block_in, 1, true);
    v.aproj = reg_linear(n, "decoder.mid.attn_1.proj_out", block_in, block_in, 1, true);
    v.mid2 = build_vae_res(n, "decoder.mid.block_2", block_in, block_in);
    v.up.resize(L);
    v.upconv.resize(L);
    for (int lvl = L - 1; lvl >= 0; --lvl) {
        const int block_out = v.ch * v.mult[lvl];
        for (int j = 0; j <= v.nres; ++j) {
            v.up[lvl].push_back(build_vae_res(n, "decoder.up." + std::to_string(lvl) + ".block." + std::to_string(j), block_in, block_out));
            block_in = block_out;
        }
        if (lvl != 0) v.upconv[lvl] = reg_linear(n, "decoder.up." + std::to_string(lvl) + ".upsample.conv", block_in, block_in, 9, true);
    }
    v.norm_out = reg_norm(n, "decoder.norm_out", block_in);
    v.conv_out = reg_linear(n, "decoder.conv_out", v.out_ch, block_in, 9, true);
}

// AutoencoderKL encode path: Encoder (ae_modules.py:364-463) + quant_conv (autoencoder.py:97-102)
static void build_vae_encoder_arch(moca_unet* n) {
    VaeW& v = n->vae;
    const int L = (int)v.mult.size();
    v.conv_in = reg_linear(n, "encoder.conv_in", v.ch, v.in_ch, 9, true);
    v.down.resize(L);
    v.downconv.resize(L);
    int block_in = v.ch;
    for (int lvl = 0; lvl < L; ++lvl) {
        const int block_out = v.ch * v.mult[lvl];
        for (int j = 0; j < v.nres; ++j) {
            v.down[lvl].push_back(build_vae_res(n, "encoder.down." + std::to_string(lvl) + ".block." + std::to_string(j), block_in, block_out));
            block_in = block_out;
        }
        if (lvl != L - 1) v.downconv[lvl] = reg_linear(n, "encoder.down." + std::to_string(lvl) + ".downsample.conv", block_in, block_in, 9, true);
    }
    v.mid1 = build_vae_res(n, "encoder.mid.block_1", block_in, block_in);
    v.attn_norm = reg_norm(n, "encoder.mid.attn_1.norm", block_in);
    v.aq = reg_linear(n, "encoder.mid.attn_1.q", block_in, block_in, 1, true);
    v.ak = reg_linear(n, "encoder.mid.attn_1.k", block_in, block_in, 1, true);
    v.av = reg_linear(n, "encoder.mid.attn_1.v", block_in, block_in, 1, true);
    v.aproj = reg_linear(n, "encoder.mid.attn_1.proj_out", block_in, block_in, 1, true);
    v.mid2 = build_vae_res(n, "encoder.mid.block_2", block_in, block_in);
    v.norm_out = reg_norm(n, "encoder.norm_out", block_in);
    v.conv_out = reg_linear(n, "encoder.conv_out", v.mom_ch, block_in, 9, true);
    v.q_w = reg_f32(n, "quant_conv.weight", (int64_t)v.mom_ch * v.mom_ch);
    v.q_b = reg_f32(n, "quant_conv.bias", v.mom_ch);
}

}  // namespace moca

// Text tower object: open_clip's CLIP text transformer as FrozenOpenCLIPEmbedder uses it (lvdm/modules/encoders/condition.py:174-235).
// layers_used blocks are run and registered (layer = "penultimate": layers - 1, condition.py:197-202, 226-229); the unused last
// block, text_projection and logit_scale of the state_dict are not part of the path.  moca_text_encode (unet_plan.cu) runs it.
extern "C" int32_t moca_text_create(moca_ctx* ctx, int32_t vocab, int32_t context_length, int32_t width, int32_t heads, int32_t layers,
                                    int32_t layers_used, moca_unet** out) {
    MOCA_REQUIRE(out != nullptr, MOCA_E_ARG, "moca_text_create: NULL argument");
    MOCA_REQUIRE(vocab > 0 && context_length > 0 && context_length <= 128 && width > 0 && width % 64 == 0 && width <= 1280 && heads > 0 &&
                 width == heads * 64 && layers > 0 && layers_used > 0 && layers_used <= layers, MOCA_E_SHAPE,
                 "moca_text_create: unsupported shape (context_length %d <= 128, width %d = 64 * heads %d <= 1280, %d of %d layers)",
                 context_length, width, heads, layers_used, layers);
    moca_unet* n = new moca_unet();
    n->ctx = ctx;
    n->dry = (ctx == nullptr);
    TextW& t = n->text;
    t.present = true;
    t.vocab = vocab; t.ctx_len = context_length; t.width = width; t.heads = heads; t.layers = layers; t.used = layers_used;
    t.tok_emb = reg_f32(n, "model.token_embedding.weight", (int64_t)vocab * width);
    t.pos_emb = reg_f32(n, "model.positional_embedding", (int64_t)context_length * width);
    t.blocks.resize(layers_used);
    for (int i = 0; i < layers_used; ++i) {
        TextLayerW& b = t.blocks[i];
        const std::string p = "model.transformer.resblocks." + std::to_string(i);
        b.ln1 = reg_norm(n, p + ".ln_1", width);
        b.in_proj = alloc_packed(n, 3 * width, width, 1, true);          // nn.MultiheadAttention: q | k | v rows of one matrix
        reg_pack(n, p + ".attn.in_proj_weight", b.in_proj, 0, 3 * width);
        reg_f32(n, p + ".attn.in_proj_bias", 3 * width, b.in_proj.b);
        b.out_proj = reg_linear(n, p + ".attn.out_proj", width, width, 1, true);
        b.ln2 = reg_norm(n, p + ".ln_2", width);
        b.c_fc = reg_linear(n, p + ".mlp.c_fc", 4 * width, width, 1, true);
        b.c_proj = reg_linear(n, p + ".mlp.c_proj", width, 4 * width, 1, true);
    }
    t.ln_final = reg_norm(n, "model.ln_final", width);
    for (void* p : n->allocs)
        if (!p) {
            set_error("moca_text_create: device allocation failed");
            return MOCA_E_CUDA;
        }
    *out = n;
    return MOCA_OK;
}

// VAE encoder object (same handle conventions as the decoder below); moca_vae_encode (unet_plan.cu) runs it.
extern "C" int32_t moca_vae_encoder_create(moca_ctx* ctx, int32_t ch, const int32_t* ch_mult, int32_t n_mult, int32_t num_res_blocks,
                                           int32_t in_channels, int32_t z_channels, int32_t double_z, moca_unet** out) {
    MOCA_REQUIRE(out && ch_mult && n_mult >= 1 && n_mult <= 6, MOCA_E_ARG, "moca_vae_encoder_create: bad argument");
    MOCA_REQUIRE(ch > 0 && ch % 64 == 0 && z_channels >= 1 && z_channels <= 16 && in_channels >= 1 && in_channels <= 16 && num_res_blocks >= 1,
                 MOCA_E_SHAPE, "moca_vae_encoder_create: ch %d must be a multiple of 64 (one GEMM k-block), z_channels / in_channels <= 16", ch);
    for (int i = 0; i < n_mult; ++i) MOCA_REQUIRE(ch_mult[i] >= 1, MOCA_E_SHAPE, "moca_vae_encoder_create: ch_mult[%d] = %d", i, ch_mult[i]);
    moca_unet* n = new moca_unet();
    n->ctx = ctx;
    n->dry = (ctx == nullptr);
    n->vae.present = true;
    n->vae.encoder = true;
    n->vae.ch = ch; n->vae.zc = z_channels; n->vae.in_ch = in_channels; n->vae.nres = num_res_blocks;
    n->vae.mom_ch = (double_z ? 2 : 1) * z_channels;
    n->vae.mult.assign(ch_mult, ch_mult + n_mult);
    build_vae_encoder_arch(n);
    for (void* p : n->allocs)
        if (!p) {
            set_error("moca_vae_encoder_create: device allocation failed");
            return MOCA_E_CUDA;
        }
    *out = n;
    return MOCA_OK;
}

// VAE decoder object: the same handle type and the same moca_unet_load / _load_many / _finalize / _destroy / _param_* calls as the
// UNet; moca_vae_decode (unet_plan.cu) runs it.  ch / ch_mult / num_res_blocks / z_channels / out_ch: the yaml's ddconfig.
extern "C" int32_t moca_vae_create(moca_ctx* ctx, int32_t ch, const int32_t* ch_mult, int32_t n_mult, int32_t num_res_blocks,
                                   int32_t z_channels, int32_t out_ch, moca_unet** out) {
    MOCA_REQUIRE(out && ch_mult && n_mult >= 1 && n_mult <= 6, MOCA_E_ARG, "moca_vae_create: bad argument");
    MOCA_REQUIRE(ch > 0 && ch % 64 == 0 && z_channels >= 1 && z_channels <= 16 && out_ch >= 1 && out_ch <= 4 && num_res_blocks >= 1, MOCA_E_SHAPE,
                 "moca_vae_create: ch %d must be a multiple of 64 (one GEMM k-block), z_channels <= 16, out_ch <= 4", ch);
    for (int i = 0; i < n_mult; ++i) MOCA_REQUIRE(ch_mult[i] >= 1, MOCA_E_SHAPE, "moca_vae_create: ch_mult[%d] = %d", i, ch_mult[i]);
    moca_unet* n = new moca_unet();
    n->ctx = ctx;
    n->dry = (ctx == nullptr);
    n->vae.present = true;
    n->vae.ch = ch; n->vae.zc = z_channels; n->vae.out_ch = out_ch; n->vae.nres = num_res_blocks;
    n->vae.mult.assign(ch_mult, ch_mult + n_mult);
    build_vae_arch(n);
    for (void* p : n->allocs)
        if (!p) {
            set_error("moca_vae_create: device allocation failed");
            return MOCA_E_CUDA;
        }
    *out = n;
    return MOCA_OK;
}

extern "C" int32_t moca_unet_create(moca_ctx* ctx, int32_t in_channels, int32_t out_channels, int32_t model_channels,
                                    int32_t num_res_blocks, const int32_t* channel_mult, int32_t n_mult,
                                    const int32_t* attention_resolutions, int32_t n_attn_res, int32_t num_head_channels,
                                    int32_t context_dim, int32_t init_attn_heads, moca_unet** out) {
    MOCA_REQUIRE(out && channel_mult && attention_resolutions, MOCA_E_ARG, "moca_unet_create: NULL argument");
    MOCA_REQUIRE(num_head_channels == 64, MOCA_E_SHAPE, "moca_unet_create: attention kernels are built for head dim 64 (got %d)", num_head_channels);
    MOCA_REQUIRE(model_channels % 64 == 0 && context_dim % 8 == 0, MOCA_E_SHAPE,
                 "moca_unet_create: model_channels %d must be a multiple of 64 and context_dim %d of 8", model_channels, context_dim);
    MOCA_REQUIRE(in_channels <= 64 && out_channels <= 128, MOCA_E_SHAPE, "moca_unet_create: in/out channels too large");
    moca_unet* n = new moca_unet();
    n->ctx = ctx;
    n->dry = (ctx == nullptr);          // ctx == NULL: host-only architecture listing (moca_unet_param_*)
    n->in_ch = in_channels; n->out_ch = out_channels; n->mc = model_channels; n->nres = num_res_blocks;
    n->dh = num_head_channels; n->ctx_dim = context_dim; n->init_heads = init_attn_heads;
    n->mult.assign(channel_mult, channel_mult + n_mult);
    n->attn_res.assign(attention_resolutions, attention_resolutions + n_attn_res);
    build_arch(n);
    for (void* p : n->allocs)
        if (!p) {
            set_error("moca_unet_create: device allocation failed");
            return MOCA_E_CUDA;
        }
    *out = n;
    return MOCA_OK;
}

extern "C" int32_t moca_unet_param_count(moca_unet* n) { return n ? (int32_t)n->slots.size() : 0; }

extern "C" int32_t moca_unet_param_info(moca_unet* n, int32_t i, char* name, int32_t name_cap, int64_t* numel) {
    MOCA_REQUIRE(n && name && numel && i >= 0 && i < (int)n->slots.size(), MOCA_E_ARG, "moca_unet_param_info: bad argument");
    auto it = n->slots.begin();
    std::advance(it, i);
    snprintf(name, name_cap, "%s", it->first.c_str());
    *numel = it->second.numel;
    return MOCA_OK;
}

extern "C" int32_t moca_unet_load(moca_unet* n, const char* name, const float* data, int64_t numel, void* stream) {
    MOCA_REQUIRE(n && name && data, MOCA_E_ARG, "moca_unet_load: NULL argument");
    MOCA_REQUIRE(!n->dry, MOCA_E_STATE, "moca_unet_load: network was created without a context (listing only)");
    auto it = n->slots.find(name);
    MOCA_REQUIRE(it != n->slots.end(), MOCA_E_ARG, "moca_unet_load: unknown parameter '%s'", name);
    LoadSlot& s = it->second;
    MOCA_REQUIRE(s.numel == numel, MOCA_E_SHAPE, "moca_unet_load: '%s' has %lld elements, expected %lld", name, (long long)numel,
                 (long long)s.numel);
    cudaStream_t st = (cudaStream_t)stream;
    if (s.kind == LoadSlot::COPY_F32) {
        copy_bias_kernel<<<cdiv(numel, 256), 256, 0, st>>>(data, s.dst_f32, (int)numel, s.geglu_bn);
        MOCA_LAUNCH_CHECK();
    } else {
        int rc = pack_weight(data, s.N, s.K, s.taps, s.Kp, s.dst, s.pitch, s.row0, s.geglu_bn, st);
        if (rc) return rc;
    }
    s.loaded = true;
    n->finalized = false;
    return MOCA_OK;
}

// The whole state_dict in ONE kernel launch: names[i] / data[i] (fp32 DEVICE pointers, all valid until the stream has run
// the kernel) / numel[i].  Same conversion as moca_unet_load.
extern "C" int32_t moca_unet_load_many(moca_unet* n, int32_t count, const char* const* names, const float* const* data,
                                       const int64_t* numel, void* stream) {
    MOCA_REQUIRE(n && names && data && numel && count > 0, MOCA_E_ARG, "moca_unet_load_many: NULL argument");
    MOCA_REQUIRE(!n->dry, MOCA_E_STATE, "moca_unet_load_many: network was created without a context (listing only)");
    std::vector<LoadJob> jobs((size_t)count);
    long long total = 0;
    for (int i = 0; i < count; ++i) {
        MOCA_REQUIRE(names[i] && data[i], MOCA_E_ARG, "moca_unet_load_many: NULL entry %d", i);
        auto it = n->slots.find(names[i]);
        MOCA_REQUIRE(it != n->slots.end(), MOCA_E_ARG, "moca_unet_load_many: unknown parameter '%s'", names[i]);
        const LoadSlot& s = it->second;
        MOCA_REQUIRE(s.numel == numel[i], MOCA_E_SHAPE, "moca_unet_load_many: '%s' has %lld elements, expected %lld", names[i],
                     (long long)numel[i], (long long)s.numel);
        LoadJob& J = jobs[i];
        memset(&J, 0, sizeof(J));
        J.src = data[i]; J.start = total; J.geglu_bn = s.geglu_bn;
        if (s.kind == LoadSlot::COPY_F32) {
            J.kind = 0; J.dst_f32 = s.dst_f32; J.N = (int)s.numel;
            total += s.numel;
        } else {
            J.kind = 1; J.dst = s.dst; J.N = s.N; J.K = s.K; J.taps = s.taps; J.Kp = s.Kp; J.row0 = s.row0; J.pitch = s.pitch;
            total += (long long)s.N * s.taps * s.Kp;
        }
    }
    cudaStream_t st = (cudaStream_t)stream;
    LoadJob* d_jobs = nullptr;
    MOCA_CHECK_CUDA(cudaMalloc(&d_jobs, sizeof(LoadJob) * (size_t)count));
    cudaError_t ce = cudaMemcpyAsync(d_jobs, jobs.data(), sizeof(LoadJob) * (size_t)count, cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) {
        const long long nchunks = (total + 1023) / 1024;
        const int blocks = (int)(nchunks < (long long)n->ctx->c.num_sms * 16 ? nchunks : (long long)n->ctx->c.num_sms * 16);
        load_many_kernel<<<blocks, 256, 0, st>>>(d_jobs, count, total);
        ce = cudaGetLastError();
    }
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);       // jobs (host vector) and d_jobs are released below
    cudaFree(d_jobs);
    MOCA_REQUIRE(ce == cudaSuccess, MOCA_E_CUDA, "moca_unet_load_many: %s", cudaGetErrorString(ce));
    for (int i = 0; i < count; ++i) n->slots[names[i]].loaded = true;
    n->finalized = false;
    return MOCA_OK;
}

extern "C" int32_t moca_unet_finalize(moca_unet* n) {
    MOCA_REQUIRE(n, MOCA_E_ARG, "moca_unet_finalize: NULL");
    for (auto& kv : n->slots)
        MOCA_REQUIRE(kv.second.loaded, MOCA_E_STATE, "moca_unet_finalize: parameter '%s' was never loaded", kv.first.c_str());
    MOCA_CHECK_CUDA(cudaDeviceSynchronize());
    n->finalized = true;
    return MOCA_OK;
}
