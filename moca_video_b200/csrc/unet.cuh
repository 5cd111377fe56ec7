// UNet runner: packed weights + per-shape execution plan (ops list, workspace, CUDA graph).
// Mirrors UNetModel (lvdm/modules/networks/openaimodel3d.py:279-578) for the option set of
// configs/inference_t2v_512_v2.0.yaml (use_linear, temporal_conv, temporal_attention, temporal_selfatt_only,
// addition_attention, fps_cond, num_head_channels heads, no relative position, no image cross-attention).
#pragma once
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "attn.cuh"
#include "common.cuh"
#include "gemm_tc.cuh"
#include "norm.cuh"

namespace moca {

struct PackedW {               // op_t [N, taps*Kp]
    op_t* w = nullptr;
    float* b = nullptr;        // fp32 [N] or null
    int N = 0, K = 0, Kp = 0, taps = 1;
};

struct NormW {
    float* g = nullptr;
    float* b = nullptr;
};

struct ResW {
    int cin = 0, cout = 0;
    NormW gn1, gn2;
    PackedW conv1, conv2, skip;
    bool has_skip = false;
    int emb_off = 0;           // column offset into the stacked emb_layers output
    NormW tgn[4];
    PackedW tconv[4];
};

struct TfW {
    bool spatial = true;
    bool linear = true;
    int C = 0, inner = 0, heads = 0;
    NormW gn, ln1, ln2, ln3;
    PackedW proj_in, proj_out;
    PackedW qkv1, out1;        // attn1: fused q|k|v [3*inner, inner], to_out
    PackedW q2, out2;          // attn2 (spatial): to_q ; k|v live in the stacked context projection at kv_off
    PackedW qkv2;              // attn2 (temporal): fused self-attention
    int kv_off = 0;
    PackedW ff1, ff2;
};

struct LayerW {
    std::string kind, name;
    int cin = 0, cout = 0;
    ResW res;
    TfW tf;
    PackedW conv;              // conv_in / down / up
};

struct BlockW {
    std::vector<LayerW> layers;
    int skip_cin = 0;
};

// ---- SD-VAE decoder (lvdm/modules/networks/ae_modules.py:466-578, the "next" row N1) ------------------------------------
struct VaeResW {               // ResnetBlock (:151-210), temb_channels = 0
    int cin = 0, cout = 0;
    NormW n1, n2;
    PackedW conv1, conv2, nin;
    bool has_nin = false;
};
struct VaeW {
    bool present = false;
    int ch = 128, zc = 4, out_ch = 3, nres = 2;
    std::vector<int> mult;
    float* pq_w = nullptr;                           // post_quant_conv [zc, zc] (1x1) and bias, applied while staging z
    float* pq_b = nullptr;
    PackedW conv_in, conv_out;
    VaeResW mid1, mid2;
    NormW attn_norm, norm_out;
    PackedW aq, ak, av, aproj;                       // AttnBlock (:26-78): 1x1 convs, one head of block_in channels
    std::vector<std::vector<VaeResW>> up;            // up[level][block], level as in the state_dict
    std::vector<PackedW> upconv;                     // up[level].upsample.conv (level > 0)
    // Encoder (:364-463) + quant_conv (autoencoder.py:97-102): conv_in / mid* / attn* / norm_out / conv_out above are the encoder's
    bool encoder = false;
    int in_ch = 3, mom_ch = 8;                       // image channels; moments = 2 * z_channels (double_z)
    std::vector<std::vector<VaeResW>> down;          // down[level][block]
    std::vector<PackedW> downconv;                   // down[level].downsample.conv (level < L - 1): pad (0,1,0,1), stride 2
    float* q_w = nullptr;                            // quant_conv [mom_ch, mom_ch] (1x1) and bias, applied while un-staging
    float* q_b = nullptr;
};

// OpenCLIP text tower: token / positional embeddings, pre-LN residual attention blocks, ln_final; names as in
// FrozenOpenCLIPEmbedder.state_dict() ("model.transformer.resblocks.{i}.attn.in_proj_weight", ...)
struct TextLayerW {
    NormW ln1, ln2;
    PackedW in_proj, out_proj, c_fc, c_proj;
};
struct TextW {
    bool present = false;
    int vocab = 49408, ctx_len = 77, width = 1024, heads = 16, layers = 24, used = 23;   // used: blocks actually run ("penultimate": layers - 1)
    float* tok_emb = nullptr;                        // [vocab, width] fp32
    float* pos_emb = nullptr;                        // [ctx_len, width] fp32
    std::vector<TextLayerW> blocks;
    NormW ln_final;
};

struct LoadSlot {
    enum Kind { COPY_F32, PACK } kind = COPY_F32;
    float* dst_f32 = nullptr;
    op_t* dst = nullptr;
    int N = 0, K = 0, taps = 1, Kp = 0, row0 = 0, geglu_bn = 0;
    int64_t pitch = 0;
    int64_t numel = 0;
    bool loaded = false;
};

struct Plan;

}  // namespace moca

struct moca_unet {
    moca_ctx* ctx = nullptr;
    int in_ch = 4, out_ch = 4, mc = 320, nres = 2, dh = 64, ctx_dim = 1024, init_heads = 8;
    std::vector<int> mult, attn_res;
    std::vector<moca::BlockW> inp, outb;
    moca::BlockW init_attn, mid;
    moca::PackedW time0, time2, fps0, fps2, emb_all, kv_all, conv_out;
    moca::NormW gn_out;
    moca::TextW text;                                // moca_text_create: this object is a text tower
    moca::VaeW vae;                                  // moca_vae_create: this object is a VAE decoder, not a UNet
    int n_emb = 0, n_kv = 0;
    std::unordered_map<std::string, moca::LoadSlot> slots;
    std::vector<void*> allocs;
    bool finalized = false;
    bool dry = false;
    bool taps = false;
    std::map<std::vector<int>, std::shared_ptr<moca::Plan>> plans;
    moca::Plan* last = nullptr;
    uint64_t use_clock = 0;
    size_t weight_bytes = 0;
};
