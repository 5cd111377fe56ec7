/*
 * moca_b200.h — C-ABI of libmoca_b200.so, the B200 (sm_100a) implementation of the MoCA-Video hot path.
 *
 * The reference (ZhangT-tech/MoCA-Video) is pure Python: its "operator API" for this path is the Python
 * method surface listed in SURVEY.md §8(b).  Each entry point below names the reference method whose body
 * it replaces (paths relative to the reference tree).  The Python shim in moca_video_b200/ binds these with
 * ctypes; INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns int32 status: MOCA_OK or a negative MOCA_E_* code; moca_last_error() gives the
 *     message of the last failure on the calling thread.  No C++ exception crosses this boundary.
 *   - all tensor pointers are DEVICE pointers unless the parameter is documented "host"; tensors are dense
 *     fp32 in the reference's own layout (latents [B,C,F,H,W]).  The library never frees caller memory.
 *   - `stream` is a cudaStream_t passed as void*; every call is asynchronous on it (no hidden sync), and
 *     kernels never draw random numbers: noise is always an input (RNG parity with the reference).
 *   - there is no CPU fallback: on a device that is not sm_100 moca_ctx_create fails with MOCA_E_ARCH.
 */
#ifndef MOCA_B200_H
#define MOCA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MOCA_OK 0
#define MOCA_E_SHAPE (-1)
#define MOCA_E_DTYPE (-2)
#define MOCA_E_ALIGN (-3)
#define MOCA_E_ARCH (-4)
#define MOCA_E_CUDA (-5)
#define MOCA_E_NCCL (-6)
#define MOCA_E_ARG (-7)
#define MOCA_E_STATE (-8)

#define MOCA_MAX_FRAMES 64
#define MOCA_MAX_MASKS 1024    /* F * masks-per-frame of the SAM branch; also the scratch size of the mask pre-passes */

typedef struct moca_ctx moca_ctx;       /* one per process / GPU; not thread-safe */
typedef struct moca_unet moca_unet;     /* packed 16-bit weights + execution plan of one UNet */

const char* moca_last_error(void);
int32_t moca_abi_version(void);
/* 16-bit operand type this build uses for every tensor-core contraction and for the "op16" buffers of the test
 * primitives below: MOCA_DTYPE_F16 (default; IEEE half, saturating conversions) or MOCA_DTYPE_BF16 (-DMOCA_OPERAND_BF16). */
#define MOCA_DTYPE_F16 0
#define MOCA_DTYPE_BF16 1
int32_t moca_operand_dtype(void);

/* ---- context ------------------------------------------------------------------------------------------ */
int32_t moca_ctx_create(int32_t device, moca_ctx** out);
int32_t moca_ctx_destroy(moca_ctx* ctx);

/* ---- K13: DDIMSampler.unet CFG combine (lvdm/models/samplers/ddim.py:362-374) fused with
 *           DDIMSampler.ddim_step (ddim.py:377-649; exact math in SURVEY.md Appendix B) -------------------
 * x, e_c, e_u, noise, x_prev, pred_x0: [B,C,F,H,W] fp32.  e_u may be NULL (no guidance; ddim.py:366).
 * coef: HOST float[4*F] rows A, A_prev, sigma, sqrt(1-A) for each frame's table index (ddim.py:409-412).
 * ts:   HOST int32[F] DDPM timestep of each frame (ddim.py:428).
 * mask: [B,1,Fm,H,W] fp32 or NULL; mask_idx: HOST int32[F], mask frame used by frame f, <0 = no blend
 *       (the reference's clobbered loop index, SURVEY Q3, is expressed by the caller through this table).
 * cond_img: [B,C,1,H,W] fp32 (ignored when mask is NULL).
 * enh_lo / enh_hi: blend gain for t <= t_thresh / t > t_thresh (ddim.py:584: 1.5 / 1.0; SAM branch :845: 2 / 0=off).
 * cover_max: masks whose sum exceeds cover_max*H*W*B are skipped; <= 0 = no coverage test (the DAVIS branch has none,
 *       ddim.py:587; the SAM branch's 80 % rule lives in moca_sam_blend).
 * F <= MOCA_MAX_FRAMES. */
int32_t moca_cfg_ddim_step(moca_ctx* ctx, const float* x, const float* e_c, const float* e_u, const float* noise,
                           float cfg_scale, const float* coef, const int32_t* ts, const float* mask, int32_t Fm,
                           const int32_t* mask_idx, const float* cond_img, double gamma, float enh_lo, float enh_hi,
                           int32_t t_thresh, float cover_max, int32_t B, int32_t C, int32_t F, int32_t H, int32_t W,
                           float* x_prev, float* pred_x0, void* stream);

/* ---- SAM branch of ddim_step (no DAVIS masks): DDIMSampler.apply_cond_img -> _apply_segmentation
 *      (ddim.py:591-606, 739-903) and calculate_iou (:905-943) ------------------------------------------------
 * Grounding-DINO / SAM-2 themselves are third-party networks: their masks are inputs.  The flow of one step is
 *   1. moca_cfg_ddim_step(gamma = 0, no mask)  -> x_prev and the momentum-corrected x0 prediction p' (pred_x0 slot)
 *   2. per frame with t <= 300, in frame order: moca_latent_to_u8(p', frame) -> image for the segmenter (:745-761);
 *      no box -> previous frame's masks (:787-793); else moca_mask_iou(new, previous) < 0.5 -> previous (:809-813)
 *   3. moca_sam_blend: masks applied in order, a mask covering more than cover_max (0.8) of the frame RESETS the frame
 *      to p' (:826-828), otherwise where(mask > 0.5, cond_img * gain (2), cur) (:845, :897-901); then
 *      pred_x0 = (1 - gamma) * cur + gamma * sigma * noise (:609).
 * moca_latent_to_u8: pred_x0 = sample 0 of [B,C,F,H,W] (the reference squeezes b = 1); image: DEVICE uint8 [H,W,3].
 * moca_mask_iou:     masks fp32 [n,H,W]; pairs = zip(a, b); *iou_mean is a HOST float; synchronises the stream
 *                    (the caller branches on it, as the reference does with .item()).
 * moca_sam_blend:    p_corr, noise, pred_x0 [B,C,F,H,W]; sigma HOST float[F]; cond_img [B,C,1,H,W];
 *                    masks fp32 [F,Nmax,H,W]; n_masks HOST int32[F] (0 = frame untouched; F*Nmax <= MOCA_MAX_MASKS). */
int32_t moca_latent_to_u8(moca_ctx* ctx, const float* pred_x0, int32_t C, int32_t F, int32_t H, int32_t W, int32_t frame,
                          uint8_t* image, void* stream);
int32_t moca_mask_iou(moca_ctx* ctx, const float* masks_a, int32_t n_a, const float* masks_b, int32_t n_b, int32_t H,
                      int32_t W, float* iou_mean, void* stream);
int32_t moca_sam_blend(moca_ctx* ctx, const float* p_corr, const float* noise, const float* sigma, const float* cond_img,
                       const float* masks, const int32_t* n_masks, int32_t Nmax, double gamma, float gain,
                       float cover_max, int32_t B, int32_t C, int32_t F, int32_t H, int32_t W, float* pred_x0,
                       void* stream);

/* ---- DDIMSampler.p_sample_ddim (ddim.py:274-359): CFG + plain DDIM update, scalar table index -------------
 * n = B*C*F*H*W elements.  coef: A, A_prev, sigma, sqrt(1-A).  use_scale selects ddim.py:349-355. */
int32_t moca_p_sample_ddim(moca_ctx* ctx, const float* x, const float* e_c, const float* e_u, const float* noise,
                           float cfg_scale, float A, float A_prev, float sigma, float sqrt_one_minus_A,
                           int32_t use_scale, float scale, float scale_prev, int64_t n, float* x_prev,
                           float* pred_x0, void* stream);

/* ---- K15: prepare_latents (scripts/evaluation/funcs.py:55-80): diagonal re-noising into the FIFO queue ----
 * z [B,C,Fz,H,W] clean latents, noise/queue [B,C,Q,H,W]; src_frame HOST int32[Q] (frame of z for queue slot q),
 * alpha HOST float[Q] (ddim_alphas of slot q).  queue[q] = alpha^0.5 * z[src] + (1-alpha)^0.5 * noise[q]. */
int32_t moca_renoise(moca_ctx* ctx, const float* z, const float* noise, const int32_t* src_frame, const float* alpha,
                     int32_t B, int32_t C, int32_t Fz, int32_t Q, int32_t H, int32_t W, float* queue, void* stream);

/* ---- K15: shift_latents (funcs.py:86-99): queue[:, :, :-1] = queue[:, :, 1:]; queue[:, :, -1] = tail -------
 * out may not alias queue.  tail [B,C,1,H,W]. */
int32_t moca_queue_shift(moca_ctx* ctx, const float* queue, const float* tail, int32_t B, int32_t C, int32_t Q,
                         int32_t H, int32_t W, float* out, void* stream);

/* ---- FIFO window gather / look-ahead write-back (funcs.py:307-316, 351-352) -------------------------------
 * gather:  win[B,C,Fw,H,W] = queue[:, :, start:start+Fw]
 * scatter: queue[:, :, start:start+Fw] = win[:, :, win_off:win_off+Fw] for a window of Fwin frames. */
int32_t moca_window_gather(moca_ctx* ctx, const float* queue, int32_t B, int32_t C, int32_t Q, int32_t H, int32_t W,
                           int32_t start, int32_t Fw, float* win, void* stream);
int32_t moca_window_scatter(moca_ctx* ctx, const float* win, int32_t Fwin, int32_t win_off, int32_t B, int32_t C,
                            int32_t Q, int32_t H, int32_t W, int32_t start, int32_t Fw, float* queue, void* stream);

/* ---- K14: FreeInit (utils/freeinit_utils.py) ---------------------------------------------------------------
 * moca_freeinit_filter: get_freq_filter (freeinit_utils.py:51-70) -> lpf [T,H,W] fp32.
 *   type: 0 gaussian (:73-91), 1 butterworth (:94-113), 2 ideal (:116-134), 3 box (:137-157).
 * moca_freeinit_mix: freq_mix_3d (freeinit_utils.py:7-47): y = Re IFFT3( ifftshift( fftshift(FFT3 x)*lpf
 *   + fftshift(FFT3 noise)*(1-lpf) ) ) over the last three dims; x, noise, y [NV,T,H,W] (NV = B*C volumes).
 *   work: device scratch of moca_freeinit_work_bytes(NV,T,H,W) bytes. */
int32_t moca_freeinit_filter(moca_ctx* ctx, int32_t type, int32_t n, double d_s, double d_t, int32_t T, int32_t H,
                             int32_t W, float* lpf, void* stream);
int64_t moca_freeinit_work_bytes(int32_t NV, int32_t T, int32_t H, int32_t W);
int32_t moca_freeinit_mix(moca_ctx* ctx, const float* x, const float* noise, const float* lpf, int32_t NV, int32_t T,
                          int32_t H, int32_t W, float* y, void* work, void* stream);

/* moca_freeinit_mix_static: the same computation as moca_freeinit_mix without caller-provided work memory and without any
 *   allocation.  It has NO counterpart in the reference; it was added for callers that
 *   cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The spectrum scratch
 *   is a file-scope device array of moca_freeinit_mix_static_max_elems() = 262144 complex values (2 MiB, resident on every GPU
 *   that loads the library until exit); NV*T*H*W beyond that returns MOCA_E_SHAPE with a message naming the limit.  All calls on
 *   a device share the array: calls that overlap on the GPU (different streams without ordering) race; same-stream calls do not.
 *   The array keeps the previous call's contents; every element a call reads is first rewritten by that call. */
int64_t moca_freeinit_mix_static_max_elems(void);
int32_t moca_freeinit_mix_static(moca_ctx* ctx, const float* x, const float* noise, const float* lpf, int32_t NV, int32_t T,
                                 int32_t H, int32_t W, float* y, void* stream);

/* ---- UNet: UNetModel (lvdm/modules/networks/openaimodel3d.py:279-578) ---------------------------------------
 * moca_unet_create: architecture integers of the yaml (configs/inference_t2v_512_v2.0.yaml:22-50).
 * moca_unet_load:   one call per state_dict tensor (name = state_dict key, fp32 DEVICE pointer, element count);
 *                   the library converts to its packed 16-bit layouts and keeps its own copy.
 * moca_unet_finalize: checks every parameter arrived.
 * moca_unet_forward: UNetModel.forward (openaimodel3d.py:534-578).
 *   x, y      [B,Cin,T,H,W] fp32 (reference layout)
 *   timesteps HOST int32[B*T]: per-(b,t) timestep (the reference's is_fifo case has B=1, openaimodel3d.py:535,548;
 *             the plain case repeats each sample's t over its T frames, :545-549)
 *   context   [B,Lmax,ctx_dim] fp32, ctx_len HOST int32[B] valid tokens per sample
 *   fps       HOST int32[B]
 * Shapes (B,T,H,W,Lmax) select a cached execution plan (workspace + CUDA graph), built on first use. */
int32_t moca_unet_create(moca_ctx* ctx, int32_t in_channels, int32_t out_channels, int32_t model_channels,
                         int32_t num_res_blocks, const int32_t* channel_mult, int32_t n_mult,
                         const int32_t* attention_resolutions, int32_t n_attn_res, int32_t num_head_channels,
                         int32_t context_dim, int32_t init_attn_heads, moca_unet** out);
int32_t moca_unet_destroy(moca_unet* net);
int32_t moca_unet_load(moca_unet* net, const char* name, const float* data, int64_t numel, void* stream);
/* The whole state_dict in one kernel launch (SURVEY §8b `moca_weights_pack`): names / data (fp32 DEVICE pointers) / numel
 * arrays of `count` entries; synchronises the stream before returning, so the tensors may be freed afterwards. */
int32_t moca_unet_load_many(moca_unet* net, int32_t count, const char* const* names, const float* const* data,
                            const int64_t* numel, void* stream);
int32_t moca_unet_finalize(moca_unet* net);
int32_t moca_unet_forward(moca_unet* net, const float* x, const int32_t* timesteps, const float* context,
                          const int32_t* ctx_len, int32_t Lmax, const int32_t* fps, int32_t B, int32_t T, int32_t H,
                          int32_t W, float* y, void* stream);
/* Debug/parity taps: number of recorded taps of the last forward built with taps enabled, and a copy of tap i as
 * fp32 [rows, channels] channels-last tokens (rows = B*T*H_l*W_l).  Enable with moca_unet_set_taps(net, 1)
 * before the first forward of a shape; taps disable CUDA-graph replay. */
int32_t moca_unet_set_taps(moca_unet* net, int32_t enable);
int32_t moca_unet_num_taps(moca_unet* net);
int32_t moca_unet_tap_info(moca_unet* net, int32_t i, char* name, int32_t name_cap, int64_t* rows, int64_t* channels);
int32_t moca_unet_tap_read(moca_unet* net, int32_t i, float* dst, void* stream);
/* Count of kernels one forward launches (for bench.py's gpu_launches). */
int32_t moca_unet_launches(moca_unet* net);
/* Algorithmic FLOPs (2*MAC, as the reference computes them) of the last planned forward. */
double moca_unet_flops(moca_unet* net);

/* Host-only architecture listing: moca_unet_create with ctx == NULL builds the parameter table without touching a
 * device; param_count / param_info enumerate the state_dict keys the loader expects (CPU tests compare them with
 * the reference's own state_dict listing). */
/* Per-module entry points: one ResBlock (with its TemporalConvBlock; lvdm/modules/networks/openaimodel3d.py:208-234, 269-276),
 * SpatialTransformer (lvdm/modules/attention.py:262-278) or TemporalTransformer (:331-373) of the loaded network, named as in the
 * reference's state_dict ("input_blocks.1.0", "middle_block.1", "init_attn.0", ...).  x, y: [B, C, T, H, W] fp32 (the reference
 * modules' [(b t), c, h, w] / [b, c, t, h, w] inputs are views of this).  emb: [B*T, 4*model_channels] fp32, the embedding
 * ResBlock.forward receives (ResBlock only).  context: [B, Lmax, context_dim] fp32 with ctx_len[B] valid tokens (SpatialTransformer
 * only; the reference repeats it per frame).  The same kernels and plan fragments as moca_unet_forward. */
int32_t moca_unet_layer_forward(moca_unet* net, const char* layer, const float* x, const float* emb, const float* context,
                                const int32_t* ctx_len, int32_t Lmax, int32_t B, int32_t T, int32_t H, int32_t W, float* y,
                                void* stream);
int32_t moca_unet_layer_channels(moca_unet* net, const char* layer, int32_t* c_in, int32_t* c_out);
int32_t moca_unet_param_count(moca_unet* net);
int32_t moca_unet_param_info(moca_unet* net, int32_t i, char* name, int32_t name_cap, int64_t* numel);
int64_t moca_unet_workspace_bytes(moca_unet* net);
/* Per-op device timing of the last planned forward (CSV "idx,us,gflop,label" into buf; returns bytes needed). */
int64_t moca_unet_profile(moca_unet* net, int32_t iters, char* buf, int64_t cap, void* stream);
int64_t moca_unet_weight_bytes(moca_unet* net);

/* ---- first-stage decoder: AutoencoderKL.decode (lvdm/models/autoencoder.py:104-107) = post_quant_conv + Decoder
 * (lvdm/modules/networks/ae_modules.py:466-578: conv_in, mid ResnetBlock :151-210 / AttnBlock :26-78 / ResnetBlock, the up levels
 * with nearest x2 Upsample :103-116, norm_out, SiLU, conv_out), called per frame by decode_first_stage_2DAE (ddpm3d.py:555-562).
 * moca_vae_create: the yaml's ddconfig integers (configs/inference_t2v_512_v2.0.yaml:60-72); ctx == NULL lists parameters only.
 *   The handle is a moca_unet: moca_unet_load / _load_many / _finalize / _destroy / _param_* / _launches / _flops / _profile apply,
 *   with the first_stage_model state_dict keys ("post_quant_conv.weight", "decoder.mid.attn_1.q.weight", ...); encoder and
 *   quant_conv keys are not part of the decode path and are not expected.
 * moca_vae_decode: z [B, z_channels, H, W] fp32 latents as the sampler leaves them; the library applies 1 / scale_factor
 *   (ddpm3d.py:559) and writes image [B, out_ch, H * 2^(n_mult-1), W * 2^(n_mult-1)] fp32.  B <= 16 frames per call. */
int32_t moca_vae_create(moca_ctx* ctx, int32_t ch, const int32_t* ch_mult, int32_t n_mult, int32_t num_res_blocks,
                        int32_t z_channels, int32_t out_ch, moca_unet** out);
int32_t moca_vae_decode(moca_unet* vae, const float* z, int32_t B, int32_t H, int32_t W, float scale_factor, float* image,
                        void* stream);
/* Encode path: AutoencoderKL.encode (lvdm/models/autoencoder.py:97-102) = Encoder (ae_modules.py:364-463; its Downsample :80-100 is
 * F.pad (0,1,0,1) + a stride-2 convolution) + quant_conv, used by encode_first_stage_2DAE (lvdm/models/ddpm3d.py:485-490) in the DAVIS
 * branch of prepare_latents (scripts/evaluation/funcs.py:38-47).
 * moca_vae_encoder_create: ddconfig integers; state_dict keys "encoder.*" and "quant_conv.*".
 * moca_vae_encode: x [B, in_channels, H, W] fp32 image -> moments [B, 2 * z_channels, H / 2^(n_mult-1), W / 2^(n_mult-1)] fp32
 *   (mean | logvar, the tensor DiagonalGaussianDistribution, lvdm/distributions.py:24-33, is constructed from; sampling the
 *   posterior and the scale_factor, ddpm3d.py:458-465, stay with the caller: the reference draws that noise on the host). */
int32_t moca_vae_encoder_create(moca_ctx* ctx, int32_t ch, const int32_t* ch_mult, int32_t n_mult, int32_t num_res_blocks,
                                int32_t in_channels, int32_t z_channels, int32_t double_z, moca_unet** out);
int32_t moca_vae_encode(moca_unet* vae, const float* x, int32_t B, int32_t H, int32_t W, float* moments, void* stream);

/* ---- text tower: FrozenOpenCLIPEmbedder.encode_with_transformer (lvdm/modules/encoders/condition.py:212-230) = open_clip's CLIP
 * text transformer (open_clip_torch 2.30.0, reference requirements.txt:196; the package is not vendored in the reference): token
 * + positional embedding, pre-LN residual blocks (nn.MultiheadAttention with the causal mask, c_fc -> GELU -> c_proj), ln_final.
 * moca_text_create: ViT-H-14 text tower = (49408, 77, 1024, 16, 24); layers_used = layers - 1 for layer "penultimate"
 *   (condition.py:197-202, 226-229), layers for "last".  ctx == NULL lists parameters only.  The handle is a moca_unet:
 *   moca_unet_load / _load_many / _finalize / _destroy / _param_* apply, keys as in FrozenOpenCLIPEmbedder.state_dict()
 *   ("model.token_embedding.weight", "model.transformer.resblocks.0.attn.in_proj_weight", ...); the unused last block,
 *   text_projection and logit_scale are not expected.
 * moca_text_encode: tokens DEVICE int32 [B, context_length] (open_clip.tokenize output; the BPE tokenizer is host string work and
 *   stays with the caller) -> z [B, context_length, width] fp32, the c_crossattn entry of the sampler's conditioning.  B <= 64. */
int32_t moca_text_create(moca_ctx* ctx, int32_t vocab, int32_t context_length, int32_t width, int32_t heads, int32_t layers,
                         int32_t layers_used, moca_unet** out);
int32_t moca_text_encode(moca_unet* text, const int32_t* tokens, int32_t B, float* z, void* stream);

/* ---- primitives exposed for unit parity tests (tests/test_gpu_ops.py): the same kernels the UNet plan launches ----
 * moca_gemm_op16: D[M,N] = A[M,K] * W[N,K]^T (+bias[N]) on the tcgen05 core; A, W op16 (moca_operand_dtype) row-major (K contiguous),
 *   D fp32 or op16 (out_op16).  K % 64 == 0.                                    (nn.Linear, attention.py:54-57)
 * moca_geglu_f32: FeedForward's GEGLU projection (attention.py:376-383) with the gate fused in the GEMM epilogue.
 * moca_conv3x3_f32: Conv2d 3x3 padding 1 stride 1|2 on channels-last tokens  (openaimodel3d.py:149-153, 66-70)
 * moca_tconv_f32:   Conv3d (3,1,1) padding (1,0,0) on tokens [B,T,HW,C]       (openaimodel3d.py:252-263)
 * moca_groupnorm_op16 / moca_layernorm_op16: normalisation (+SiLU) producing the 16-bit GEMM operand
 *                                                          (basics.py:76-87; attention.py:199-201)
 * moca_attention_op16: softmax(q k^T / 8) v, head dim 64                      (attention.py:92-114)
 * These helpers allocate scratch and synchronise; they are a test surface, not the hot path. */
int32_t moca_gemm_op16(moca_ctx* ctx, const void* A, const void* Wt, const float* bias, int32_t M, int32_t N, int32_t K,
                       void* D, int32_t out_op16, void* stream);
int32_t moca_geglu_f32(moca_ctx* ctx, const void* A_op16, const float* w, const float* bias, int32_t M, int32_t N, int32_t K,
                       void* out_op16, void* stream);
int32_t moca_conv3x3_f32(moca_ctx* ctx, const float* x_nhwc, const float* w_oihw, const float* bias, int32_t NF,
                         int32_t H, int32_t W, int32_t Cin, int32_t Cout, int32_t stride, float* y_nhwc, void* stream);
int32_t moca_tconv_f32(moca_ctx* ctx, const float* x, const float* w, const float* bias, int32_t B, int32_t T, int32_t HW,
                       int32_t Cin, int32_t Cout, float* y, void* stream);
int32_t moca_groupnorm_op16(moca_ctx* ctx, const float* x, const float* gamma, const float* beta, int64_t rows, int32_t C,
                            int32_t rows_per_stat, float eps, int32_t silu, void* out_op16, void* stream);
int32_t moca_layernorm_op16(moca_ctx* ctx, const float* x, const float* gamma, const float* beta, int64_t rows, int32_t C,
                            void* out_op16, void* stream);
int32_t moca_attention_op16(moca_ctx* ctx, const void* q, const void* k, const void* v, int32_t batch, int32_t heads,
                            int32_t Lq, int32_t Lk, int32_t Lk_valid, int32_t q_stride, int32_t kv_stride, void* out,
                            int32_t o_stride, int32_t temporal_inner, void* stream);

/* ---- multi-GPU queue hand-off (replaces the mp.Queue/CUDA-IPC exchange of
 *      scripts/evaluation/funcs_mp.py:197-236) -----------------------------------------------------------------
 * One process per GPU.  moca_comm_unique_id (rank 0) -> the 128 bytes reach every rank through any channel ->
 * moca_comm_init(ctx, id, nranks, rank) creates the NCCL communicator of this context (libnccl.so.2 is resolved with dlopen,
 * in a PyTorch process the copy torch loaded).  moca_queue_handoff is the per-iteration exchange of the updated half-windows:
 * an all-gather of `count` floats per rank on the caller's stream (packed: count floats, gathered: nranks * count).
 * moca_halfwin_pack / _unpack move the half-windows of the windows a rank owns in and out of those staging buffers. */
int32_t moca_comm_unique_id(void* id_out_128_bytes);
int32_t moca_comm_init(moca_ctx* ctx, const void* id_128_bytes, int32_t nranks, int32_t rank);
int32_t moca_queue_handoff(moca_ctx* ctx, const float* packed, float* gathered, int64_t count, void* stream);
int32_t moca_comm_destroy(moca_ctx* ctx);
int32_t moca_comm_size(moca_ctx* ctx);          /* 0: no communicator */
int32_t moca_halfwin_pack(moca_ctx* ctx, const float* xprev, int32_t Fwin, int32_t half, int32_t B, int32_t C, int32_t H,
                          int32_t W, float* packed, void* stream);
int32_t moca_halfwin_unpack(moca_ctx* ctx, const float* packed, int32_t n_windows, const int32_t* window_start,
                            int32_t half, int32_t B, int32_t C, int32_t Q, int32_t H, int32_t W, float* queue,
                            void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MOCA_B200_H */
