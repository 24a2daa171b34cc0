/* srb.h — C ABI of libsrb.so, the B200 (sm_100a) implementation of the HybridSuperResolutionModel hot path.
 *
 * The reference (GDev96/SuperResolution) is pure Python and has no FFI of its own; its only seam for this path is
 * the nn.Module surface of src/architecture.py:30-82 plus the small functions next to it.  Every entry point below
 * names the reference interface it replaces.  A maintainer binds it with ctypes.
 *
 * Conventions: plain pointers and sizes, no torch types.  All pointers are DEVICE pointers unless a name ends in
 * _host.  Every function returns 0 (SRB_OK) or a negative SRB_E_* code; srb_last_error() gives the message
 * (thread-local).  Work is enqueued on the given cudaStream_t (passed as void*), no internal synchronisation, no
 * hidden allocation: the caller owns outputs and the workspace.
 */
#ifndef SRB_H_
#define SRB_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRB_OK 0
#define SRB_E_SHAPE (-1)     /* unsupported geometry (full model: H, W multiples of 4; stage 2 alone: of 8) */
#define SRB_E_DTYPE (-2)
#define SRB_E_WORKSPACE (-3) /* workspace / packed-weight buffer too small */
#define SRB_E_CUDA (-4)      /* CUDA runtime / driver error, text in srb_last_error() */
#define SRB_E_ARG (-5)

/* element types of host/device image buffers */
#define SRB_F32 0
#define SRB_BF16 1
#define SRB_U16 2

/* smoothing modes of AntiCheckerboardLayer (reference src/architecture.py:9-27,67-72) */
#define SRB_SMOOTH_NONE 0
#define SRB_SMOOTH_LIGHT 1    /* 3x3 /16   */
#define SRB_SMOOTH_BALANCED 2 /* 5x5 /256  */
#define SRB_SMOOTH_STRONG 3   /* 7x7 /1600 */

int srb_version(void);
const char* srb_last_error(void);

/* ---- model weights -------------------------------------------------------------------------------------------
 * Replaces: nn.Module parameter storage of HybridSuperResolutionModel (state_dict contract).
 * params: array of n device pointers in the canonical order returned by srb_param_name(i) (fp32 tensors, contiguous,
 * PyTorch layout) followed by the two int64 index buffers.  packed receives bf16 / fp32 kernel-ready copies. */
int srb_param_count(void);
const char* srb_param_name(int i);       /* state-dict key, e.g. "stage1.conv_first.weight" */
int64_t srb_param_numel(int i);
size_t srb_packed_bytes(void);
int srb_pack_weights(const void* const* params, int n, void* packed, size_t packed_bytes, void* stream);

/* ---- forward --------------------------------------------------------------------------------------------------
 * Replaces: HybridSuperResolutionModel.forward (src/architecture.py:74-82): stage1 (RRDBNet) -> s1 -> stage2 (HAT)
 * -> s2 -> sf.  x: [B,1,H,W] (x_dtype F32 in [0,1], or U16 which is divided by 65535 like src/dataset.py:34).
 * y: [B,1,4H,4W] F32 (or U16: clip, *65535, truncate like scripts/Modello_5.py:29-33).
 * stage_mask: bit0 run stage 1, bit1 run stage 2 (both set = full model); 7 = full model stopping after s2 (the caller
 * then resizes with srb_resample_separable and applies sf with srb_blur: the output_size branch of :80-82). */
size_t srb_workspace_bytes(int B, int H, int W);
int srb_forward(const void* x, int x_dtype, void* y, int y_dtype, int B, int H, int W, const void* packed,
                void* workspace, size_t workspace_bytes, int smoothing, int stage_mask, void* stream);
/* Profiling variant (bench only): same work, but every launch is bracketed by CUDA events on `stream`; the call
 * synchronises the stream and returns per-class totals.  Classes: 0 conv3x3 (conv_tc_kernel, tcgen05 implicit GEMM),
 * 1 linear (tcgen05), 2 HAB attention branch (hab_tc_kernel: LN + qkv + W-MSA + proj, all tcgen05), 3 layernorm,
 * 4 other, 5 MLP branch (tcgen05), 6 OCAB attention core, 7 RRDB chain kernel (conv_chain_kernel, tcgen05).
 * class_flops = algorithmic FLOPs (2*MACs). */
#define SRB_PROF_CLASSES 8
int srb_forward_profile(const void* x, int x_dtype, void* y, int y_dtype, int B, int H, int W, const void* packed,
                        void* workspace, size_t workspace_bytes, int smoothing, int stage_mask, void* stream,
                        double* class_ms, double* class_flops, int* class_launches);
/* number of kernels one srb_forward enqueues for this geometry (for bench "gpu_launches") */
int srb_forward_launch_count(int B, int H, int W, int smoothing, int stage_mask);

/* ---- granular ops (tests / building blocks) -------------------------------------------------------------------
 * 3x3 (taps=9) or 1x1 (taps=1) NHWC bf16 implicit-GEMM convolution on tcgen05 with the fused epilogue
 *   v = acc + bias; v = act(v); v = v*alpha1 + res1; v = v*alpha2 + res2*res2_cs
 * Replaces: nn.Conv2d(…,3,1,1) / nn.Linear call sites inside RRDBNet / HAT. */
int srb_conv_nhwc_bf16(const void* in, int in_pitch, const void* w_packed, const float* bias, void* out_bf16,
                       int out_pitch, float* out_f32, int out32_pitch, int B, int H, int W, int Cin, int Cout,
                       int taps, int act, float slope, float alpha1, const void* res1, int res1_f32, int res1_pitch,
                       float alpha2, const void* res2, int res2_f32, int res2_pitch, const float* res2_cs,
                       int pixel_shuffle, void* stream);
/* Training-path variant of the same kernel: v = acc + bias; pre_out (bf16, optional) = v; v = act(v) (ACT_GELU = exact-erf
 * GELU); v *= GELU'(gelu_grad_of) (optional: the data-gradient conv that feeds a GELU backward); v = v * alpha + res1 with
 * alpha = alpha1_samples[image] (optional fp32 [B] on the device: the DropPath factor of a residual branch) or 1; then
 * v += res2 * res2_cs[image][channel] (optional bf16 tensor and fp32 [B][Cout] scale: HAB's conv_scale * CAB term).  Keeps
 * the pre-activation the backward needs without a separate GELU pass over HBM in either direction.  Cout % 32 == 0. */
int srb_conv_nhwc_bf16_train(const void* in, int in_pitch, const void* w_packed, const float* bias, void* out_bf16,
                             int out_pitch, int B, int H, int W, int Cin, int Cout, int taps, int act, float slope,
                             const void* res1, int res1_f32, int res1_pitch, void* pre_out, int pre_pitch,
                             const void* gelu_grad_of, int gelu_grad_pitch, const float* alpha1_samples, const void* res2,
                             int res2_pitch, const float* res2_cs, void* stream);
/* Attention branch of a HAB block (upstream hat_arch.py HAB.forward: norm1 -> (shifted) window_partition -> qkv ->
 * softmax(q k^T * head_dim^-0.5 + relative_position_bias + shift mask) v -> proj -> window_reverse -> + shortcut;
 * reached from src/architecture.py:78) as ONE kernel over tokens x: bf16 [B,H,W,96]:
 *   xn = LayerNorm(x) (needed by the parallel CAB branch),  t = x + proj(W-MSA(xn)).
 * window 8, 6 heads, shift 0 or 4 (cyclic shift + the -100 region mask of calculate_mask); H % 8 == 0, W % 16 == 0.
 * wqkv_packed bf16 [288][96], wproj_packed bf16 [96][96] (srb_pack_conv_weight, taps = 1).
 * bias_frag_log2e = srb_expand_rel_bias(table [225][6] fp32, index int64 [64*64] (relative_position_index_SA),
 * NK = 64, scale = log2(e)): the gathered bias in the kernel-ready order of the consuming kernel (opaque to callers). */
int srb_expand_rel_bias(const float* table, int T, const int64_t* index, int NK, void* out_bf16, float scale,
                        void* stream);
int srb_hab_attention(const void* x, const float* ln_g, const float* ln_b, const void* wqkv_packed, const float* bqkv,
                      const void* bias_frag_log2e, const void* wproj_packed, const float* bproj, void* xn, void* t,
                      int B, int H, int W, int shift, void* stream);
/* Token-wise MLP branch of a HAT block (HAB / OCAB; upstream hat_arch.py `x = x + mlp(norm2(x))`, reached from
 * src/architecture.py:78) as ONE tcgen05 kernel:   y0 = t + cab*cs[b,:] (cab/cs may be NULL: y0 = t);
 *   y = y0 + fc2(GELU(fc1(LayerNorm(y0)))).   t, cab, y: bf16 [ntok,96] (y may alias t); cs: fp32 [B,96];
 * w1_folded bf16 [192][96] / b1_folded fp32 [192]: fc1 with the LayerNorm affine folded in (srb_fold_ln_linear);
 * w2_packed bf16 [96][192] (srb_pack_conv_weight, taps=1); ntok and tok_per_img must be multiples of 128.
 * srb_fold_ln_linear: W'[n][k] = W[n][k]*gamma[k] (bf16), b'[n] = b[n] + sum_k W[n][k]*beta[k]; fp32 device inputs,
 * W [N][K] as stored by nn.Linear. */
int srb_fold_ln_linear(const float* w, const float* b, const float* gamma, const float* beta, void* w_folded,
                       float* b_folded, int N, int K, void* stream);
int srb_mlp_branch(const void* t, const void* cab, const float* cs, const void* w1_folded, const float* b1_folded,
                   const void* w2_packed, const float* b2, void* y, size_t ntok, size_t tok_per_img, void* stream);
/* srb_mlp_branch plus a second output xn_out = LayerNorm(y) * ln_g + ln_b (bf16 [ntok,96]): the norm1 of the block that
 * follows (OCAB.norm1 after the last HAB of a residual group), computed from the output rows while they are stored. */
int srb_mlp_branch_ln(const void* t, const void* cab, const float* cs, const void* w1_folded, const float* b1_folded,
                      const void* w2_packed, const float* b2, void* y, const float* ln_g, const float* ln_b,
                      void* xn_out, size_t ntok, size_t tok_per_img, void* stream);
/* fp32 [Cout,Cin,kh,kw] (PyTorch) -> bf16 [kh*kw][Cout][Cin] */
int srb_pack_conv_weight(const float* w, void* w_packed, int Cout, int Cin, int taps, void* stream);

/* ---- loss / metrics / normalisation / stitching --------------------------------------------------------------
 * srb_weighted_l1   replaces CombinedLoss.forward (src/losses.py:9-21): loss[0] = mean(|p-t| * (t>0.02 ? 500 : 1)).
 *                   grad (optional, may be NULL) = dLoss/dpred = sign(p-t)*w/n.  scratch: srb_loss_scratch_bytes().
 * srb_psnr_ssim     replaces Metrics.update / ssim_torch (src/metrics.py:5-32) for fp32 [B,1,H,W] pairs:
 *                   out2[0] = sum_b 10*log10(1/(mse_b+1e-8)) after clamp[0,1];  out2[1] = SSIM batch mean * B.
 * srb_log_normalize_u16 replaces the per-pixel transform of scripts/Dataset_step4_normalization.py:161-179:
 *                   u16(clip((log1p(max(nan_to_num(x),0)) - gmin)/denom, 0, 1)*65535), denom = f32(gmax-gmin+1e-8).
 * srb_blend_tiles / srb_blend_finalize: halo-blend stitch of tile outputs into rows [row0,row1) of a frame
 *                   (whole-field tiled inference; no reference counterpart: Modello_5 iterates patches).
 *                   acc/wsum: fp32 [(row1-row0) x FW], zeroed by the caller.  oy/ox: device int32 tile origins. */
size_t srb_loss_scratch_bytes(void);
int srb_weighted_l1(const float* pred, const float* target, size_t n, float* loss, float* grad, void* scratch,
                    void* stream);
/* srb_charbonnier_astro  the non-VGG terms of the stale loss variant (Backup/src/losses.py:6-15,37-48) in one
 *                   pass: out3 = {char = mean sqrt(d^2 + 1e-6), l1_raw = mean |d|, astro = mean sqrt(d^2 + 1e-6)(1 + 5 t)};
 *                   grad (optional) = d(l1_w char + astro_w astro)/dpred.  scratch: srb_loss_scratch_bytes(). */
int srb_charbonnier_astro(const float* pred, const float* target, size_t n, float l1_w, float astro_w, float* out3,
                          float* grad, void* scratch, void* stream);
size_t srb_psnr_ssim_scratch_bytes(int B, int H, int W);
int srb_psnr_ssim(const float* pred, const float* target, int B, int H, int W, float* out2, void* scratch,
                  void* stream);
/* srb_psnr_ssim_noalloc: srb_psnr_ssim for callers that may not allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same inputs, outputs and results as srb_psnr_ssim; the scratch is a static __device__ array of
 * SRB_PSNR_SSIM_NOALLOC_MAX_BYTES inside the library (held by every GPU that loads it, until exit), so the call makes
 * no cudaMalloc / cudaFree and takes no scratch pointer.  Shapes with srb_psnr_ssim_scratch_bytes(B,H,W) above the
 * limit (e.g. more than 252 pairs of 512x512) are rejected with SRB_E_WORKSPACE.  All calls on one device share the
 * array: calls that may overlap on the GPU (different streams without a dependency) race; serialise them. */
#define SRB_PSNR_SSIM_NOALLOC_MAX_BYTES (256 * 1024)
int srb_psnr_ssim_noalloc(const float* pred, const float* target, int B, int H, int W, float* out2, void* stream);
int srb_log_normalize_u16(const float* x, uint16_t* out, size_t n, float gmin, float denom, void* stream);
/* srb_to_u16          save_as_tiff16's conversion (scripts/Modello_5.py:29-33): u16(clip(x, 0, 1) * 65535), truncating cast
 * srb_preview_strip   the [LR | SR | HR] comparison strip of scripts/Modello_5.py:83-85 / Modello_supporto.py:187-190:
 *                     cat(nearest-upsampled lr [B,h,w] -> [B,H,W], sr, hr; width).clamp(0,1) -> [B,H,3W]; out_u8 = 0: fp32,
 *                     1: uint8 with torchvision save_image's rounding (x*255 + 0.5, clamp, truncate) */
int srb_to_u16(const float* x, uint16_t* out, size_t n, void* stream);
int srb_preview_strip(const float* lr, const float* sr, const float* hr, void* out, int out_u8, int B, int h, int w, int H,
                      int W, void* stream);
/* srb_robust_hist   one level of the exact percentile search behind calculate_robust_stats
 *                   (scripts/Dataset_step4_normalization.py:55-91): over the valid pixels of x (finite, >= thresh > 0)
 *                   ADD into hist (u64, zeroed by the caller) the counts of  level 0: bits[30:21] (1 x 2048 bins);
 *                   level 1: bits[20:10] of the pixels with bits[31:21] == prefixes[j] (nprefix x 2048 bins);
 *                   level 2: bits[9:0] of the pixels with bits[31:10] == prefixes[j].  x 16-byte aligned; 1..4 prefixes. */
int srb_robust_hist(const float* x, size_t n, float thresh, int level, const uint32_t* prefixes, int nprefix,
                    unsigned long long* hist, void* stream);
int srb_blend_tiles(const float* tiles, const int* oy, const int* ox, int T, int th, int tw, int overlap, float* acc,
                    float* wsum, int FH, int FW, int row0, int row1, void* stream);
int srb_blend_finalize(const float* acc, const float* wsum, void* out, int out_dtype, size_t n, void* stream);
/* srb_gather_tiles: out[t] = frame[iy[t] : iy[t]+tile, ix[t] : ix[t]+tile] (frame u16 or f32 [FH,FW]; iy/ix device int32):
 *                   the tile batch of the whole-field driver in one launch. */
int srb_gather_tiles(const void* frame, int dtype, int FH, int FW, const int* iy, const int* ix, int T, int tile, void* out,
                     void* stream);


/* ---- backward of the conv / linear layers (autograd of nn.Conv2d(..,3,1,1) / nn.Linear inside
 * RRDBNet / HAT, reached from scripts/Modello_supporto.py:133) -------------------------------------------------------
 * data gradient:   dX = srb_conv_nhwc_bf16(dY, w = srb_pack_conv_weight_dgrad(W), Cin' = Cout, Cout' = Cin); accumulate
 *                  into an existing gradient with res1 = out (in place).
 * srb_pack_conv_weight_dgrad  fp32 [Cout,Cin,kh,kw] -> bf16 [kh*kw][Cin][Cout] with the taps reversed.
 * srb_conv_wgrad   dW[Cout,Cin,kh,kw] (fp32, PyTorch layout) = sum_px dY[px,co] * X[px+tap,ci]  (+= when accumulate);
 *                  db[Cout] = sum_px dY (optional).  x / dy: bf16 NHWC with channel pitches; W % 32 == 0;
 *                  Cin, Cout multiples of 8.  Deterministic (split-K partials reduced in fixed order).
 *                  scratch: srb_conv_wgrad_scratch_bytes(...) (0 = unsupported geometry). */
int srb_pack_conv_weight_dgrad(const float* w, void* w_packed, int Cout, int Cin, int taps, void* stream);
size_t srb_conv_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout, int taps);
int srb_conv_wgrad(const void* x, int x_pitch, const void* dy, int dy_pitch, float* dw, float* db, void* scratch, int B,
                   int H, int W, int Cin, int Cout, int taps, int accumulate, void* stream);

/* srb_upsample2x_nearest  F.interpolate(scale_factor=2, mode='nearest') of a 64-channel bf16 NHWC tensor
 *                         ([B,h,w,in_pitch] -> dense [B,2h,2w,64]); RRDBNet.forward upsampling (rrdbnet_arch.py).
 * srb_conv3x3_cout1       conv_last of both stages: bf16 [B,H,W,64] -> fp32 [B,H,W], w fp32 [1,64,3,3] (H % 8, W % 32). */
int srb_upsample2x_nearest(const void* in, int in_pitch, void* out, int B, int h, int w, void* stream);
int srb_conv3x3_cout1(const void* in, int in_pitch, const float* w, const float* bias, float* out, int B, int H, int W,
                      void* stream);
/* element-wise backward operators of the RRDBNet training path; tensors are NHWC slices [npx][C] with a channel pitch
 * srb_lrelu_bwd       g *= (y > 0 ? 1 : slope) in place, y = the saved LeakyReLU OUTPUT (bf16 both; C % 8 == 0)
 * srb_axpby           out = a*x + b*y (y may be NULL); each operand bf16 (f32 flag 0) or fp32 (1); out may alias x / y
 * srb_upsample2x_bwd  backward of F.interpolate(scale_factor=2, mode='nearest'): out[b,y,x,:] = sum of the 2x2 block */
int srb_lrelu_bwd(void* g, int g_pitch, const void* y, int y_pitch, size_t npx, int C, float slope, void* stream);
int srb_axpby(void* out, int out_f32, int out_pitch, float a, const void* x, int x_f32, int x_pitch, float b,
              const void* y, int y_f32, int y_pitch, size_t npx, int C, void* stream);
/* srb_axpby_samples   out = a_samples[px / npx_per_sample] * x + b * y: the per-sample DropPath factor of a residual branch
 *                     (upstream drop_path(): keep-mask / keep_prob per sample) fused with the residual add; a_samples on device */
int srb_axpby_samples(void* out, int out_f32, int out_pitch, const float* a_samples, size_t npx_per_sample, const void* x,
                      int x_f32, int x_pitch, float b, const void* y, int y_f32, int y_pitch, size_t npx, int C, void* stream);
int srb_upsample2x_bwd(const void* g, int g_pitch, void* out, int out_pitch, int B, int h, int w, int C, void* stream);

/* srb_blur               AntiCheckerboardLayer.forward (reference src/architecture.py:9-27): depthwise binomial blur with zero
 *                        padding on fp32 [B,H,W]; mode = SRB_SMOOTH_LIGHT (3x3 /16) | BALANCED (5x5 /256) | STRONG (7x7 /1600).
 *                        out_dtype SRB_F32, or SRB_U16 = clip[0,1]*65535 truncated (scripts/Modello_5.py:29-33).  The
 *                        operator is symmetric, so its backward is the same call on the gradient.
 * srb_resample_separable out[b,oy,ox] = sum_{i<K} sum_{j<K} wy[oy*K+i] * wx[ox*K+j] * in[b, iy[oy*K+i], ix[ox*K+j]]  (fp32):
 *                        the bicubic resize branch F.interpolate(mode='bicubic', align_corners=False) of
 *                        src/architecture.py:80-81 with K = 4 and the host-built tap tables, and its backward with the
 *                        transposed tables (K = the largest number of outputs one input feeds).  K <= 64. */
int srb_blur(const float* in, void* out, int out_dtype, int mode, int B, int H, int W, void* stream);
int srb_resample_separable(const float* in, float* out, int B, int Hi, int Wi, int Ho, int Wo, const int* iy,
                           const float* wy, const int* ix, const float* wx, int K, void* stream);

/* HAT (stage 2) backward building blocks (assembled into the stage-2 training path by superresolution_b200/training_hat.py):
 * srb_layernorm96      y = LayerNorm(x)*gamma + beta over 96 channels, eps 1e-5 (nn.LayerNorm of HAT), bf16 [ntok,96]
 * srb_layernorm96_bwd  dx = LN backward of g (+ res, bf16 or fp32, optional: the shortcut's gradient), dgamma / dbeta
 *                      (+= when accumulate); deterministic two-level reduction; scratch: srb_layernorm96_bwd_scratch_bytes()
 * srb_gelu             grad == NULL: out = GELU_erf(pre);  else out = grad * GELU_erf'(pre)  (out may alias grad) */
/* srb_window_attention      (S)W-MSA core of HAB from a saved qkv tensor (bf16 [B,H,W,288] = q|k|v, 6 heads x 16):
 *                           out[B,H,W,96] = softmax(q k^T / 4 + bias[h] (+ shift mask)) v per 8x8 window; shift 0 or 4;
 *                           bias fp32 [6][64][64] = relative_position_bias_table gathered through the index buffer.
 * srb_window_attention_bwd  dqkv[B,H,W,288] from dout[B,H,W,96] (P recomputed) and dtable[225][6] += gradient of the
 *                           relative position bias TABLE (index: int32 [64][64]; atomicAdd: zero it first). */
int srb_window_attention(const void* qkv, const float* bias, void* out, int B, int H, int W, int shift, void* stream);
int srb_window_attention_bwd(const void* qkv, const float* bias, const int* index, const void* dout, void* dqkv,
                             float* dtable, int B, int H, int W, int shift, void* stream);
/* srb_overlap_attention      OCAB attention core from a saved qkv tensor (bf16 [B,H,W,288]): 8x8 query windows attend to the
 *                            12x12 key/value window around them (zero vectors outside the image, still in the softmax);
 *                            bias fp32 [6][64][144] = OCAB relative_position_bias_table gathered through its index buffer.
 * srb_overlap_attention_bwd  dqkv[B,H,W,288] from dout[B,H,W,96]; dK/dV of overlapping windows summed per pixel in fixed
 *                            order; dtable[361][6] += (index: int32 [64][144]; atomicAdd: zero it first).
 *                            scratch: srb_overlap_attention_bwd_scratch_bytes(B, H, W). */
/* srb_overlap_attention_infer  the OCAB core the inference forward (srb_forward) launches: same math, log2 domain.
 *                            qkv: the q part must arrive scaled by head_dim^-0.5 * log2(e) (srb_forward packs the q rows of
 *                            overlap_attn.qkv that way); bias_frag_log2e = srb_expand_rel_bias(table [361][6], index
 *                            int64 [64*144], NK = 144, scale = log2(e)).  out bf16 [B,H,W,96]. */
int srb_overlap_attention_infer(const void* qkv, const void* bias_frag_log2e, void* out, int B, int H, int W, void* stream);
size_t srb_overlap_attention_bwd_scratch_bytes(int B, int H, int W);
int srb_overlap_attention(const void* qkv, const float* bias, void* out, int B, int H, int W, void* stream);
int srb_overlap_attention_bwd(const void* qkv, const float* bias, const int* index, const void* dout, void* dqkv,
                              float* dtable, void* scratch, int B, int H, int W, void* stream);
/* srb_channel_attention      ChannelAttention of CAB: out = y * sigmoid(W2 relu(W1 mean_px(y) + b1) + b2), y bf16 [B,HW,96],
 *                            w1 fp32 [3,96], w2 fp32 [96,3]; saves pooled [B,96], hid [B,3], scale [B,96] for the backward.
 *                            out may be NULL: only the statistics are produced (the caller applies scale[b][c] in a later
 *                            epilogue, e.g. res2_cs of srb_conv_nhwc_bf16_train).
 * srb_channel_attention_bwd  dy (direct + pooled term) and dw1/db1/dw2/db2 (+= when accumulate); dpooled [B,96] scratch out.
 *                            g_scale multiplies the incoming gradient g (HAB's conv_scale 0.01 without a pass over g).
 *                            scratch: srb_channel_attention_scratch_bytes(B, HW).  Deterministic. */
size_t srb_channel_attention_scratch_bytes(int B, int HW);
int srb_channel_attention(const void* y, const float* w1, const float* b1, const float* w2, const float* b2, void* out,
                          float* pooled, float* hid, float* scale, void* scratch, int B, int HW, void* stream);
int srb_channel_attention_bwd(const void* y, const void* g, const float* w1, const float* w2, const float* pooled,
                              const float* hid, const float* scale, void* dy, float* dw1, float* db1, float* dw2,
                              float* db2, float* dpooled, void* scratch, int B, int HW, int accumulate, float g_scale,
                              void* stream);
size_t srb_layernorm96_bwd_scratch_bytes(void);
int srb_layernorm96(const void* x, const float* gamma, const float* beta, void* y, size_t ntok, void* stream);
int srb_layernorm96_bwd(const void* x, const void* g, const float* gamma, const void* res, int res_f32, void* dx,
                        float* dgamma, float* dbeta, void* scratch, size_t ntok, int accumulate, void* stream);
int srb_gelu(void* out, int out_pitch, const void* pre, int pre_pitch, const void* grad, int grad_pitch, size_t npx, int C,
             void* stream);

/* ---- optimizer half of the training step ------------------------------------------------
 * Replaces GradScaler.unscale_ + clip_grad_norm_(1.0) + torch.optim.AdamW.step of scripts/Modello_supporto.py:107,
 * 133-141 on FLAT fp32 buffers (all parameters / gradients / moments of the model back to back).
 * srb_grad_sumsq   info2[0] = sum of squares of grad[0..n), info2[1] = 1 if any element is inf/nan else 0
 *                  (deterministic two-level reduction; scratch: srb_grad_norm_scratch_bytes()).
 * srb_adamw_step   g = grad * coef, coef = inv_scale * min(1, max_norm / (sqrt(info2[0]) * inv_scale + 1e-6))
 *                  (no clipping when max_norm <= 0 or info2 == NULL); the step is skipped when info2[1] != 0;
 *                  then torch.optim.AdamW: p *= 1 - lr*wd; m = lerp(m, g, 1-beta1); v = beta2*v + (1-beta2)*g*g;
 *                  p -= lr/(1-beta1^step) * m / (sqrt(v)/sqrt(1-beta2^step) + eps).  step counts from 1.
 *                  shadow_bf16 (optional): bf16 copy of the updated parameters written in the same pass.
 *                  No host synchronisation; the multi-rank gradient mean is folded into inv_scale (1/world). */
size_t srb_grad_norm_scratch_bytes(void);
int srb_grad_sumsq(const float* grad, size_t n, float* info2, void* scratch, void* stream);
int srb_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr, float beta1,
                   float beta2, float eps, float weight_decay, int step, float inv_scale, float max_norm,
                   const float* info2, void* shadow_bf16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SRB_H_ */
