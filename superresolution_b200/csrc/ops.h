// Internal (C++) declarations of every kernel launcher in libsrb.so.
#pragma once
#include "common.h"

namespace srb {

// api.cu
int pack_conv_weight(const float* w, bf16* o, int Cout, int Cin, int taps, cudaStream_t s);
// fp32 [Cout][Cin] linear weight -> bf16 with rows [0, nrows) multiplied by `scale` before the rounding; bias likewise
int pack_linear_scaled(const float* w, const float* b, bf16* w_out, float* b_out, int Cout, int Cin, int nrows, float scale,
                       cudaStream_t s);

// small_ops.cu
int rrdb_conv_first(const void* x, int x_dtype, const float* w, const float* bias, bf16* out, int out_pitch,
                    float* trunk, float* feat, int B, int h, int w_, cudaStream_t s);
int conv_cout1(const bf16* in, int pitch, const float* w, const float* bias, float* out, int B, int H, int W,
               cudaStream_t s);
int blur(const float* in, void* out, int out_dtype, int mode, int B, int H, int W, cudaStream_t s);
int resample_separable(const float* in, float* out, int B, int Hi, int Wi, int Ho, int Wo, const int* iy, const float* wy,
                       const int* ix, const float* wx, int K, cudaStream_t s);
int convert_out(const float* in, void* out, int out_dtype, size_t n, cudaStream_t s);
int preview_strip(const float* lr, const float* sr, const float* hr, void* out, int out_u8, int B, int h, int w, int H, int W,
                  cudaStream_t s);

// robust_stats.cu
int robust_hist(const float* x, size_t n, float thresh, int level, const uint32_t* prefixes, int nprefix,
                unsigned long long* hist, cudaStream_t s);

// small_ops2.cu
int hat_conv_first_ln(const float* img, const float* w, const float* bias, const float* gamma, const float* beta,
                      bf16* f0, bf16* xn, int B, int H, int W, cudaStream_t s);
int layernorm96(const bf16* in, const float* gamma, const float* beta, bf16* out, size_t ntok, cudaStream_t s);
int combine_ln96(const bf16* t, const bf16* cab, const float* cs, const float* gamma, const float* beta, bf16* y,
                 bf16* xn, size_t ntok, size_t tok_per_img, cudaStream_t s);
int upsample2x(const bf16* in, int in_pitch, bf16* out, int B, int h, int w, cudaStream_t s);
int cab_pool_slabs(int H, int W);  // pool scratch = B * slabs * 96 floats
int cab_channel_scale(const float* pool, int nslab, float* cs, const float* w1, const float* b1, const float* w2,
                      const float* b2, float conv_scale, int B, int HW, cudaStream_t s);

// hab_tc.cu — fused HAB attention branch: xn = LN1(x); t = x + proj((S)W-MSA(xn)) + bproj (one pass over x), the qkv /
// proj GEMMs and the attention core on tcgen05 (TMEM accumulators) over 128-token tiles; bias_frag_log2e:
// expand_rel_bias_bf16(..., NK = 64, scale = log2(e))
int hab_attention_tc(const bf16* x, const float* ln_g, const float* ln_b, const bf16* wqkv, const float* bqkv,
                     const bf16* bias_frag_log2e, const bf16* wproj, const float* bproj, bf16* xn, bf16* t, int B, int H,
                     int W, int shift, cudaStream_t s);
// attn.cu — OCAB: 8x8 query windows x 12x12 zero-padded key/value windows (NK = 144), no shift, no mask.  Log2 domain:
// the q part of qkv arrives scaled by head_dim^-0.5 * log2(e) (scale_rows at pack time), bias_frag holds bias * log2(e).
int ocab_attention(const bf16* qkv, const bf16* bias_frag, bf16* out, int B, int H, int W, cudaStream_t s);
// table fp32 [T][6], index int64 [64*NK] (python-style negative wrap) -> kernel-ready bias [6*64*NK]: [6][8][64][8] operand planes
// for NK = 64 (hab_tc.cu), mma.sync C-fragment order for NK = 144 (attn.cu)
int expand_rel_bias_bf16(const float* table, int T, const int64_t* index, int NK, bf16* out, float scale,
                         cudaStream_t s);  // out = table[index] * scale
int expand_rel_bias_f32(const float* table, int T, const int64_t* index, int NK, float* out, cudaStream_t s);

// mlp_tc.cu:  y0 = t + cab*cs (cab may be null);  y = y0 + fc2(GELU(fc1(LayerNorm(y0))))   (y may alias t)
// w1f / b1f: fc1 with the LayerNorm affine folded in (fold_ln_linear)
int mlp_fused(const bf16* t, const bf16* cab, const float* cs, const bf16* w1f, const float* b1f, const bf16* w2,
              const float* b2, bf16* y, size_t ntok, size_t tok_per_img, cudaStream_t s,
              const float* ln2_g = nullptr, const float* ln2_b = nullptr, bf16* xn2 = nullptr);
// W'[n][k] = W[n][k]*gamma[k] -> bf16, b'[n] = b[n] + sum_k W[n][k]*beta[k]   (fp32 inputs)
int fold_ln_linear(const float* w, const float* b, const float* gamma, const float* beta, bf16* w_out, float* b_out,
                   int N, int K, cudaStream_t s);

// metrics.cu
size_t reduce_scratch_bytes();
int weighted_l1(const float* pred, const float* target, size_t n, float* loss, float* grad, float* scratch,
                cudaStream_t s);
int charbonnier_astro(const float* pred, const float* target, size_t n, float l1_w, float astro_w, float* out3, float* grad,
                      float* scratch, cudaStream_t s);
size_t psnr_ssim_scratch_bytes(int B, int H, int W);
int psnr_ssim(const float* pred, const float* target, int B, int H, int W, float* out2, float* scratch,
              cudaStream_t s);
int psnr_ssim_noalloc(const float* pred, const float* target, int B, int H, int W, float* out2, cudaStream_t s);
int log_normalize_u16(const float* x, uint16_t* out, size_t n, float gmin, float denom, cudaStream_t s);

// conv_bwd.cu
int pack_conv_weight_dgrad(const float* w, bf16* o, int Cout, int Cin, int taps, cudaStream_t s);
// wgrad_tc.cu: the tcgen05 weight-gradient kernel behind conv_wgrad (0 splits = shape not covered)
int wgrad_tc_splits(int B, int H, int W, int Cin, int Cout, int taps);
int wgrad_tc_launch(const bf16* x, int x_pitch, const bf16* dy, int dy_pitch, float* partial, float* bpartial, int ksplit,
                    int B, int H, int W, int Cin, int Cout, int taps, cudaStream_t s);
size_t conv_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout, int taps);
int conv_wgrad(const bf16* x, int x_pitch, const bf16* dy, int dy_pitch, float* dw, float* db, void* scratch, int B,
               int H, int W, int Cin, int Cout, int taps, int accumulate, cudaStream_t s);

// train_ops.cu
int lrelu_bwd(bf16* g, int g_pitch, const bf16* y, int y_pitch, size_t npx, int C, float slope, cudaStream_t s);
int axpby(void* out, int out_f32, int out_pitch, float a, const void* x, int x_f32, int x_pitch, float b, const void* y,
          int y_f32, int y_pitch, size_t npx, int C, cudaStream_t s, const float* a_vec = nullptr,
          size_t npx_per_sample = 0);
int upsample2x_bwd(const bf16* g, int g_pitch, bf16* out, int out_pitch, int B, int h, int w, int C, cudaStream_t s);

size_t layernorm96_bwd_scratch_bytes();
int layernorm96_bwd(const bf16* x, const bf16* g, const float* gamma, const void* res, int res_f32, bf16* dx,
                    float* dgamma, float* dbeta, void* scratch, size_t ntok, int accumulate, cudaStream_t s);
int gelu_fwd_bwd(bf16* out, int out_pitch, const bf16* pre, int pre_pitch, const bf16* g, int g_pitch, size_t npx, int C,
                 cudaStream_t s);

// win_attn.cu
int win_attn_fwd(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, int shift, cudaStream_t s);
int win_attn_bwd(const bf16* qkv, const float* bias, const int* index, const bf16* dout, bf16* dqkv, float* dtable, int B,
                 int H, int W, int shift, cudaStream_t s);

// ocab_attn_train.cu
size_t ocab_attn_bwd_scratch_bytes(int B, int H, int W);
int ocab_attn_fwd(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, cudaStream_t s);
int ocab_attn_bwd(const bf16* qkv, const float* bias, const int* index, const bf16* dout, bf16* dqkv, float* dtable,
                  void* scratch, int B, int H, int W, cudaStream_t s);

// chan_attn.cu
size_t chan_attn_scratch_bytes(int B, int HW);
int chan_attn_fwd(const bf16* y, const float* w1, const float* b1, const float* w2, const float* b2, bf16* out,
                  float* pooled, float* hid, float* s, void* scratch, int B, int HW, cudaStream_t st);
int chan_attn_bwd(const bf16* y, const bf16* g, const float* w1, const float* w2, const float* pooled, const float* hid,
                  const float* s, bf16* dy, float* dw1, float* db1, float* dw2, float* db2, float* dpooled,
                  void* scratch, int B, int HW, int accumulate, float g_scale, cudaStream_t st);

// optim.cu
size_t grad_norm_scratch_bytes();
int grad_sumsq(const float* g, size_t n, float* info, void* scratch, cudaStream_t s);
int adamw_step(float* p, const float* g, float* m, float* v, size_t n, float lr, float beta1, float beta2, float eps,
               float weight_decay, int step, float inv_scale, float max_norm, const float* info, void* shadow_bf16,
               cudaStream_t s);

// blend.cu
int blend_tiles(const float* tiles, const int* oy, const int* ox, int T, int th, int tw, int ov, float* acc,
                float* wsum, int FH, int FW, int row0, int row1, cudaStream_t s);
int blend_finalize(const float* acc, const float* wsum, void* out, int out_dtype, size_t n, cudaStream_t s);
int gather_tiles(const void* frame, int dtype, int FH, int FW, const int* iy, const int* ix, int T, int tile, void* out,
                 cudaStream_t s);

// forward.cu
int model_param_count();
const char* model_param_name(int i);
int64_t model_param_numel(int i);
size_t model_packed_bytes();
int model_pack(const void* const* params, int n, void* packed, size_t packed_bytes, cudaStream_t s);
size_t model_workspace_bytes(int B, int H, int W);
int model_forward(const void* x, int x_dtype, void* y, int y_dtype, int B, int H, int W, const void* packed,
                  void* workspace, size_t workspace_bytes, int smoothing, int stage_mask, cudaStream_t s, bool dry,
                  int* launches, double* prof_ms = nullptr, double* prof_flops = nullptr, int* prof_counts = nullptr);

}  // namespace srb
