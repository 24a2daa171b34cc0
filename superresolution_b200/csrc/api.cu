// extern "C" surface of libsrb.so (declared in include/srb.h).  Thin: argument checks + launches.
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "common.h"
#include "ops.h"

namespace srb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return SRB_E_CUDA;
}

// fp32 [Cout][Cin][taps] (PyTorch conv weight, taps = kh*kw contiguous) -> bf16 [taps][Cout][Cin]
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, bf16* __restrict__ o, int Cout, int Cin,
                                        int taps) {
    const int n = Cout * Cin * taps;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int ci = i % Cin, co = (i / Cin) % Cout, t = i / (Cin * Cout);
        o[i] = __float2bfloat16_rn(w[((size_t)co * Cin + ci) * taps + t]);
    }
}

int pack_conv_weight(const float* w, bf16* o, int Cout, int Cin, int taps, cudaStream_t s) {
    const int n = Cout * Cin * taps;
    pack_conv_weight_kernel<<<(n + 255) / 256, 256, 0, s>>>(w, o, Cout, Cin, taps);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

__global__ void pack_linear_scaled_kernel(const float* __restrict__ w, const float* __restrict__ b, bf16* __restrict__ wo,
                                          float* __restrict__ bo, int Cout, int Cin, int nrows, float scale) {
    const int n = Cout * Cin;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int co = i / Cin;
        wo[i] = __float2bfloat16_rn(co < nrows ? w[i] * scale : w[i]);
        if (i < Cout) bo[i] = i < nrows ? b[i] * scale : b[i];
    }
}

int pack_linear_scaled(const float* w, const float* b, bf16* w_out, float* b_out, int Cout, int Cin, int nrows, float scale,
                       cudaStream_t s) {
    const int n = Cout * Cin;
    pack_linear_scaled_kernel<<<(n + 255) / 256, 256, 0, s>>>(w, b, w_out, b_out, Cout, Cin, nrows, scale);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb

using namespace srb;
namespace srb { extern long long* g_fa_dbg; }

extern "C" {

int srb_version(void) { return 100; }
const char* srb_last_error(void) { return g_err; }

int srb_pack_conv_weight(const float* w, void* w_packed, int Cout, int Cin, int taps, void* stream) {
    if (!w || !w_packed || Cout <= 0 || Cin <= 0 || taps <= 0) {
        set_error("pack_conv_weight: bad args");
        return SRB_E_ARG;
    }
    return pack_conv_weight(w, (bf16*)w_packed, Cout, Cin, taps, (cudaStream_t)stream);
}

int srb_conv_nhwc_bf16(const void* in, int in_pitch, const void* w_packed, const float* bias, void* out_bf16,
                       int out_pitch, float* out_f32, int out32_pitch, int B, int H, int W, int Cin, int Cout,
                       int taps, int act, float slope, float alpha1, const void* res1, int res1_f32, int res1_pitch,
                       float alpha2, const void* res2, int res2_f32, int res2_pitch, const float* res2_cs,
                       int pixel_shuffle, void* stream) {
    if (!in || !w_packed || (!out_bf16 && !out_f32) || B <= 0) {
        set_error("conv: bad args");
        return SRB_E_ARG;
    }
    ConvArgs a;
    memset(&a, 0, sizeof(a));
    a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.Cout = Cout; a.taps = taps;
    a.in = (const bf16*)in; a.in_pitch = in_pitch; a.w = (const bf16*)w_packed; a.bias = bias;
    a.act = act; a.slope = slope;
    a.alpha1 = alpha1; a.res1 = res1; a.res1_f32 = res1_f32; a.res1_pitch = res1_pitch;
    a.alpha2 = alpha2; a.res2 = res2; a.res2_f32 = res2_f32; a.res2_pitch = res2_pitch; a.res2_cs = res2_cs;
    a.out = (bf16*)out_bf16; a.out_pitch = out_pitch; a.out32 = out_f32; a.out32_pitch = out32_pitch;
    a.pixel_shuffle = pixel_shuffle;
    a.dyn_weights = 1;  // granular entry: the caller may have packed the weights just before on the same stream
    return conv_tc_launch(a, (cudaStream_t)stream);
}

int srb_conv_nhwc_bf16_train(const void* in, int in_pitch, const void* w_packed, const float* bias, void* out_bf16,
                             int out_pitch, int B, int H, int W, int Cin, int Cout, int taps, int act, float slope,
                             const void* res1, int res1_f32, int res1_pitch, void* pre_out, int pre_pitch,
                             const void* gelu_grad_of, int gelu_grad_pitch, const float* alpha1_samples, const void* res2,
                             int res2_pitch, const float* res2_cs, void* stream) {
    if (!in || !w_packed || !out_bf16 || B <= 0) { set_error("conv_train: bad args"); return SRB_E_ARG; }
    if (alpha1_samples && !res1) { set_error("conv_train: alpha1_samples scales the conv result in front of res1"); return SRB_E_ARG; }
    if (Cout % 32 || ((uintptr_t)pre_out & 15) || ((uintptr_t)gelu_grad_of & 15) || pre_pitch % 8 || gelu_grad_pitch % 8) {
        set_error("conv_train: Cout %% 32, 16-byte aligned pre_out / gelu_grad_of with pitches %% 8 required");
        return SRB_E_ARG;
    }
    ConvArgs a;
    memset(&a, 0, sizeof(a));
    a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.Cout = Cout; a.taps = taps;
    a.in = (const bf16*)in; a.in_pitch = in_pitch; a.w = (const bf16*)w_packed; a.bias = bias;
    a.act = act; a.slope = slope;
    a.alpha1 = 1.f; a.res1 = res1; a.res1_f32 = res1_f32; a.res1_pitch = res1_pitch; a.alpha2 = 1.f;
    a.out = (bf16*)out_bf16; a.out_pitch = out_pitch;
    a.pre_out = (bf16*)pre_out; a.pre_pitch = pre_pitch;
    a.gelu_grad_of = (const bf16*)gelu_grad_of; a.gelu_grad_pitch = gelu_grad_pitch;
    a.alpha1_samples = alpha1_samples;
    a.res2 = res2; a.res2_f32 = 0; a.res2_pitch = res2_pitch; a.res2_cs = res2_cs;
    a.dyn_weights = 1;
    return conv_tc_launch(a, (cudaStream_t)stream);
}

int srb_expand_rel_bias(const float* table, int T, const int64_t* index, int NK, void* out_bf16, float scale,
                        void* stream) {
    if (!table || !index || !out_bf16 || T <= 0 || (NK != 64 && NK != 144)) {
        set_error("expand_rel_bias: bad args (NK must be 64 or 144)");
        return SRB_E_ARG;
    }
    return expand_rel_bias_bf16(table, T, index, NK, (bf16*)out_bf16, scale, (cudaStream_t)stream);
}

int srb_hab_attention(const void* x, const float* ln_g, const float* ln_b, const void* wqkv_packed, const float* bqkv,
                      const void* bias_frag_log2e, const void* wproj_packed, const float* bproj, void* xn, void* t,
                      int B, int H, int W, int shift, void* stream) {
    if (!x || !ln_g || !ln_b || !wqkv_packed || !bqkv || !bias_frag_log2e || !wproj_packed || !bproj || !xn || !t ||
        B <= 0 || (shift != 0 && shift != 4)) {
        set_error("hab_attention: bad args (shift must be 0 or 4)");
        return SRB_E_ARG;
    }
    return hab_attention_tc((const bf16*)x, ln_g, ln_b, (const bf16*)wqkv_packed, bqkv, (const bf16*)bias_frag_log2e,
                            (const bf16*)wproj_packed, bproj, (bf16*)xn, (bf16*)t, B, H, W, shift, (cudaStream_t)stream);
}

int srb_fold_ln_linear(const float* w, const float* b, const float* gamma, const float* beta, void* w_folded,
                       float* b_folded, int N, int K, void* stream) {
    if (!w || !b || !gamma || !beta || !w_folded || !b_folded || N <= 0 || K <= 0) {
        set_error("fold_ln_linear: bad args");
        return SRB_E_ARG;
    }
    return fold_ln_linear(w, b, gamma, beta, (bf16*)w_folded, b_folded, N, K, (cudaStream_t)stream);
}

int srb_mlp_branch(const void* t, const void* cab, const float* cs, const void* w1_folded, const float* b1_folded,
                   const void* w2_packed, const float* b2, void* y, size_t ntok, size_t tok_per_img, void* stream) {
    if (!t || !w1_folded || !b1_folded || !w2_packed || !b2 || !y || ntok == 0 || (cab && !cs)) {
        set_error("mlp_branch: bad args");
        return SRB_E_ARG;
    }
    return mlp_fused((const bf16*)t, (const bf16*)cab, cs, (const bf16*)w1_folded, b1_folded, (const bf16*)w2_packed,
                     b2, (bf16*)y, ntok, tok_per_img, (cudaStream_t)stream);
}
int srb_mlp_branch_ln(const void* t, const void* cab, const float* cs, const void* w1_folded, const float* b1_folded,
                      const void* w2_packed, const float* b2, void* y, const float* ln_g, const float* ln_b,
                      void* xn_out, size_t ntok, size_t tok_per_img, void* stream) {
    if (!t || !w1_folded || !b1_folded || !w2_packed || !b2 || !y || ntok == 0 || (cab && !cs) || !ln_g || !ln_b ||
        !xn_out) {
        set_error("mlp_branch_ln: bad args");
        return SRB_E_ARG;
    }
    return mlp_fused((const bf16*)t, (const bf16*)cab, cs, (const bf16*)w1_folded, b1_folded, (const bf16*)w2_packed,
                     b2, (bf16*)y, ntok, tok_per_img, (cudaStream_t)stream, ln_g, ln_b, (bf16*)xn_out);
}

int srb_param_count(void) { return model_param_count(); }
const char* srb_param_name(int i) { return model_param_name(i); }
int64_t srb_param_numel(int i) { return model_param_numel(i); }
size_t srb_packed_bytes(void) { return model_packed_bytes(); }
int srb_pack_weights(const void* const* params, int n, void* packed, size_t packed_bytes, void* stream) {
    if (!params || !packed) { set_error("pack_weights: null pointer"); return SRB_E_ARG; }
    return model_pack(params, n, packed, packed_bytes, (cudaStream_t)stream);
}
size_t srb_workspace_bytes(int B, int H, int W) {
    if (B <= 0 || H <= 0 || W <= 0) return 0;
    return model_workspace_bytes(B, H, W);
}
int srb_forward(const void* x, int x_dtype, void* y, int y_dtype, int B, int H, int W, const void* packed,
                void* workspace, size_t workspace_bytes, int smoothing, int stage_mask, void* stream) {
    return model_forward(x, x_dtype, y, y_dtype, B, H, W, packed, workspace, workspace_bytes, smoothing, stage_mask,
                         (cudaStream_t)stream, false, nullptr);
}
int srb_forward_profile(const void* x, int x_dtype, void* y, int y_dtype, int B, int H, int W, const void* packed,
                        void* workspace, size_t workspace_bytes, int smoothing, int stage_mask, void* stream,
                        double* class_ms, double* class_flops, int* class_launches) {
    if (!class_ms || !class_flops || !class_launches) { set_error("forward_profile: null output"); return SRB_E_ARG; }
    return model_forward(x, x_dtype, y, y_dtype, B, H, W, packed, workspace, workspace_bytes, smoothing, stage_mask,
                         (cudaStream_t)stream, false, nullptr, class_ms, class_flops, class_launches);
}
int srb_forward_launch_count(int B, int H, int W, int smoothing, int stage_mask) {
    int n = 0;
    int rc = model_forward(nullptr, SRB_F32, nullptr, SRB_F32, B, H, W, nullptr, nullptr, 0, smoothing, stage_mask, 0,
                           true, &n);
    return rc == SRB_OK ? n : rc;
}

size_t srb_loss_scratch_bytes(void) { return reduce_scratch_bytes(); }
int srb_weighted_l1(const float* pred, const float* target, size_t n, float* loss, float* grad, void* scratch,
                    void* stream) {
    if (!pred || !target || !loss || !scratch || n == 0) { set_error("weighted_l1: bad args"); return SRB_E_ARG; }
    return weighted_l1(pred, target, n, loss, grad, (float*)scratch, (cudaStream_t)stream);
}
int srb_charbonnier_astro(const float* pred, const float* target, size_t n, float l1_w, float astro_w, float* out3,
                          float* grad, void* scratch, void* stream) {
    if (!pred || !target || !out3 || !scratch || n == 0) { set_error("charbonnier_astro: bad args"); return SRB_E_ARG; }
    return charbonnier_astro(pred, target, n, l1_w, astro_w, out3, grad, (float*)scratch, (cudaStream_t)stream);
}
size_t srb_psnr_ssim_scratch_bytes(int B, int H, int W) { return psnr_ssim_scratch_bytes(B, H, W); }
int srb_psnr_ssim(const float* pred, const float* target, int B, int H, int W, float* out2, void* scratch,
                  void* stream) {
    if (!pred || !target || !out2 || !scratch || B <= 0) { set_error("psnr_ssim: bad args"); return SRB_E_ARG; }
    return psnr_ssim(pred, target, B, H, W, out2, (float*)scratch, (cudaStream_t)stream);
}
int srb_psnr_ssim_noalloc(const float* pred, const float* target, int B, int H, int W, float* out2, void* stream) {
    if (!pred || !target || !out2) { set_error("psnr_ssim_noalloc: bad args"); return SRB_E_ARG; }
    return psnr_ssim_noalloc(pred, target, B, H, W, out2, (cudaStream_t)stream);
}
int srb_pack_conv_weight_dgrad(const float* w, void* w_packed, int Cout, int Cin, int taps, void* stream) {
    if (!w || !w_packed || Cout <= 0 || Cin <= 0 || (taps != 9 && taps != 1)) { set_error("pack_conv_weight_dgrad: bad args"); return SRB_E_ARG; }
    return pack_conv_weight_dgrad(w, (bf16*)w_packed, Cout, Cin, taps, (cudaStream_t)stream);
}
size_t srb_conv_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout, int taps) {
    return conv_wgrad_scratch_bytes(B, H, W, Cin, Cout, taps);
}
int srb_conv_wgrad(const void* x, int x_pitch, const void* dy, int dy_pitch, float* dw, float* db, void* scratch, int B,
                   int H, int W, int Cin, int Cout, int taps, int accumulate, void* stream) {
    if (!x || !dy || !dw || !scratch || B <= 0 || H <= 0 || W <= 0) { set_error("conv_wgrad: bad args"); return SRB_E_ARG; }
    return conv_wgrad((const bf16*)x, x_pitch, (const bf16*)dy, dy_pitch, dw, db, scratch, B, H, W, Cin, Cout, taps,
                      accumulate, (cudaStream_t)stream);
}
int srb_upsample2x_nearest(const void* in, int in_pitch, void* out, int B, int h, int w, void* stream) {
    if (!in || !out || B <= 0 || h <= 0 || w <= 0 || in_pitch % 8) { set_error("upsample2x_nearest: bad args"); return SRB_E_ARG; }
    return upsample2x((const bf16*)in, in_pitch, (bf16*)out, B, h, w, (cudaStream_t)stream);
}
int srb_conv3x3_cout1(const void* in, int in_pitch, const float* w, const float* bias, float* out, int B, int H, int W,
                      void* stream) {
    if (!in || !w || !bias || !out || B <= 0 || in_pitch % 8) { set_error("conv3x3_cout1: bad args"); return SRB_E_ARG; }
    return conv_cout1((const bf16*)in, in_pitch, w, bias, out, B, H, W, (cudaStream_t)stream);
}
int srb_to_u16(const float* x, uint16_t* out, size_t n, void* stream) {
    if (!x || !out || n == 0) { set_error("to_u16: bad args"); return SRB_E_ARG; }
    return convert_out(x, out, SRB_U16, n, (cudaStream_t)stream);
}
int srb_preview_strip(const float* lr, const float* sr, const float* hr, void* out, int out_u8, int B, int h, int w, int H,
                      int W, void* stream) {
    if (!lr || !sr || !hr || !out || B <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) { set_error("preview_strip: bad args"); return SRB_E_ARG; }
    return preview_strip(lr, sr, hr, out, out_u8, B, h, w, H, W, (cudaStream_t)stream);
}
int srb_robust_hist(const float* x, size_t n, float thresh, int level, const uint32_t* prefixes, int nprefix,
                    unsigned long long* hist, void* stream) {
    if (!x || !hist || n == 0) { set_error("robust_hist: bad args"); return SRB_E_ARG; }
    return robust_hist(x, n, thresh, level, prefixes, nprefix, hist, (cudaStream_t)stream);
}
int srb_blur(const float* in, void* out, int out_dtype, int mode, int B, int H, int W, void* stream) {
    if (!in || !out || in == out || B <= 0 || H <= 0 || W <= 0) { set_error("blur: bad args"); return SRB_E_ARG; }
    if (out_dtype != SRB_F32 && out_dtype != SRB_U16) { set_error("blur: out_dtype must be f32 or u16"); return SRB_E_DTYPE; }
    return blur(in, out, out_dtype, mode, B, H, W, (cudaStream_t)stream);
}
int srb_resample_separable(const float* in, float* out, int B, int Hi, int Wi, int Ho, int Wo, const int* iy,
                           const float* wy, const int* ix, const float* wx, int K, void* stream) {
    if (!in || !out || !iy || !wy || !ix || !wx || B <= 0 || Hi <= 0 || Wi <= 0 || Ho <= 0 || Wo <= 0 || K <= 0 || K > 64) {
        set_error("resample_separable: bad args");
        return SRB_E_ARG;
    }
    return resample_separable(in, out, B, Hi, Wi, Ho, Wo, iy, wy, ix, wx, K, (cudaStream_t)stream);
}
int srb_lrelu_bwd(void* g, int g_pitch, const void* y, int y_pitch, size_t npx, int C, float slope, void* stream) {
    if (!g || !y || npx == 0 || C <= 0) { set_error("lrelu_bwd: bad args"); return SRB_E_ARG; }
    return lrelu_bwd((bf16*)g, g_pitch, (const bf16*)y, y_pitch, npx, C, slope, (cudaStream_t)stream);
}
int srb_axpby(void* out, int out_f32, int out_pitch, float a, const void* x, int x_f32, int x_pitch, float b,
              const void* y, int y_f32, int y_pitch, size_t npx, int C, void* stream) {
    if (!out || !x || npx == 0 || C <= 0) { set_error("axpby: bad args"); return SRB_E_ARG; }
    return axpby(out, out_f32, out_pitch, a, x, x_f32, x_pitch, b, y, y_f32, y_pitch, npx, C, (cudaStream_t)stream);
}
int srb_axpby_samples(void* out, int out_f32, int out_pitch, const float* a_samples, size_t npx_per_sample, const void* x,
                      int x_f32, int x_pitch, float b, const void* y, int y_f32, int y_pitch, size_t npx, int C, void* stream) {
    if (!out || !x || !a_samples || npx == 0 || C <= 0) { set_error("axpby_samples: bad args"); return SRB_E_ARG; }
    return axpby(out, out_f32, out_pitch, 1.f, x, x_f32, x_pitch, b, y, y_f32, y_pitch, npx, C, (cudaStream_t)stream, a_samples,
                 npx_per_sample);
}
int srb_upsample2x_bwd(const void* g, int g_pitch, void* out, int out_pitch, int B, int h, int w, int C, void* stream) {
    if (!g || !out || B <= 0 || h <= 0 || w <= 0) { set_error("upsample2x_bwd: bad args"); return SRB_E_ARG; }
    return upsample2x_bwd((const bf16*)g, g_pitch, (bf16*)out, out_pitch, B, h, w, C, (cudaStream_t)stream);
}
size_t srb_layernorm96_bwd_scratch_bytes(void) { return layernorm96_bwd_scratch_bytes(); }
int srb_layernorm96(const void* x, const float* gamma, const float* beta, void* y, size_t ntok, void* stream) {
    if (!x || !gamma || !beta || !y || ntok == 0) { set_error("layernorm96: bad args"); return SRB_E_ARG; }
    return layernorm96((const bf16*)x, gamma, beta, (bf16*)y, ntok, (cudaStream_t)stream);
}
int srb_layernorm96_bwd(const void* x, const void* g, const float* gamma, const void* res, int res_f32, void* dx,
                        float* dgamma, float* dbeta, void* scratch, size_t ntok, int accumulate, void* stream) {
    if (!x || !g || !gamma || !dx || !dgamma || !dbeta || !scratch || ntok == 0) { set_error("layernorm96_bwd: bad args"); return SRB_E_ARG; }
    return layernorm96_bwd((const bf16*)x, (const bf16*)g, gamma, res, res_f32, (bf16*)dx, dgamma, dbeta, scratch, ntok,
                           accumulate, (cudaStream_t)stream);
}
int srb_gelu(void* out, int out_pitch, const void* pre, int pre_pitch, const void* grad, int grad_pitch, size_t npx, int C,
             void* stream) {
    if (!out || !pre || npx == 0 || C <= 0) { set_error("gelu: bad args"); return SRB_E_ARG; }
    return gelu_fwd_bwd((bf16*)out, out_pitch, (const bf16*)pre, pre_pitch, (const bf16*)grad, grad_pitch, npx, C,
                        (cudaStream_t)stream);
}
int srb_window_attention(const void* qkv, const float* bias, void* out, int B, int H, int W, int shift, void* stream) {
    if (!qkv || !bias || !out) { set_error("window_attention: bad args"); return SRB_E_ARG; }
    return win_attn_fwd((const bf16*)qkv, bias, (bf16*)out, B, H, W, shift, (cudaStream_t)stream);
}
int srb_window_attention_bwd(const void* qkv, const float* bias, const int* index, const void* dout, void* dqkv,
                             float* dtable, int B, int H, int W, int shift, void* stream) {
    if (!qkv || !bias || !index || !dout || !dqkv || !dtable) { set_error("window_attention_bwd: bad args"); return SRB_E_ARG; }
    return win_attn_bwd((const bf16*)qkv, bias, index, (const bf16*)dout, (bf16*)dqkv, dtable, B, H, W, shift,
                        (cudaStream_t)stream);
}
size_t srb_overlap_attention_bwd_scratch_bytes(int B, int H, int W) { return ocab_attn_bwd_scratch_bytes(B, H, W); }
int srb_overlap_attention_infer(const void* qkv, const void* bias_frag_log2e, void* out, int B, int H, int W, void* stream) {
    if (!qkv || !bias_frag_log2e || !out || B <= 0) { set_error("overlap_attention_infer: bad args"); return SRB_E_ARG; }
    return ocab_attention((const bf16*)qkv, (const bf16*)bias_frag_log2e, (bf16*)out, B, H, W, (cudaStream_t)stream);
}
int srb_overlap_attention(const void* qkv, const float* bias, void* out, int B, int H, int W, void* stream) {
    if (!qkv || !bias || !out) { set_error("overlap_attention: bad args"); return SRB_E_ARG; }
    return ocab_attn_fwd((const bf16*)qkv, bias, (bf16*)out, B, H, W, (cudaStream_t)stream);
}
int srb_overlap_attention_bwd(const void* qkv, const float* bias, const int* index, const void* dout, void* dqkv,
                              float* dtable, void* scratch, int B, int H, int W, void* stream) {
    if (!qkv || !bias || !index || !dout || !dqkv || !dtable || !scratch) { set_error("overlap_attention_bwd: bad args"); return SRB_E_ARG; }
    return ocab_attn_bwd((const bf16*)qkv, bias, index, (const bf16*)dout, (bf16*)dqkv, dtable, scratch, B, H, W,
                         (cudaStream_t)stream);
}
size_t srb_channel_attention_scratch_bytes(int B, int HW) { return chan_attn_scratch_bytes(B, HW); }
int srb_channel_attention(const void* y, const float* w1, const float* b1, const float* w2, const float* b2, void* out,
                          float* pooled, float* hid, float* scale, void* scratch, int B, int HW, void* stream) {
    if (!y || !w1 || !b1 || !w2 || !b2 || !pooled || !hid || !scale || !scratch || B <= 0 || HW <= 0) {  // out may be NULL
        set_error("channel_attention: bad args");
        return SRB_E_ARG;
    }
    return chan_attn_fwd((const bf16*)y, w1, b1, w2, b2, (bf16*)out, pooled, hid, scale, scratch, B, HW, (cudaStream_t)stream);
}
int srb_channel_attention_bwd(const void* y, const void* g, const float* w1, const float* w2, const float* pooled,
                              const float* hid, const float* scale, void* dy, float* dw1, float* db1, float* dw2,
                              float* db2, float* dpooled, void* scratch, int B, int HW, int accumulate, float g_scale,
                              void* stream) {
    if (!y || !g || !w1 || !w2 || !pooled || !hid || !scale || !dy || !dw1 || !db1 || !dw2 || !db2 || !dpooled ||
        !scratch || B <= 0 || HW <= 0) {
        set_error("channel_attention_bwd: bad args");
        return SRB_E_ARG;
    }
    return chan_attn_bwd((const bf16*)y, (const bf16*)g, w1, w2, pooled, hid, scale, (bf16*)dy, dw1, db1, dw2, db2, dpooled,
                         scratch, B, HW, accumulate, g_scale, (cudaStream_t)stream);
}
size_t srb_grad_norm_scratch_bytes(void) { return grad_norm_scratch_bytes(); }
int srb_grad_sumsq(const float* grad, size_t n, float* info2, void* scratch, void* stream) {
    if (!grad || !info2 || !scratch || n == 0) { set_error("grad_sumsq: bad args"); return SRB_E_ARG; }
    return grad_sumsq(grad, n, info2, scratch, (cudaStream_t)stream);
}
int srb_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr, float beta1,
                   float beta2, float eps, float weight_decay, int step, float inv_scale, float max_norm,
                   const float* info2, void* shadow_bf16, void* stream) {
    if (!param || !grad || !exp_avg || !exp_avg_sq || n == 0) { set_error("adamw_step: bad args"); return SRB_E_ARG; }
    return adamw_step(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step, inv_scale,
                      max_norm, info2, shadow_bf16, (cudaStream_t)stream);
}
int srb_log_normalize_u16(const float* x, uint16_t* out, size_t n, float gmin, float denom, void* stream) {
    if (n == 0) return SRB_OK;  // empty input is legal (reference: empty array in, empty array out)
    if (!x || !out) { set_error("log_normalize: null pointer"); return SRB_E_ARG; }
    return log_normalize_u16(x, out, n, gmin, denom, (cudaStream_t)stream);
}
int srb_blend_tiles(const float* tiles, const int* oy, const int* ox, int T, int th, int tw, int overlap, float* acc,
                    float* wsum, int FH, int FW, int row0, int row1, void* stream) {
    if (!tiles || !oy || !ox || !acc || !wsum || T <= 0) { set_error("blend_tiles: bad args"); return SRB_E_ARG; }
    return blend_tiles(tiles, oy, ox, T, th, tw, overlap, acc, wsum, FH, FW, row0, row1, (cudaStream_t)stream);
}
int srb_gather_tiles(const void* frame, int dtype, int FH, int FW, const int* iy, const int* ix, int T, int tile, void* out,
                     void* stream) {
    if (!frame || !iy || !ix || !out || T <= 0) { set_error("gather_tiles: bad args"); return SRB_E_ARG; }
    return gather_tiles(frame, dtype, FH, FW, iy, ix, T, tile, out, (cudaStream_t)stream);
}
int srb_blend_finalize(const float* acc, const float* wsum, void* out, int out_dtype, size_t n, void* stream) {
    if (!acc || !wsum || !out) { set_error("blend_finalize: null pointer"); return SRB_E_ARG; }
    return blend_finalize(acc, wsum, out, out_dtype, n, (cudaStream_t)stream);
}

/* development aid (not in include/srb.h): device buffer [148][8] receiving per-phase cycle totals of the fused
 * attention kernel; pass NULL to switch off */
void srb_debug_set_phase_buffer(long long* dev_buf) { g_fa_dbg = dev_buf; }

}  // extern "C"
