// More HBM-bound kernels: HAT conv_first(1->96)+patch_embed LayerNorm, LayerNorm(96), nearest x2 upsample,
// CAB channel attention (global average pool + squeeze/excite MLP).  SURVEY.md §2.2 rows K6, K9, K10, K13.
#include "common.h"
#include "ops.h"

namespace srb {

// 16-lane group reduction (groups are aligned halves of a warp)
__device__ __forceinline__ float group16_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// K9+K10: HAT.conv_first (1->96, 3x3, pad 1) on the fp32 stage-1 image, fused with patch_embed.norm.
// 16 lanes per pixel, lanes 0..11 own 8 channels each.  Writes f0 (long skip, pre-norm) and x = LN(f0), both bf16
// [B,H,W,96].  LayerNorm statistics in fp32 over the fp32 conv result (eps 1e-5, biased variance).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
hat_conv_first_ln_kernel(const float* __restrict__ img, const float* __restrict__ w /*[96][1][3][3]*/,
                         const float* __restrict__ bias, const float* __restrict__ gamma,
                         const float* __restrict__ beta, bf16* __restrict__ f0, bf16* __restrict__ xn, int B, int H,
                         int W) {
    const int sub = threadIdx.x & 15;
    const bool act = sub < 12;
    const int c0 = (act ? sub : 0) * 8;
    float wr[72], br[8], gr[8], ber[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
#pragma unroll
        for (int t = 0; t < 9; ++t) wr[j * 9 + t] = __ldg(w + (c0 + j) * 9 + t);
        br[j] = __ldg(bias + c0 + j);
        gr[j] = __ldg(gamma + c0 + j);
        ber[j] = __ldg(beta + c0 + j);
    }
    const size_t npix = (size_t)B * H * W;
    const size_t stride = ((size_t)gridDim.x * blockDim.x) >> 4;
    // all 16 lanes of a group iterate together (npix is a multiple of the per-iteration pixel count or guarded)
    for (size_t pix0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 4; pix0 < npix + stride; pix0 += stride) {
        const bool valid = pix0 < npix;
        const size_t pix = valid ? pix0 : 0;
        const int x = pix % W, y = (pix / W) % H;
        const size_t b = pix / ((size_t)W * H);
        float in[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            const int yy = y + t / 3 - 1, xx = x + t % 3 - 1;
            in[t] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(img + (b * H + yy) * W + xx) : 0.f;
        }
        float v[8];
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float a = br[j];
#pragma unroll
            for (int t = 0; t < 9; ++t) a = fmaf(in[t], wr[j * 9 + t], a);
            v[j] = a;
            s += a;
        }
        s = group16_sum(act ? s : 0.f);
        const float mean = s * (1.f / 96.f);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) q += (v[j] - mean) * (v[j] - mean);
        q = group16_sum(act ? q : 0.f);
        const float rstd = rsqrtf(q * (1.f / 96.f) + 1e-5f);
        if (valid && act) {
            uint4 o0, o1;
            __nv_bfloat162* h0 = (__nv_bfloat162*)&o0;
            __nv_bfloat162* h1 = (__nv_bfloat162*)&o1;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                h0[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
                h1[k] = __floats2bfloat162_rn((v[2 * k] - mean) * rstd * gr[2 * k] + ber[2 * k],
                                              (v[2 * k + 1] - mean) * rstd * gr[2 * k + 1] + ber[2 * k + 1]);
            }
            *(uint4*)(f0 + pix * 96 + c0) = o0;
            *(uint4*)(xn + pix * 96 + c0) = o1;
        }
    }
}

int hat_conv_first_ln(const float* img, const float* w, const float* bias, const float* gamma, const float* beta,
                      bf16* f0, bf16* xn, int B, int H, int W, cudaStream_t s) {
    const size_t threads = (size_t)B * H * W * 16;
    int grid = (int)((threads + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    hat_conv_first_ln_kernel<<<grid, 256, 0, s>>>(img, w, bias, gamma, beta, f0, xn, B, H, W);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// K10: LayerNorm over 96 channels (eps 1e-5), bf16 [ntok,96] -> bf16 [ntok,96]; 16 lanes per token (12 active).
// Under the reference's autocast LayerNorm runs in fp32 and its output is consumed by bf16 GEMMs: same here.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
layernorm96_kernel(const bf16* __restrict__ in, const float* __restrict__ gamma, const float* __restrict__ beta,
                   bf16* __restrict__ out, size_t ntok) {
    const int sub = threadIdx.x & 15;
    const bool act = sub < 12;
    const int c0 = (act ? sub : 0) * 8;
    float gr[8], br[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        gr[j] = __ldg(gamma + c0 + j);
        br[j] = __ldg(beta + c0 + j);
    }
    const size_t stride = ((size_t)gridDim.x * blockDim.x) >> 4;
    for (size_t t0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 4; t0 < ntok + stride; t0 += stride) {
        const bool valid = t0 < ntok;
        const size_t tok = valid ? t0 : 0;
        float v[8];
        {
            const uint4 raw = *(const uint4*)(in + tok * 96 + c0);
            const __nv_bfloat162* h = (const __nv_bfloat162*)&raw;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = __bfloat1622float2(h[k]);
                v[2 * k] = f.x;
                v[2 * k + 1] = f.y;
            }
        }
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[j];
        s = group16_sum(act ? s : 0.f);
        const float mean = s * (1.f / 96.f);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) q += (v[j] - mean) * (v[j] - mean);
        q = group16_sum(act ? q : 0.f);
        const float rstd = rsqrtf(q * (1.f / 96.f) + 1e-5f);
        if (valid && act) {
            uint4 o;
            __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                h[k] = __floats2bfloat162_rn((v[2 * k] - mean) * rstd * gr[2 * k] + br[2 * k],
                                             (v[2 * k + 1] - mean) * rstd * gr[2 * k + 1] + br[2 * k + 1]);
            *(uint4*)(out + tok * 96 + c0) = o;
        }
    }
}

// HAB combine + norm2 in one pass:  y = t + cab * cs[b,:]  (cs = sigmoid(...) * conv_scale from the CAB channel
// attention),  xn = LayerNorm(y).  Replaces the add/mul elementwise kernels and the norm2 LayerNorm of the reference.
__global__ void __launch_bounds__(256)
combine_ln96_kernel(const bf16* __restrict__ t, const bf16* __restrict__ cab, const float* __restrict__ cs,
                    const float* __restrict__ gamma, const float* __restrict__ beta, bf16* __restrict__ y,
                    bf16* __restrict__ xn, size_t ntok, size_t tok_per_img) {
    const int sub = threadIdx.x & 15;
    const bool act = sub < 12;
    const int c0 = (act ? sub : 0) * 8;
    float gr[8], br[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        gr[j] = __ldg(gamma + c0 + j);
        br[j] = __ldg(beta + c0 + j);
    }
    const size_t stride = ((size_t)gridDim.x * blockDim.x) >> 4;
    for (size_t t0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 4; t0 < ntok + stride; t0 += stride) {
        const bool valid = t0 < ntok;
        const size_t tok = valid ? t0 : 0;
        const size_t b = tok / tok_per_img;
        float v[8];
        {
            const uint4 ra = *(const uint4*)(t + tok * 96 + c0);
            const uint4 rb = *(const uint4*)(cab + tok * 96 + c0);
            const float4 s0 = __ldg((const float4*)(cs + b * 96 + c0));
            const float4 s1 = __ldg((const float4*)(cs + b * 96 + c0 + 4));
            const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
            const __nv_bfloat162* ha = (const __nv_bfloat162*)&ra;
            const __nv_bfloat162* hb = (const __nv_bfloat162*)&rb;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 fa = __bfloat1622float2(ha[k]), fb = __bfloat1622float2(hb[k]);
                v[2 * k] = fmaf(fb.x, sc[2 * k], fa.x);
                v[2 * k + 1] = fmaf(fb.y, sc[2 * k + 1], fa.y);
            }
        }
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[j];
        s = group16_sum(act ? s : 0.f);
        const float mean = s * (1.f / 96.f);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) q += (v[j] - mean) * (v[j] - mean);
        q = group16_sum(act ? q : 0.f);
        const float rstd = rsqrtf(q * (1.f / 96.f) + 1e-5f);
        if (valid && act) {
            uint4 o0, o1;
            __nv_bfloat162* h0 = (__nv_bfloat162*)&o0;
            __nv_bfloat162* h1 = (__nv_bfloat162*)&o1;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                h0[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
                h1[k] = __floats2bfloat162_rn((v[2 * k] - mean) * rstd * gr[2 * k] + br[2 * k],
                                              (v[2 * k + 1] - mean) * rstd * gr[2 * k + 1] + br[2 * k + 1]);
            }
            *(uint4*)(y + tok * 96 + c0) = o0;
            *(uint4*)(xn + tok * 96 + c0) = o1;
        }
    }
}

int combine_ln96(const bf16* t, const bf16* cab, const float* cs, const float* gamma, const float* beta, bf16* y,
                 bf16* xn, size_t ntok, size_t tok_per_img, cudaStream_t s) {
    const size_t threads = ntok * 16;
    int grid = (int)((threads + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    combine_ln96_kernel<<<grid, 256, 0, s>>>(t, cab, cs, gamma, beta, y, xn, ntok, tok_per_img);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int layernorm96(const bf16* in, const float* gamma, const float* beta, bf16* out, size_t ntok, cudaStream_t s) {
    const size_t threads = ntok * 16;
    int grid = (int)((threads + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    layernorm96_kernel<<<grid, 256, 0, s>>>(in, gamma, beta, out, ntok);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// K6: F.interpolate(scale_factor=2, mode='nearest') on NHWC bf16 with C=64: out[b,y,x,:] = in[b,y/2,x/2,:]
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
upsample2x_kernel(const bf16* __restrict__ in, int in_pitch, bf16* __restrict__ out, int B, int h, int w) {
    const size_t n = (size_t)B * (2 * h) * (2 * w) * 8;  // 16-byte chunks, 8 per pixel
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int ck = i & 7;
        const size_t pix = i >> 3;
        const int x = pix % (2 * w), y = (pix / (2 * w)) % (2 * h);
        const size_t b = pix / ((size_t)4 * w * h);
        *(uint4*)(out + pix * 64 + ck * 8) =
            __ldg((const uint4*)(in + ((b * h + (y >> 1)) * w + (x >> 1)) * in_pitch + ck * 8));
    }
}

int upsample2x(const bf16* in, int in_pitch, bf16* out, int B, int h, int w, cudaStream_t s) {
    const size_t n = (size_t)B * 4 * h * w * 8;
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    upsample2x_kernel<<<grid, 256, 0, s>>>(in, in_pitch, out, B, h, w);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// K13: ChannelAttention of CAB: AdaptiveAvgPool2d(1) over [B,HW,96], then 1x1 96->3, ReLU, 1x1 3->96, sigmoid.
// Stage A lives in the cab.2 conv epilogue (conv_tc.cu, ConvArgs::colsum): fp32 channel sums per 16 x 16 output tile,
// partial[b][tile][96]. Stage B (here): fixed-order sum of the tiles, the tiny MLP, and cs[b][c] = sigmoid(...) *
// conv_scale so the HAB combine is a single FMA.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(96)
cab_finish_kernel(const float* __restrict__ partial, int nslab, float inv_hw, const float* __restrict__ w1 /*[3][96]*/,
                  const float* __restrict__ b1, const float* __restrict__ w2 /*[96][3]*/,
                  const float* __restrict__ b2, float conv_scale, float* __restrict__ cs) {
    __shared__ float mean[96];
    __shared__ float hid[3];
    const int b = blockIdx.x, c = threadIdx.x;
    float t4[4] = {0.f, 0.f, 0.f, 0.f};  // four interleaved chains, fixed order
    for (int i = 0; i < nslab; ++i) t4[i & 3] += partial[((size_t)b * nslab + i) * 96 + c];
    mean[c] = ((t4[0] + t4[1]) + (t4[2] + t4[3])) * inv_hw;
    __syncthreads();
    if (c < 3) {
        float s = b1[c];
        for (int k = 0; k < 96; ++k) s = fmaf(w1[c * 96 + k], mean[k], s);
        hid[c] = fmaxf(s, 0.f);
    }
    __syncthreads();
    float s = b2[c];
#pragma unroll
    for (int k = 0; k < 3; ++k) s = fmaf(w2[c * 3 + k], hid[k], s);
    cs[b * 96 + c] = conv_scale / (1.f + expf(-s));
}

// One slab = one 16 x 16 output tile of the cab.2 conv (conv_tc <96,2,9>), which writes the per-tile sums itself.
int cab_pool_slabs(int HW) { return (HW + 255) / 256; }

int cab_channel_scale(const float* pool, int nslab, float* cs, const float* w1, const float* b1, const float* w2,
                      const float* b2, float conv_scale, int B, int HW, cudaStream_t s) {
    cab_finish_kernel<<<B, 96, 0, s>>>(pool, nslab, 1.f / (float)HW, w1, b1, w2, b2, conv_scale, cs);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
