// More HBM-bound kernels: HAT conv_first(1->96)+patch_embed LayerNorm, LayerNorm(96), nearest x2 upsample,
// CAB channel attention (global average pool + squeeze/excite MLP).  SURVEY.md §2.2 rows K6, K9, K10, K13.
#include "common.h"
#include "ops.h"

namespace srb {

// 16-lane group reduction (groups are aligned halves of a warp)
// sum over the 16 lanes of a half-warp; the mask names only that half, so the two groups of a warp may leave their
// grid-stride loops on different iterations
__device__ __forceinline__ float group16_sum(float v) {
    const unsigned m = 0xffffu << (threadIdx.x & 16);
    v += __shfl_xor_sync(m, v, 8);
    v += __shfl_xor_sync(m, v, 4);
    v += __shfl_xor_sync(m, v, 2);
    v += __shfl_xor_sync(m, v, 1);
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// K9+K10: HAT.conv_first (1->96, 3x3, pad 1) on the fp32 stage-1 image, fused with patch_embed.norm.
// 16 lanes per pixel, lanes 0..11 own 8 channels each.  Writes f0 (long skip, pre-norm) and x = LN(f0), both bf16
// [B,H,W,96].  LayerNorm statistics in fp32 over the fp32 conv result (eps 1e-5, biased variance).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 2)
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
    // Each 16-lane group walks a run of 8 consecutive pixels of one row with a sliding 3x3 window: 3 new loads per
    // pixel instead of 9, index arithmetic once per run (32-bit; host checks B*H*W < 2^31 and W % 8 == 0).
    const unsigned segs_per_row = (unsigned)W >> 3;
    const unsigned nseg = (unsigned)B * H * segs_per_row;
    const unsigned stride = (gridDim.x * blockDim.x) >> 4;
    for (unsigned seg = (blockIdx.x * blockDim.x + threadIdx.x) >> 4; seg < nseg; seg += stride) {
        const unsigned rowi = seg / segs_per_row;
        const int xs = (int)(seg - rowi * segs_per_row) * 8, y = (int)(rowi % (unsigned)H);
        const unsigned b = rowi / (unsigned)H;
        const float* r1 = img + ((size_t)b * H + y) * W;
        const float* r0 = r1 - W;
        const float* r2 = r1 + W;
        const bool up = y > 0, dn = y + 1 < H;
        float in[9];  // in[3 * row + col]
        in[0] = in[3] = in[6] = 0.f;
        if (xs > 0) {
            in[0] = up ? __ldg(r0 + xs - 1) : 0.f;
            in[3] = __ldg(r1 + xs - 1);
            in[6] = dn ? __ldg(r2 + xs - 1) : 0.f;
        }
        in[1] = up ? __ldg(r0 + xs) : 0.f;
        in[4] = __ldg(r1 + xs);
        in[7] = dn ? __ldg(r2 + xs) : 0.f;
#pragma unroll 1
        for (int i = 0; i < 8; ++i) {
            const int x = xs + i;
            const bool rt = x + 1 < W;
            in[2] = (rt && up) ? __ldg(r0 + x + 1) : 0.f;
            in[5] = rt ? __ldg(r1 + x + 1) : 0.f;
            in[8] = (rt && dn) ? __ldg(r2 + x + 1) : 0.f;
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
            if (act) {
                uint4 o0, o1;
                __nv_bfloat162* h0 = (__nv_bfloat162*)&o0;
                __nv_bfloat162* h1 = (__nv_bfloat162*)&o1;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    h0[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
                    h1[k] = __floats2bfloat162_rn((v[2 * k] - mean) * rstd * gr[2 * k] + ber[2 * k],
                                                  (v[2 * k + 1] - mean) * rstd * gr[2 * k + 1] + ber[2 * k + 1]);
                }
                const size_t pix = ((size_t)b * H + y) * W + x;
                *(uint4*)(f0 + pix * 96 + c0) = o0;
                *(uint4*)(xn + pix * 96 + c0) = o1;
            }
            in[0] = in[1]; in[1] = in[2];
            in[3] = in[4]; in[4] = in[5];
            in[6] = in[7]; in[7] = in[8];
        }
    }
}

int hat_conv_first_ln(const float* img, const float* w, const float* bias, const float* gamma, const float* beta,
                      bf16* f0, bf16* xn, int B, int H, int W, cudaStream_t s) {
    const size_t threads = (size_t)B * H * W * 16;
    if (threads >= ((size_t)1 << 31)) { set_error("hat_conv_first_ln: B*H*W too large for 32-bit indexing"); return SRB_E_SHAPE; }
    if (W % 8) { set_error("hat_conv_first_ln: W=%d must be a multiple of 8 (the HAT window)", W); return SRB_E_SHAPE; }
    int grid = (int)((threads / 8 + 255) / 256);
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
    constexpr int U = 4;  // tokens in flight per 16-lane group (HBM-bound: keep several 16-byte loads outstanding)
    const size_t stride = ((size_t)gridDim.x * blockDim.x) >> 4;
    for (size_t t0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 4; t0 < ntok + stride * U; t0 += stride * U) {
        uint4 raw[U];
        bool valid[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t tk = t0 + u * stride;
            valid[u] = tk < ntok;
            raw[u] = valid[u] ? __ldcs((const uint4*)(in + tk * 96 + c0)) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float v[8];
            const __nv_bfloat162* h = (const __nv_bfloat162*)&raw[u];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = __bfloat1622float2(h[k]);
                v[2 * k] = f.x;
                v[2 * k + 1] = f.y;
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
            if (valid[u] && act) {
                uint4 o;
                __nv_bfloat162* ho = (__nv_bfloat162*)&o;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ho[k] = __floats2bfloat162_rn((v[2 * k] - mean) * rstd * gr[2 * k] + br[2 * k],
                                                  (v[2 * k + 1] - mean) * rstd * gr[2 * k + 1] + br[2 * k + 1]);
                *(uint4*)(out + (t0 + u * stride) * 96 + c0) = o;
            }
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
    const unsigned n = (unsigned)B * (2 * h) * (2 * w) * 8;  // 16-byte chunks, 8 per pixel (host checks < 2^31)
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int ck = i & 7;
        const unsigned pix = i >> 3;
        const unsigned rowi = pix / (unsigned)(2 * w);
        const unsigned x = pix - rowi * (2 * w), y = rowi % (unsigned)(2 * h), b = rowi / (unsigned)(2 * h);
        *(uint4*)(out + (size_t)pix * 64 + ck * 8) =
            __ldg((const uint4*)(in + (((size_t)b * h + (y >> 1)) * w + (x >> 1)) * in_pitch + ck * 8));
    }
}

int upsample2x(const bf16* in, int in_pitch, bf16* out, int B, int h, int w, cudaStream_t s) {
    const size_t n = (size_t)B * 4 * h * w * 8;
    if (n >= ((size_t)1 << 31)) { set_error("upsample2x: tensor too large for 32-bit indexing"); return SRB_E_SHAPE; }
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
__global__ void __launch_bounds__(768)
cab_finish_kernel(const float* __restrict__ partial, int nslab, float inv_hw, const float* __restrict__ w1 /*[3][96]*/,
                  const float* __restrict__ b1, const float* __restrict__ w2 /*[96][3]*/,
                  const float* __restrict__ b2, float conv_scale, float* __restrict__ cs) {
    // 8 slab groups x 96 channels: group g sums slabs g, g+8, ... (independent loads, unrolled), then the 8 group
    // sums are added in fixed order -> deterministic, and the 256-slab sum is 32 loads deep instead of 256
    __shared__ float part[8][96];
    __shared__ float mean[96];
    __shared__ float hid[3];
    const int b = blockIdx.x, c = threadIdx.x % 96, g = threadIdx.x / 96;
    float tot = 0.f;
    const float* src = partial + (size_t)b * nslab * 96 + c;
#pragma unroll 8
    for (int i = g; i < nslab; i += 8) tot += __ldg(src + (size_t)i * 96);
    part[g][c] = tot;
    __syncthreads();
    if (threadIdx.x < 96) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += part[k][c];
        mean[c] = t * inv_hw;
    }
    __syncthreads();
    if (threadIdx.x < 96) {  // warp w computes hidden unit w: 3 products per lane, fixed-order butterfly
        const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
        float s = w1[w * 96 + l] * mean[l];
        s = fmaf(w1[w * 96 + l + 32], mean[l + 32], s);
        s = fmaf(w1[w * 96 + l + 64], mean[l + 64], s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (l == 0) hid[w] = fmaxf(s + b1[w], 0.f);
    }
    __syncthreads();
    if (threadIdx.x < 96) {
        float s = b2[c];
#pragma unroll
        for (int k = 0; k < 3; ++k) s = fmaf(w2[c * 3 + k], hid[k], s);
        cs[b * 96 + c] = conv_scale / (1.f + expf(-s));
    }
}

// One slab = one 16 x 16 output tile of the cab.2 conv (conv_tc <96,2,9>), which writes the per-tile sums itself.
// = the 16 x 16 spatial tiles of conv_tc<96, 2, 9> (cab.2), whose epilogue leaves one channel-sum row per tile
int cab_pool_slabs(int H, int W) { return ((H + 15) / 16) * ((W + 15) / 16); }

int cab_channel_scale(const float* pool, int nslab, float* cs, const float* w1, const float* b1, const float* w2,
                      const float* b2, float conv_scale, int B, int HW, cudaStream_t s) {
    cab_finish_kernel<<<B, 768, 0, s>>>(pool, nslab, 1.f / (float)HW, w1, b1, w2, b2, conv_scale, cs);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
