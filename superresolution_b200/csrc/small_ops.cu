// HBM-bound / CUDA-core kernels around the tensor-core convs: input scaling + pixel-unshuffle + conv_first(4->64),
// HAT conv_first(1->96)+LayerNorm, 64->1 convs, binomial blur, nearest x2, LayerNorm(96), CAB channel attention.
// Each cites the reference call site it replaces.
#include "common.h"
#include "ops.h"
#include "ptx.cuh"

namespace srb {

__device__ __forceinline__ float load_pix(const void* x, int dtype, size_t i) {
    // K1: src/dataset.py:31-34 (u16 / 65535.0) fused into the first conv's loads
    if (dtype == SRB_U16) return (float)((const uint16_t*)x)[i] / 65535.0f;
    if (dtype == SRB_BF16) return __bfloat162float(((const bf16*)x)[i]);
    return ((const float*)x)[i];
}

// ---------------------------------------------------------------------------------------------------------------
// K2: pixel_unshuffle(2) + conv_first (4->64, 3x3, pad 1) of RRDBNet (App. A.1).  One thread = one output pixel x 8
// output channels.  Writes bf16 into channels [0,64) of the dense-block buffer and the fp32 trunk (two copies: the
// running trunk and the long-skip "feat").
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rrdb_conv_first_kernel(const void* __restrict__ x, int x_dtype, const float* __restrict__ w /*[64][4][3][3]*/,
                       const float* __restrict__ bias, bf16* __restrict__ out, int out_pitch,
                       float* __restrict__ trunk, float* __restrict__ feat, int B, int h, int w_) {
    // weights transposed to [k = cin*9 + tap][64 cout]: a lane reads its 8 output channels of one k as two 16-byte
    // loads (the [cout][k] order put all 8 channel groups of a warp on one bank: 8-way conflicts on every load)
    __shared__ __align__(16) float sw[36 * 64];
    __shared__ float sb[64];
    for (int i = threadIdx.x; i < 64 * 36; i += blockDim.x) sw[(i % 36) * 64 + i / 36] = w[i];
    if (threadIdx.x < 64) sb[threadIdx.x] = bias[threadIdx.x];
    __syncthreads();
    const unsigned total = (unsigned)B * h * w_ * 8;  // host checks < 2^31
    const int W = 2 * w_, H = 2 * h;
    for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int cg = idx & 7;
        const unsigned pix = idx >> 3;
        const unsigned rowi = pix / (unsigned)w_;
        const int xo = (int)(pix - rowi * w_), yo = (int)(rowi % (unsigned)h), b = (int)(rowi / (unsigned)h);
        float in[36];
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const int yy = yo + ky - 1, xx = xo + kx - 1;
                    float v = 0.f;
                    if (yy >= 0 && yy < h && xx >= 0 && xx < w_)
                        v = load_pix(x, x_dtype, ((size_t)b * H + 2 * yy + (c >> 1)) * W + 2 * xx + (c & 1));
                    in[c * 9 + ky * 3 + kx] = v;
                }
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = sb[cg * 8 + j];
#pragma unroll
        for (int k = 0; k < 36; ++k) {
            const float4 w0 = *(const float4*)(sw + k * 64 + cg * 8), w1 = *(const float4*)(sw + k * 64 + cg * 8 + 4);
            acc[0] = fmaf(in[k], w0.x, acc[0]); acc[1] = fmaf(in[k], w0.y, acc[1]);
            acc[2] = fmaf(in[k], w0.z, acc[2]); acc[3] = fmaf(in[k], w0.w, acc[3]);
            acc[4] = fmaf(in[k], w1.x, acc[4]); acc[5] = fmaf(in[k], w1.y, acc[5]);
            acc[6] = fmaf(in[k], w1.z, acc[6]); acc[7] = fmaf(in[k], w1.w, acc[7]);
        }
        uint4 o;
        __nv_bfloat162* hh = (__nv_bfloat162*)&o;
#pragma unroll
        for (int k = 0; k < 4; ++k) hh[k] = __floats2bfloat162_rn(acc[2 * k], acc[2 * k + 1]);
        *(uint4*)(out + (size_t)pix * out_pitch + cg * 8) = o;
        float4* tp = (float4*)(trunk + (size_t)pix * 64 + cg * 8);
        tp[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        tp[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        float4* fp = (float4*)(feat + (size_t)pix * 64 + cg * 8);
        fp[0] = tp[0];
        fp[1] = tp[1];
    }
}

int rrdb_conv_first(const void* x, int x_dtype, const float* w, const float* bias, bf16* out, int out_pitch,
                    float* trunk, float* feat, int B, int h, int w_, cudaStream_t s) {
    const size_t total = (size_t)B * h * w_ * 8;
    if (total >= ((size_t)1 << 31)) { set_error("rrdb_conv_first: tensor too large for 32-bit indexing"); return SRB_E_SHAPE; }
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    rrdb_conv_first_kernel<<<grid, 256, 0, s>>>(x, x_dtype, w, bias, out, out_pitch, trunk, feat, B, h, w_);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// K7/K17: 3x3 conv with ONE output channel (conv_last of both stages), NHWC bf16 [B,H,W,64] -> fp32 [B,H,W].
// The haloed patch of an 8 x 32 output tile is staged in shared memory once (coalesced 16-byte chunks, zero fill = conv
// padding, 144-byte pixel pitch = conflict-free ldmatrix); the 576-term dot products then run as an implicit GEMM on
// mma.sync.m16n8k16: M = 16 consecutive output pixels of a row, K = 9 taps x 64 channels (36 k-steps), N = 8 of which
// column 0 carries the bf16-rounded weights and column 1 their fp32 residual (hi + lo = the fp32 weight to ~16 bits), so
// out = D[:,0] + D[:,1] + bias.  The CUDA-core form spent ~1300 instructions per thread and tile (576 FMA + as many
// bf16 unpack operations per output) and was instruction-bound at 3.6x the HBM time; this one issues 72 MMAs + 72
// ldmatrix per warp and tile.
// ---------------------------------------------------------------------------------------------------------------
constexpr int C1_TH = 8, C1_TW = 32;                          // output tile: warp w = output row w, two m-tiles of 16 px
constexpr int C1_PH = C1_TH + 2, C1_PW = C1_TW + 2;           // input patch with halo
constexpr int C1_PIX = 144;                                   // bytes per staged pixel (64 bf16 + 16 B skew)
constexpr int C1_SMEM = C1_PH * C1_PW * C1_PIX;

__global__ void __launch_bounds__(256, 2)
conv_cout1_kernel(const bf16* __restrict__ in, int pitch, const float* __restrict__ w /*[1][64][3][3]*/,
                  const float* __restrict__ bias, float* __restrict__ out, int B, int H, int W) {
    extern __shared__ __align__(16) uint8_t c1_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // B fragments of the 36 k-steps (tap t, channel chunk c): b0 = B[k = 2 (lane % 4) + {0, 1}][n = lane / 4],
    // b1 = the same with k + 8; n = 0: bf16(w), n = 1: bf16(w - bf16(w)), n >= 2: 0
    uint32_t bf[36][2];
    {
        const int n = lane >> 2, kq = 2 * (lane & 3);
#pragma unroll
        for (int t = 0; t < 9; ++t)
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    float v[2];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float wv = __ldg(w + (16 * c + kq + 8 * h + e) * 9 + t);
                        const float hi = __bfloat162float(__float2bfloat16_rn(wv));
                        v[e] = n == 0 ? hi : (n == 1 ? wv - hi : 0.f);
                    }
                    __nv_bfloat162 pk = __floats2bfloat162_rn(v[0], v[1]);
                    bf[t * 4 + c][h] = *(uint32_t*)&pk;
                }
    }
    const float b0 = bias[0];
    const int tiles_x = (W + C1_TW - 1) / C1_TW, tiles_y = (H + C1_TH - 1) / C1_TH;  // ragged last tiles are masked
    const int ntiles = B * tiles_x * tiles_y;
    // patch of `tile` -> buffer `buf` by cp.async (zero fill = conv padding by plain stores): the next tile's patch is in
    // flight while the current one is multiplied (two 48 KB buffers, two blocks per SM)
    auto stage = [&](int tile, int buf) {
        const int tx = tile % tiles_x, ty = (tile / tiles_x) % tiles_y, b = tile / (tiles_x * tiles_y);
        const int y0 = ty * C1_TH - 1, x0 = tx * C1_TW - 1;
        uint8_t* dst0 = c1_smem + buf * C1_SMEM;
        constexpr int kIt = (C1_PH * C1_PW * 8 + 255) / 256;
#pragma unroll
        for (int k = 0; k < kIt; ++k) {
            const int i = threadIdx.x + 256 * k;
            if (i >= C1_PH * C1_PW * 8) break;
            const int ck = i & 7, p = i >> 3;
            const int py = p / C1_PW, px = p % C1_PW;
            const int y = y0 + py, x = x0 + px;
            uint8_t* dst = dst0 + p * C1_PIX + ck * 16;
            if (y >= 0 && y < H && x >= 0 && x < W)
                cp_async_16(dst, (const uint4*)(in + (((size_t)b * H + y) * W + x) * pitch) + ck);
            else
                *(uint4*)dst = make_uint4(0, 0, 0, 0);
        }
        cp_async_commit_group();
    };
    if ((int)blockIdx.x < ntiles) stage(blockIdx.x, 0);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int tx = tile % tiles_x, ty = (tile / tiles_x) % tiles_y, b = tile / (tiles_x * tiles_y);
        const int nxt = tile + gridDim.x;
        if (nxt < ntiles) {
            stage(nxt, (it + 1) & 1);  // (that buffer was read in iteration it - 1: the barrier at its end protects it)
            cp_async_wait_group<1>();
        } else {
            cp_async_wait_group<0>();
        }
        __syncthreads();
        const uint8_t* patch = c1_smem + (it & 1) * C1_SMEM;
        // warp = output row `warp`; m-tile mt = output pixels [16 mt, 16 mt + 16); A fragment of (tap, chunk): ldmatrix.x4
        // over rows = pixels (pitch C1_PIX), the four 8x8 matrices = (rows 0-7 | 8-15) x (k 0-7 | 8-15)
        float acc[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) acc[mt][0] = acc[mt][1] = acc[mt][2] = acc[mt][3] = 0.f;
        const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(patch);
        const uint32_t lane_off = (uint32_t)(((lane & 7) + ((lane >> 3) & 1) * 8) * C1_PIX + (lane >> 4) * 16);
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            const uint32_t row_base = smem_base + (uint32_t)(((warp + t / 3) * C1_PW + t % 3) * C1_PIX) + lane_off;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    uint32_t a0, a1, a2, a3;
                    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                                 : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3)
                                 : "r"(row_base + (uint32_t)(16 * mt * C1_PIX + 32 * c)));
                    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
                                 "{%0,%1,%2,%3};"
                                 : "+f"(acc[mt][0]), "+f"(acc[mt][1]), "+f"(acc[mt][2]), "+f"(acc[mt][3])
                                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(bf[t * 4 + c][0]), "r"(bf[t * 4 + c][1]));
                }
            }
        }
        // D fragment: lanes with lane % 4 == 0 hold columns 0, 1 of rows lane / 4 (acc[.][0..1]) and lane / 4 + 8 ([2..3])
        if ((lane & 3) == 0) {
            const int oy = ty * C1_TH + warp;
            if (oy < H) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    const int ox = tx * C1_TW + 16 * mt + (lane >> 2);
                    float* op = out + ((size_t)b * H + oy) * W;
                    if (ox < W) op[ox] = acc[mt][0] + acc[mt][1] + b0;
                    if (ox + 8 < W) op[ox + 8] = acc[mt][2] + acc[mt][3] + b0;
                }
            }
        }
        __syncthreads();  // every warp has read this buffer before the next iteration refills it
    }
}

int conv_cout1(const bf16* in, int pitch, const float* w, const float* bias, float* out, int B, int H, int W,
               cudaStream_t s) {
    static std::atomic<unsigned long long> attr{0};
    SRB_TRY(opt_in_dynamic_smem(conv_cout1_kernel, 2 * C1_SMEM, attr));
    const int ntiles = B * ((H + C1_TH - 1) / C1_TH) * ((W + C1_TW - 1) / C1_TW);
    int grid = ntiles < num_sms() * 2 ? ntiles : num_sms() * 2;  // two resident blocks per SM (registers)
    conv_cout1_kernel<<<grid, 256, 2 * C1_SMEM, s>>>(in, pitch, w, bias, out, B, H, W);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// K8: AntiCheckerboardLayer (src/architecture.py:9-27): depthwise binomial k x k blur, zero padding, on fp32 [B,H,W].
// Optional fused output conversion to u16 (scripts/Modello_5.py:29-33: clip, *65535, truncate).
// ---------------------------------------------------------------------------------------------------------------
template <int K>
__global__ void __launch_bounds__(256)
blur_kernel(const float* __restrict__ in, void* __restrict__ out, int out_dtype, int B, int H, int W) {
    constexpr int R = K / 2;
    // binomial rows: 3:[1,2,1]/4  5:[1,4,6,4,1]/16  7:[1,6,15,20,15,6,1]/64 ; the reference divides the integer
    // outer product by 16 / 256 / 1600 — note 7x7 uses 1600, not 4096 (src/architecture.py:13).
    const float row[7] = {K == 3 ? 1.f : K == 5 ? 1.f : 1.f, K == 3 ? 2.f : K == 5 ? 4.f : 6.f,
                          K == 3 ? 1.f : K == 5 ? 6.f : 15.f, K == 5 ? 4.f : 20.f, K == 5 ? 1.f : 15.f, 6.f, 1.f};
    const float norm = K == 3 ? 16.f : K == 5 ? 256.f : 1600.f;
    const unsigned n = (unsigned)B * H * W;  // host checks < 2^31
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned rowi = i / (unsigned)W;
        const int x = (int)(i - rowi * W), y = (int)(rowi % (unsigned)H);
        const size_t base = (size_t)i - (size_t)y * W - x;
        float s = 0.f;
#pragma unroll
        for (int dy = -R; dy <= R; ++dy) {
            const int yy = y + dy;
            if (yy < 0 || yy >= H) continue;
#pragma unroll
            for (int dx = -R; dx <= R; ++dx) {
                const int xx = x + dx;
                if (xx < 0 || xx >= W) continue;
                s = fmaf(in[base + (size_t)yy * W + xx], (row[dy + R] * row[dx + R]) / norm, s);
            }
        }
        if (out_dtype == SRB_U16) {
            const float c = fminf(fmaxf(s, 0.f), 1.f) * 65535.f;
            ((uint16_t*)out)[i] = (uint16_t)c;
        } else {
            ((float*)out)[i] = s;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// K8b: separable K-tap resample on fp32 [B,H,W] (bicubic resize branch of src/architecture.py:80-81 and its transpose).
// One thread per output pixel; the x taps of a warp's 32 neighbouring outputs overlap, so the loads coalesce through L1.
// Accumulation order matches ATen's upsample_bicubic2d: rows outer, columns inner (sum_i wy_i * (sum_j wx_j * in)).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
resample_kernel(const float* __restrict__ in, float* __restrict__ out, int B, int Hi, int Wi, int Ho, int Wo,
                const int* __restrict__ iy, const float* __restrict__ wy, const int* __restrict__ ix,
                const float* __restrict__ wx, int K) {
    const unsigned n = (unsigned)B * Ho * Wo;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned r = i / (unsigned)Wo;
        const int ox = (int)(i - r * Wo), oy = (int)(r % (unsigned)Ho), b = (int)(r / (unsigned)Ho);
        const float* src = in + (size_t)b * Hi * Wi;
        float acc = 0.f;
        for (int a = 0; a < K; ++a) {
            const float wa = wy[oy * K + a];
            if (wa == 0.f) continue;
            const float* row = src + (size_t)iy[oy * K + a] * Wi;
            float s = 0.f;
            for (int c = 0; c < K; ++c) s = fmaf(row[ix[ox * K + c]], wx[ox * K + c], s);
            acc = fmaf(s, wa, acc);
        }
        out[i] = acc;
    }
}

int resample_separable(const float* in, float* out, int B, int Hi, int Wi, int Ho, int Wo, const int* iy, const float* wy,
                       const int* ix, const float* wx, int K, cudaStream_t s) {
    const size_t n = (size_t)B * Ho * Wo;
    if (n >= ((size_t)1 << 31) || (size_t)B * Hi * Wi >= ((size_t)1 << 31)) {
        set_error("resample_separable: tensor too large for 32-bit indexing");
        return SRB_E_SHAPE;
    }
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    resample_kernel<<<grid, 256, 0, s>>>(in, out, B, Hi, Wi, Ho, Wo, iy, wy, ix, wx, K);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int blur(const float* in, void* out, int out_dtype, int mode, int B, int H, int W, cudaStream_t s) {
    const size_t n = (size_t)B * H * W;
    if (n >= ((size_t)1 << 31)) { set_error("blur: tensor too large for 32-bit indexing"); return SRB_E_SHAPE; }
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    if (mode == SRB_SMOOTH_LIGHT) blur_kernel<3><<<grid, 256, 0, s>>>(in, out, out_dtype, B, H, W);
    else if (mode == SRB_SMOOTH_BALANCED) blur_kernel<5><<<grid, 256, 0, s>>>(in, out, out_dtype, B, H, W);
    else if (mode == SRB_SMOOTH_STRONG) blur_kernel<7><<<grid, 256, 0, s>>>(in, out, out_dtype, B, H, W);
    else { set_error("blur: bad mode %d", mode); return SRB_E_ARG; }
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// fp32 [n] -> fp32 / u16 (used when smoothing == none and the caller wants u16 or a copy into y)
__global__ void convert_out_kernel(const float* __restrict__ in, void* __restrict__ out, int out_dtype, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = in[i];
        if (out_dtype == SRB_U16) ((uint16_t*)out)[i] = (uint16_t)(fminf(fmaxf(v, 0.f), 1.f) * 65535.f);
        else ((float*)out)[i] = v;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// K8c: validation / preview strip of scripts/Modello_5.py:83-85 and Modello_supporto.py:187-190:
//   comp = cat(F.interpolate(lr, size=(H, W), mode='nearest'), sr, hr, dim=3).clamp(0, 1)   -> [B, H, 3W]
// out_dtype F32 = comp itself; U8 = torchvision save_image's quantisation (x * 255 + 0.5, clamp, truncate).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
preview_strip_kernel(const float* __restrict__ lr, const float* __restrict__ sr, const float* __restrict__ hr,
                     void* __restrict__ out, int out_u8, int B, int h, int w, int H, int W, float sy, float sx) {
    const unsigned n = (unsigned)B * H * 3 * W;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned row = i / (3u * W);
        const int c = (int)(i - row * 3u * W), y = (int)(row % (unsigned)H), b = (int)(row / (unsigned)H);
        float v;
        if (c < W) {  // ATen nearest: src = min(floor(dst * in / out), in - 1), the ratio in fp32
            const int yy = min((int)floorf(y * sy), h - 1), xx = min((int)floorf(c * sx), w - 1);
            v = lr[((size_t)b * h + yy) * w + xx];
        } else if (c < 2 * W) {
            v = sr[((size_t)b * H + y) * W + c - W];
        } else {
            v = hr[((size_t)b * H + y) * W + c - 2 * W];
        }
        v = fminf(fmaxf(v, 0.f), 1.f);
        if (out_u8) ((uint8_t*)out)[i] = (uint8_t)fminf(fmaxf(v * 255.f + 0.5f, 0.f), 255.f);
        else ((float*)out)[i] = v;
    }
}

int preview_strip(const float* lr, const float* sr, const float* hr, void* out, int out_u8, int B, int h, int w, int H, int W,
                  cudaStream_t s) {
    const size_t n = (size_t)B * H * 3 * W;
    if (n >= ((size_t)1 << 31)) { set_error("preview_strip: tensor too large for 32-bit indexing"); return SRB_E_SHAPE; }
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    preview_strip_kernel<<<grid, 256, 0, s>>>(lr, sr, hr, out, out_u8, B, h, w, H, W, (float)h / (float)H, (float)w / (float)W);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int convert_out(const float* in, void* out, int out_dtype, size_t n, cudaStream_t s) {
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    convert_out_kernel<<<grid, 256, 0, s>>>(in, out, out_dtype, n);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
