// Element-wise backward operators of the RRDBNet training path (autograd of LeakyReLU,
// residual scaling, F.interpolate(nearest, 2x) inside BasicSR RRDBNet).  All tensors are NHWC slices: [npx][C] with a
// channel pitch, so a slice of a dense-block gradient buffer is addressed in place.  HBM-bound, 16-byte accesses.
#include <cuda_bf16.h>

#include "common.h"
#include "ops.h"
#include "ptx.cuh"

namespace srb {

namespace {

// g[px][c] *= (y[px][c] > 0 ? 1 : slope)   — LeakyReLU backward from the saved OUTPUT (sign(y) == sign(pre-activation))
__global__ void __launch_bounds__(256)
lrelu_bwd_kernel(bf16* __restrict__ g, int g_pitch, const bf16* __restrict__ y, int y_pitch, size_t npx, int C,
                 float slope) {
    const int cg = C >> 3;
    const size_t n = npx * cg;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t px = i / cg;
        const int c = (int)(i - px * cg) * 8;
        uint4 gv = *(const uint4*)(g + px * g_pitch + c);
        const uint4 yv = __ldg((const uint4*)(y + px * y_pitch + c));
        __nv_bfloat162* gh = (__nv_bfloat162*)&gv;
        const __nv_bfloat162* yh = (const __nv_bfloat162*)&yv;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float2 gf = __bfloat1622float2(gh[k]);
            const float2 yf = __bfloat1622float2(yh[k]);
            gf.x = yf.x > 0.f ? gf.x : gf.x * slope;
            gf.y = yf.y > 0.f ? gf.y : gf.y * slope;
            gh[k] = __floats2bfloat162_rn(gf.x, gf.y);
        }
        *(uint4*)(g + px * g_pitch + c) = gv;
    }
}

__device__ __forceinline__ float4 load4(const void* p, int f32, size_t idx) {
    if (f32) return *(const float4*)((const float*)p + idx);
    const uint2 raw = *(const uint2*)((const bf16*)p + idx);
    const float2 a = __bfloat1622float2(*(const __nv_bfloat162*)&raw.x), b = __bfloat1622float2(*(const __nv_bfloat162*)&raw.y);
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void store4(void* p, int f32, size_t idx, float4 v) {
    if (f32) {
        *(float4*)((float*)p + idx) = v;
    } else {
        uint2 raw;
        *(__nv_bfloat162*)&raw.x = __floats2bfloat162_rn(v.x, v.y);
        *(__nv_bfloat162*)&raw.y = __floats2bfloat162_rn(v.z, v.w);
        *(uint2*)((bf16*)p + idx) = raw;
    }
}

// out = a * x + b * y  (y optional); every operand bf16 or fp32 with its own channel pitch; out may alias x or y
__global__ void __launch_bounds__(256)
axpby_kernel(void* out, int out_f32, int out_pitch, float a0, const float* __restrict__ a_vec, size_t npx_per_sample,
             const void* x, int x_f32, int x_pitch, float b, const void* y, int y_f32, int y_pitch, size_t npx, int C) {
    const int cg = C >> 2;
    const size_t n = npx * cg;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t px = i / cg;
        const int c = (int)(i - px * cg) * 4;
        const float a = a_vec ? a0 * __ldg(a_vec + px / npx_per_sample) : a0;  // per-sample factor (DropPath)
        float4 v = load4(x, x_f32, px * x_pitch + c);
        v.x *= a; v.y *= a; v.z *= a; v.w *= a;
        if (y) {
            const float4 w = load4(y, y_f32, px * y_pitch + c);
            v.x = fmaf(b, w.x, v.x); v.y = fmaf(b, w.y, v.y); v.z = fmaf(b, w.z, v.z); v.w = fmaf(b, w.w, v.w);
        }
        store4(out, out_f32, px * out_pitch + c, v);
    }
}

// backward of nearest 2x upsampling: out[b,y,x,c] = sum_{i,j in {0,1}} g[b,2y+i,2x+j,c]  (fp32 sum, bf16 storage)
__global__ void __launch_bounds__(256)
upsample2x_bwd_kernel(const bf16* __restrict__ g, int g_pitch, bf16* __restrict__ out, int out_pitch, int B, int h, int w,
                      int C) {
    const int cg = C >> 3;
    const size_t n = (size_t)B * h * w * cg;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t px = i / cg;
        const int c = (int)(i - px * cg) * 8;
        const int x = (int)(px % w), y = (int)((px / w) % h);
        const size_t b = px / ((size_t)w * h);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int ij = 0; ij < 4; ++ij) {
            const size_t gp = ((b * 2 * h + 2 * y + (ij >> 1)) * (size_t)(2 * w) + 2 * x + (ij & 1));
            const uint4 raw = __ldg((const uint4*)(g + gp * g_pitch + c));
            const __nv_bfloat162* hh = (const __nv_bfloat162*)&raw;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = __bfloat1622float2(hh[k]);
                acc[2 * k] += f.x;
                acc[2 * k + 1] += f.y;
            }
        }
        uint4 o;
        __nv_bfloat162* oh = (__nv_bfloat162*)&o;
#pragma unroll
        for (int k = 0; k < 4; ++k) oh[k] = __floats2bfloat162_rn(acc[2 * k], acc[2 * k + 1]);
        *(uint4*)(out + px * out_pitch + c) = o;
    }
}

int grid_for(size_t n) {
    size_t g = (n + 255) / 256;
    const size_t cap = (size_t)num_sms() * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

int lrelu_bwd(bf16* g, int g_pitch, const bf16* y, int y_pitch, size_t npx, int C, float slope, cudaStream_t s) {
    if (C % 8 || g_pitch % 8 || y_pitch % 8 || (((uintptr_t)g | (uintptr_t)y) & 15)) {
        set_error("lrelu_bwd: C and pitches must be multiples of 8, pointers 16-byte aligned");
        return SRB_E_ARG;
    }
    lrelu_bwd_kernel<<<grid_for(npx * (C / 8)), 256, 0, s>>>(g, g_pitch, y, y_pitch, npx, C, slope);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int axpby(void* out, int out_f32, int out_pitch, float a, const void* x, int x_f32, int x_pitch, float b, const void* y,
          int y_f32, int y_pitch, size_t npx, int C, cudaStream_t s, const float* a_vec, size_t npx_per_sample) {
    if (a_vec && (npx_per_sample == 0 || npx % npx_per_sample)) { set_error("axpby: npx must be a multiple of npx_per_sample"); return SRB_E_ARG; }
    if (C % 4 || out_pitch % 4 || x_pitch % 4 || (y && y_pitch % 4) ||
        (((uintptr_t)out | (uintptr_t)x | (uintptr_t)y) & 7)) {
        set_error("axpby: C and pitches must be multiples of 4, pointers 8-byte aligned (16 for fp32)");
        return SRB_E_ARG;
    }
    if ((out_f32 && ((uintptr_t)out & 15)) || (x_f32 && ((uintptr_t)x & 15)) || (y && y_f32 && ((uintptr_t)y & 15))) {
        set_error("axpby: fp32 operands must be 16-byte aligned");
        return SRB_E_ARG;
    }
    axpby_kernel<<<grid_for(npx * (C / 4)), 256, 0, s>>>(out, out_f32, out_pitch, a, a_vec, npx_per_sample, x, x_f32, x_pitch,
                                                         b, y, y_f32, y_pitch, npx, C);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int upsample2x_bwd(const bf16* g, int g_pitch, bf16* out, int out_pitch, int B, int h, int w, int C, cudaStream_t s) {
    if (C % 8 || g_pitch % 8 || out_pitch % 8 || (((uintptr_t)g | (uintptr_t)out) & 15)) {
        set_error("upsample2x_bwd: C and pitches must be multiples of 8, pointers 16-byte aligned");
        return SRB_E_ARG;
    }
    upsample2x_bwd_kernel<<<grid_for((size_t)B * h * w * (C / 8)), 256, 0, s>>>(g, g_pitch, out, out_pitch, B, h, w, C);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb

// ---------------------------------------------------------------------------------------------------------------
// HAT building blocks of the stage-2 training path (training_hat.py): LayerNorm(96) backward, exact-erf GELU forward/backward
// ---------------------------------------------------------------------------------------------------------------
namespace srb {

namespace {

__device__ __forceinline__ float half16_sum(float v) {  // sum over the 16 lanes of a half-warp
    const unsigned m = 0xffffu << (threadIdx.x & 16);
    v += __shfl_xor_sync(m, v, 8);
    v += __shfl_xor_sync(m, v, 4);
    v += __shfl_xor_sync(m, v, 2);
    v += __shfl_xor_sync(m, v, 1);
    return v;
}

constexpr int LNB_BLOCKS = 148 * 4;

// y = LayerNorm(x) * gamma + beta over 96 channels (eps 1e-5, biased variance), g = dL/dy:
//   dx = (g*gamma - mean(g*gamma) - xhat * mean(g*gamma*xhat)) * rstd (+ res),  dgamma = sum_tok g*xhat,  dbeta = sum_tok g
// 16 lanes per token (12 active, 8 channels each).  partial[blk][2][96]: per-block sums of dgamma / dbeta.
__global__ void __launch_bounds__(256)
layernorm96_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ g, const float* __restrict__ gamma,
                       const void* __restrict__ res, int res_f32, bf16* __restrict__ dx, size_t ntok,
                       float* __restrict__ partial) {
    __shared__ float red[16][2][96];
    const int sub = threadIdx.x & 15, grp = threadIdx.x >> 4;
    const bool act = sub < 12;
    const int c0 = (act ? sub : 0) * 8;
    float gam[8], dg[8], db[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { gam[j] = __ldg(gamma + c0 + j); dg[j] = 0.f; db[j] = 0.f; }
    const size_t stride = ((size_t)gridDim.x * blockDim.x) >> 4;
    constexpr int U = 2;  // tokens in flight per 16-lane group: x, g and res of both are requested before either is used
    for (size_t t0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 4; t0 < ntok; t0 += U * stride) {
        uint4 xr[U], gr[U], rr0[U], rr1[U];
        bool valid[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t tok = t0 + u * stride;
            valid[u] = tok < ntok;
            const size_t off = (valid[u] ? tok : t0) * 96 + c0;
            xr[u] = __ldcs((const uint4*)(x + off));
            gr[u] = __ldcs((const uint4*)(g + off));
            rr0[u] = rr1[u] = make_uint4(0, 0, 0, 0);
            if (res && act) {
                if (res_f32) {
                    rr0[u] = __ldcs((const uint4*)((const float*)res + off));
                    rr1[u] = __ldcs((const uint4*)((const float*)res + off + 4));
                } else {
                    rr0[u] = __ldcs((const uint4*)((const bf16*)res + off));
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!valid[u]) continue;  // uniform over the 16-lane group (the shuffles below name the half-warp)
            const size_t tok = t0 + u * stride;
            float xv[8], gv[8];
            {
                const __nv_bfloat162* xh = (const __nv_bfloat162*)&xr[u];
                const __nv_bfloat162* gh = (const __nv_bfloat162*)&gr[u];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 a = __bfloat1622float2(xh[k]), b = __bfloat1622float2(gh[k]);
                    xv[2 * k] = a.x; xv[2 * k + 1] = a.y; gv[2 * k] = b.x; gv[2 * k + 1] = b.y;
                }
            }
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) s += xv[j];
            const float mean = half16_sum(act ? s : 0.f) * (1.f / 96.f);
            float q = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) q += (xv[j] - mean) * (xv[j] - mean);
            const float rstd = rsqrtf(half16_sum(act ? q : 0.f) * (1.f / 96.f) + 1e-5f);
            float m1 = 0.f, m2 = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                xv[j] = (xv[j] - mean) * rstd;  // xhat
                const float gg = gv[j] * gam[j];
                m1 += gg;
                m2 = fmaf(gg, xv[j], m2);
                if (act) {
                    dg[j] = fmaf(gv[j], xv[j], dg[j]);
                    db[j] += gv[j];
                }
            }
            m1 = half16_sum(act ? m1 : 0.f) * (1.f / 96.f);
            m2 = half16_sum(act ? m2 : 0.f) * (1.f / 96.f);
            if (act) {
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = (gv[j] * gam[j] - m1 - xv[j] * m2) * rstd;
                if (res) {
                    if (res_f32) {
                        const float4 r0 = *(const float4*)&rr0[u], r1 = *(const float4*)&rr1[u];
                        o[0] += r0.x; o[1] += r0.y; o[2] += r0.z; o[3] += r0.w; o[4] += r1.x; o[5] += r1.y; o[6] += r1.z; o[7] += r1.w;
                    } else {
                        const __nv_bfloat162* rh = (const __nv_bfloat162*)&rr0[u];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = __bfloat1622float2(rh[k]);
                            o[2 * k] += f.x; o[2 * k + 1] += f.y;
                        }
                    }
                }
                uint4 ov;
                __nv_bfloat162* oh = (__nv_bfloat162*)&ov;
#pragma unroll
                for (int k = 0; k < 4; ++k) oh[k] = __floats2bfloat162_rn(o[2 * k], o[2 * k + 1]);
                __stcs((uint4*)(dx + tok * 96 + c0), ov);
            }
        }
    }
    if (act) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { red[grp][0][c0 + j] = dg[j]; red[grp][1][c0 + j] = db[j]; }
    }
    __syncthreads();
    if (threadIdx.x < 192) {
        const int w = threadIdx.x / 96, c = threadIdx.x % 96;
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k) s += red[k][w][c];
        partial[((size_t)blockIdx.x * 2 + w) * 96 + c] = s;
    }
}

// The same backward as a bulk-copy pipeline (bf16 or no residual): a producer warp streams 64-token tiles of x, g and
// res (12 KB each, contiguous) into a ring of shared-memory stages with cp.async.bulk; eight consumer warps take 8 tokens
// each.  The register-resident kernel above cannot keep more than one token per 16-lane group in flight at 64 registers
// (3.5 TB/s at 1 M tokens); here the bytes in flight are the ring (2 CTAs x 108 KB per SM).
constexpr int LNQ_TOK = 64, LNQ_STAGES = 3, LNQ_TILE_BYTES = LNQ_TOK * 192, LNQ_STAGE_BYTES = 3 * LNQ_TILE_BYTES;
constexpr int LNQ_SMEM = LNQ_STAGES * LNQ_STAGE_BYTES + 2 * LNQ_STAGES * 8 + 128;

__global__ void __launch_bounds__(288, 2)
layernorm96_bwd_pipe_kernel(const bf16* __restrict__ x, const bf16* __restrict__ g, const float* __restrict__ gamma,
                            const bf16* __restrict__ res, bf16* __restrict__ dx, size_t ntok, float* __restrict__ partial) {
    extern __shared__ uint8_t lnq_raw[];
    uint8_t* stages = (uint8_t*)(((uintptr_t)lnq_raw + 127) & ~(uintptr_t)127);
    uint64_t* full = (uint64_t*)(stages + LNQ_STAGES * LNQ_STAGE_BYTES);
    uint64_t* empty = full + LNQ_STAGES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < LNQ_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 8); }
        fence_barrier_init();
    }
    __syncthreads();
    const size_t ntiles = (ntok + LNQ_TOK - 1) / LNQ_TOK;
    const int sub = lane & 15, half = lane >> 4;
    const bool act = sub < 12;
    const int c0 = (act ? sub : 0) * 8;
    float gam[8], dg[8], db[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { gam[j] = __ldg(gamma + c0 + j); dg[j] = 0.f; db[j] = 0.f; }
    if (warp == 8) {
        if (lane == 0) {
            uint32_t it = 0;
            for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
                const uint32_t s = it % LNQ_STAGES;
                mbar_wait(&empty[s], ((it / LNQ_STAGES) & 1) ^ 1);
                const size_t t0 = tile * LNQ_TOK;
                const uint32_t bytes = (uint32_t)(min((size_t)LNQ_TOK, ntok - t0) * 192);
                uint8_t* st = stages + s * LNQ_STAGE_BYTES;
                mbar_arrive_expect_tx(&full[s], bytes * (res ? 3 : 2));
                bulk_load_1d(st, x + t0 * 96, bytes, &full[s]);
                bulk_load_1d(st + LNQ_TILE_BYTES, g + t0 * 96, bytes, &full[s]);
                if (res) bulk_load_1d(st + 2 * LNQ_TILE_BYTES, res + t0 * 96, bytes, &full[s]);
            }
        }
    } else {
        uint32_t it = 0;
        for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
            const uint32_t s = it % LNQ_STAGES;
            mbar_wait(&full[s], (it / LNQ_STAGES) & 1);
            const uint8_t* st = stages + s * LNQ_STAGE_BYTES;
            const size_t t0 = tile * LNQ_TOK;
#pragma unroll 2
            for (int k = 0; k < 4; ++k) {
                const size_t tok = t0 + warp * 8 + 2 * k + half;
                const bool valid = tok < ntok;  // uniform over the 16-lane group
                const int off = (warp * 8 + 2 * k + half) * 192 + c0 * 2;
                const uint4 xr = *(const uint4*)(st + off), gr = *(const uint4*)(st + LNQ_TILE_BYTES + off);
                const uint4 rr = res ? *(const uint4*)(st + 2 * LNQ_TILE_BYTES + off) : make_uint4(0, 0, 0, 0);
                float xv[8], gv[8];
                {
                    const __nv_bfloat162* xh = (const __nv_bfloat162*)&xr;
                    const __nv_bfloat162* gh = (const __nv_bfloat162*)&gr;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float2 a2 = __bfloat1622float2(xh[q]), b2 = __bfloat1622float2(gh[q]);
                        xv[2 * q] = a2.x; xv[2 * q + 1] = a2.y; gv[2 * q] = b2.x; gv[2 * q + 1] = b2.y;
                    }
                }
                float sum = 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) sum += xv[j];
                const float mean = half16_sum(act ? sum : 0.f) * (1.f / 96.f);
                float q2 = 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) q2 += (xv[j] - mean) * (xv[j] - mean);
                const float rstd = rsqrtf(half16_sum(act ? q2 : 0.f) * (1.f / 96.f) + 1e-5f);
                float m1 = 0.f, m2 = 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xv[j] = (xv[j] - mean) * rstd;  // xhat
                    const float gg = gv[j] * gam[j];
                    m1 += gg;
                    m2 = fmaf(gg, xv[j], m2);
                    if (act && valid) {
                        dg[j] = fmaf(gv[j], xv[j], dg[j]);
                        db[j] += gv[j];
                    }
                }
                m1 = half16_sum(act ? m1 : 0.f) * (1.f / 96.f);
                m2 = half16_sum(act ? m2 : 0.f) * (1.f / 96.f);
                if (act && valid) {
                    float o[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] = (gv[j] * gam[j] - m1 - xv[j] * m2) * rstd;
                    const __nv_bfloat162* rh = (const __nv_bfloat162*)&rr;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float2 f = __bfloat1622float2(rh[q]);
                        o[2 * q] += f.x; o[2 * q + 1] += f.y;
                    }
                    uint4 ov;
                    __nv_bfloat162* oh = (__nv_bfloat162*)&ov;
#pragma unroll
                    for (int q = 0; q < 4; ++q) oh[q] = __floats2bfloat162_rn(o[2 * q], o[2 * q + 1]);
                    __stcs((uint4*)(dx + tok * 96 + c0), ov);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);  // this warp has read its 8 tokens: the stage may be refilled
        }
    }
    // per-block dgamma / dbeta: the ring is idle now (every issued copy has been consumed) and takes the fold
    __syncthreads();
    float* red = (float*)stages;  // [16 groups][2][96]
    if (warp < 8 && act) {
        const int grp = warp * 2 + half;
#pragma unroll
        for (int j = 0; j < 8; ++j) { red[(grp * 2 + 0) * 96 + c0 + j] = dg[j]; red[(grp * 2 + 1) * 96 + c0 + j] = db[j]; }
    }
    __syncthreads();
    if (threadIdx.x < 192) {
        const int w = threadIdx.x / 96, c = threadIdx.x % 96;
        float s2 = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k) s2 += red[(k * 2 + w) * 96 + c];
        partial[((size_t)blockIdx.x * 2 + w) * 96 + c] = s2;
    }
}

__global__ void __launch_bounds__(128)
layernorm96_bwd_final_kernel(const float* __restrict__ partial, int nblk, float* __restrict__ dgamma,
                             float* __restrict__ dbeta, int accumulate) {
    __shared__ double sh[128];
    const int w = blockIdx.x / 96, c = blockIdx.x % 96;
    double s = 0.0;
    for (int i = threadIdx.x; i < nblk; i += 128) s += (double)partial[((size_t)i * 2 + w) * 96 + c];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int k = 64; k > 0; k >>= 1) {
        if (threadIdx.x < k) sh[threadIdx.x] += sh[threadIdx.x + k];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        float* o = (w == 0 ? dgamma : dbeta) + c;
        *o = accumulate ? *o + (float)sh[0] : (float)sh[0];
    }
}

__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_exact_grad(float x) {
    return 0.5f * (1.f + erff(x * 0.70710678118654752f)) + x * 0.3989422804014327f * __expf(-0.5f * x * x);
}

// mode 0: out = GELU(pre)      mode 1: out = g * GELU'(pre)   (out may alias g); [npx][C] slices with channel pitches
__global__ void __launch_bounds__(256)
gelu_kernel(bf16* __restrict__ out, int out_pitch, const bf16* __restrict__ pre, int pre_pitch, const bf16* g, int g_pitch,
            size_t npx, int C, int mode) {
    const int cg = C >> 3;
    const size_t n = npx * cg;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t px = i / cg;
        const int c = (int)(i - px * cg) * 8;
        const uint4 pv = __ldg((const uint4*)(pre + px * pre_pitch + c));
        uint4 gv = make_uint4(0, 0, 0, 0);
        if (mode) gv = *(const uint4*)(g + px * g_pitch + c);
        const __nv_bfloat162* ph = (const __nv_bfloat162*)&pv;
        const __nv_bfloat162* gh = (const __nv_bfloat162*)&gv;
        uint4 ov;
        __nv_bfloat162* oh = (__nv_bfloat162*)&ov;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 p = __bfloat1622float2(ph[k]);
            if (mode) {
                const float2 gg = __bfloat1622float2(gh[k]);
                oh[k] = __floats2bfloat162_rn(gg.x * gelu_exact_grad(p.x), gg.y * gelu_exact_grad(p.y));
            } else {
                oh[k] = __floats2bfloat162_rn(gelu_exact(p.x), gelu_exact(p.y));
            }
        }
        *(uint4*)(out + px * out_pitch + c) = ov;
    }
}

}  // namespace

size_t layernorm96_bwd_scratch_bytes() { return (size_t)LNB_BLOCKS * 2 * 96 * sizeof(float); }

int layernorm96_bwd(const bf16* x, const bf16* g, const float* gamma, const void* res, int res_f32, bf16* dx,
                    float* dgamma, float* dbeta, void* scratch, size_t ntok, int accumulate, cudaStream_t s) {
    if ((((uintptr_t)x | (uintptr_t)g | (uintptr_t)dx | (uintptr_t)res) & 15)) {
        set_error("layernorm96_bwd: pointers must be 16-byte aligned");
        return SRB_E_ARG;
    }
    size_t want = (ntok * 16 + 255) / 256;
    int grid = (int)(want < 1 ? 1 : (want > LNB_BLOCKS ? LNB_BLOCKS : want));
    if (ntok >= 16384 && !(res && res_f32)) {
        static std::atomic<unsigned long long> smem_set{0};
        SRB_TRY(opt_in_dynamic_smem(layernorm96_bwd_pipe_kernel, LNQ_SMEM, smem_set));
        const size_t ntiles = (ntok + LNQ_TOK - 1) / LNQ_TOK;
        const size_t cap = (size_t)num_sms() * 2 < (size_t)LNB_BLOCKS ? (size_t)num_sms() * 2 : (size_t)LNB_BLOCKS;
        grid = (int)(ntiles < cap ? ntiles : cap);
        layernorm96_bwd_pipe_kernel<<<grid, 288, LNQ_SMEM, s>>>(x, g, gamma, (const bf16*)res, dx, ntok, (float*)scratch);
    } else {
        layernorm96_bwd_kernel<<<grid, 256, 0, s>>>(x, g, gamma, res, res_f32, dx, ntok, (float*)scratch);
    }
    SRB_CUDA(cudaGetLastError());
    layernorm96_bwd_final_kernel<<<192, 128, 0, s>>>((const float*)scratch, grid, dgamma, dbeta, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int gelu_fwd_bwd(bf16* out, int out_pitch, const bf16* pre, int pre_pitch, const bf16* g, int g_pitch, size_t npx, int C,
                 cudaStream_t s) {
    if (C % 8 || out_pitch % 8 || pre_pitch % 8 || (g && g_pitch % 8) ||
        (((uintptr_t)out | (uintptr_t)pre | (uintptr_t)g) & 15)) {
        set_error("gelu: C and pitches must be multiples of 8, pointers 16-byte aligned");
        return SRB_E_ARG;
    }
    gelu_kernel<<<grid_for(npx * (C / 8)), 256, 0, s>>>(out, out_pitch, pre, pre_pitch, g, g_pitch, npx, C, g ? 1 : 0);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
