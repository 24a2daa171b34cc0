// Element-wise backward operators of the RRDBNet training path (autograd of LeakyReLU,
// residual scaling, F.interpolate(nearest, 2x) inside BasicSR RRDBNet).  All tensors are NHWC slices: [npx][C] with a
// channel pitch, so a slice of a dense-block gradient buffer is addressed in place.  HBM-bound, 16-byte accesses.
#include <cuda_bf16.h>

#include "common.h"
#include "ops.h"

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
axpby_kernel(void* out, int out_f32, int out_pitch, float a, const void* x, int x_f32, int x_pitch, float b,
             const void* y, int y_f32, int y_pitch, size_t npx, int C) {
    const int cg = C >> 2;
    const size_t n = npx * cg;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t px = i / cg;
        const int c = (int)(i - px * cg) * 4;
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
          int y_f32, int y_pitch, size_t npx, int C, cudaStream_t s) {
    if (C % 4 || out_pitch % 4 || x_pitch % 4 || (y && y_pitch % 4) ||
        (((uintptr_t)out | (uintptr_t)x | (uintptr_t)y) & 7)) {
        set_error("axpby: C and pitches must be multiples of 4, pointers 8-byte aligned (16 for fp32)");
        return SRB_E_ARG;
    }
    if ((out_f32 && ((uintptr_t)out & 15)) || (x_f32 && ((uintptr_t)x & 15)) || (y && y_f32 && ((uintptr_t)y & 15))) {
        set_error("axpby: fp32 operands must be 16-byte aligned");
        return SRB_E_ARG;
    }
    axpby_kernel<<<grid_for(npx * (C / 4)), 256, 0, s>>>(out, out_f32, out_pitch, a, x, x_f32, x_pitch, b, y, y_f32,
                                                         y_pitch, npx, C);
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
