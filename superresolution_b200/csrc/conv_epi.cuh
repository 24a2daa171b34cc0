// Fused epilogue arithmetic shared by the tensor-core conv kernels (conv_tc.cu, conv_chain.cu).
#pragma once
#include "act.cuh"
#include "common.h"

namespace srb {

// GELU'(x) = Phi(x) + x phi(x) with Phi from the tanh fit of act.cuh (|error| ~1e-4, below the bf16 rounding of the
// gradient it multiplies): 10 FP32-pipe operations + 2 MUFU.  The TRAIN epilogues are bound by this arithmetic (8 epilogue
// warps per SM do the work a 64-warp element-wise kernel would), so the cheaper forms are used there.
__device__ __forceinline__ float gelu_fast_grad(float x) {
    const float xx = x * x;
    const float x2 = fminf(xx, 64.f);
    float p = fmaf(x2, -3.51516789e-4f, 3.70056460e-2f);
    p = fmaf(p, x2, 7.97507884e-1f);
    const float t = tanh_approx(x * p);
    const float phi = 0.3989422804014327f * __expf(-0.5f * xx);
    return fmaf(x, phi, fmaf(0.5f, t, 0.5f));
}

// bias + activation + residuals on one 32-channel slice of one pixel (all indices compile-time after unrolling)
// sbias: the layer's bias staged in shared memory by the kernel prologue (zeros when the layer has none); a global
// __ldg per unit sat on the epilogue's critical path (~500 cycles per 128 x 32 block).
// g_row (TRAIN) / r1_row: this thread's 64-byte row of the saved GELU pre-activation / of the bf16 res1 operand, already
// staged in shared memory by TMA (16-byte chunk j at row + ((j ^ sw) << 4)); null = read it from global memory here
template <bool TRAIN = false>
__device__ __forceinline__ void epilogue_math(float (&v)[32], const ConvArgs& a, size_t pix, int cb, int b,
                                              const float* sbias = nullptr, const uint8_t* g_row = nullptr,
                                              const uint8_t* r1_row = nullptr, uint32_t g_sw = 0, bool pre_done = false) {
    // pre_done (TRAIN): the caller has added the bias and sent the pre-activation out already (TMA store)
    if (TRAIN && pre_done) {
    } else if (sbias) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float4 bb = *(const float4*)(sbias + cb + j);
            v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
        }
    } else if (a.bias) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float4 bb = __ldg((const float4*)(a.bias + cb + j));
            v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
        }
    }
    if (TRAIN && a.pre_out && !pre_done) {
        uint4* pp = (uint4*)(a.pre_out + pix * a.pre_pitch + cb);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint4 o;
            __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
            for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[8 * j + 2 * k], v[8 * j + 2 * k + 1]);
            pp[j] = o;
        }
    }
    if (a.act == ACT_LRELU) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : v[j] * a.slope;
    } else if (a.act == ACT_GELU) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_fast(v[j]);  // tanh fit of the exact (erf) GELU, see act.cuh
    }
    if (TRAIN && a.gelu_grad_of) {
        const uint4* gp = (const uint4*)(a.gelu_grad_of + pix * a.gelu_grad_pitch + cb);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint4 rr = g_row ? *(const uint4*)(g_row + ((j ^ g_sw) << 4)) : __ldg(gp + j);
            const __nv_bfloat162* h = (const __nv_bfloat162*)&rr;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = __bfloat1622float2(h[k]);
                v[8 * j + 2 * k] *= gelu_fast_grad(f.x);
                v[8 * j + 2 * k + 1] *= gelu_fast_grad(f.y);
            }
        }
    }
    if (a.res1) {
        const float al = (TRAIN && a.alpha1_samples) ? __ldg(a.alpha1_samples + b) : a.alpha1;
        if (a.res1_f32) {
            const float4* rp = (const float4*)((const float*)a.res1 + pix * a.res1_pitch + cb);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 rr = rp[j];
                v[4 * j] = fmaf(v[4 * j], al, rr.x);
                v[4 * j + 1] = fmaf(v[4 * j + 1], al, rr.y);
                v[4 * j + 2] = fmaf(v[4 * j + 2], al, rr.z);
                v[4 * j + 3] = fmaf(v[4 * j + 3], al, rr.w);
            }
        } else {
            const uint4* rp = (const uint4*)((const bf16*)a.res1 + pix * a.res1_pitch + cb);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 rr = r1_row ? *(const uint4*)(r1_row + ((j ^ g_sw) << 4)) : rp[j];
                const __nv_bfloat162* h = (const __nv_bfloat162*)&rr;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = __bfloat1622float2(h[k]);
                    v[8 * j + 2 * k] = fmaf(v[8 * j + 2 * k], al, f.x);
                    v[8 * j + 2 * k + 1] = fmaf(v[8 * j + 2 * k + 1], al, f.y);
                }
            }
        }
    }
    if (a.res2) {
        if (a.res2_f32) {
            const float4* rp = (const float4*)((const float*)a.res2 + pix * a.res2_pitch + cb);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 rr = rp[j];
                float4 c4 = make_float4(1.f, 1.f, 1.f, 1.f);
                if (a.res2_cs) c4 = __ldg((const float4*)(a.res2_cs + (size_t)b * a.Cout + cb + 4 * j));
                v[4 * j] = fmaf(v[4 * j], a.alpha2, rr.x * c4.x);
                v[4 * j + 1] = fmaf(v[4 * j + 1], a.alpha2, rr.y * c4.y);
                v[4 * j + 2] = fmaf(v[4 * j + 2], a.alpha2, rr.z * c4.z);
                v[4 * j + 3] = fmaf(v[4 * j + 3], a.alpha2, rr.w * c4.w);
            }
        } else {
            const uint4* rp = (const uint4*)((const bf16*)a.res2 + pix * a.res2_pitch + cb);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 rr = rp[j];
                const __nv_bfloat162* h = (const __nv_bfloat162*)&rr;
                float4 c0 = make_float4(1.f, 1.f, 1.f, 1.f), c1 = c0;
                if (a.res2_cs) {
                    c0 = __ldg((const float4*)(a.res2_cs + (size_t)b * a.Cout + cb + 8 * j));
                    c1 = __ldg((const float4*)(a.res2_cs + (size_t)b * a.Cout + cb + 8 * j + 4));
                }
                const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = __bfloat1622float2(h[k]);
                    v[8 * j + 2 * k] = fmaf(v[8 * j + 2 * k], a.alpha2, f.x * cs[2 * k]);
                    v[8 * j + 2 * k + 1] = fmaf(v[8 * j + 2 * k + 1], a.alpha2, f.y * cs[2 * k + 1]);
                }
            }
        }
    }
}

}  // namespace srb
