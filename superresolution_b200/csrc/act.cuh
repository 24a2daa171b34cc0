// Activation helpers shared by the tensor-core epilogues.
#pragma once
#include <cuda_runtime.h>

namespace srb {

__device__ __forceinline__ float tanh_approx(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// nn.GELU() (exact, erf form) evaluated as 0.5 x (1 + tanh(x (c0 + c1 x^2 + c2 x^4))): the odd polynomial is a
// minimax fit of atanh(erf(x / sqrt 2)) (max |gelu_fast - exact GELU| = 2.5e-5 over R in exact arithmetic, ~1e-4 with
// the 2^-11 relative error of MUFU.TANH) — an order of magnitude below the bf16 rounding of the value that is stored.
// 7 FP32-pipe operations + 1 MUFU per element instead of ~14 + 2 for the Abramowitz-Stegun erf.
__device__ __forceinline__ float gelu_fast(float x) {
    const float x2 = fminf(x * x, 64.f);  // the fit holds on |x| <= 8; beyond it tanh has saturated
    float p = fmaf(x2, -3.51516789e-4f, 3.70056460e-2f);
    p = fmaf(p, x2, 7.97507884e-1f);
    const float t = tanh_approx(x * p);
    const float hx = 0.5f * x;
    return fmaf(hx, t, hx);
}

}  // namespace srb
