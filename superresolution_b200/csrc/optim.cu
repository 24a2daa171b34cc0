// Optimizer half of the training step (reference scripts/Modello_supporto.py:107,133-141:
// GradScaler.unscale_ -> clip_grad_norm_(1.0) -> AdamW(lr, weight_decay=0).step).  The reference runs these as three
// passes of per-tensor kernels over 1285 parameter tensors; here the parameters, gradients and both moments live in
// flat fp32 buffers and the step is two HBM-bound kernels with no host synchronisation:
//   grad_sumsq : deterministic two-level sum of squares of the (still scaled) gradient + a non-finite flag
//   adamw_step : coef = inv_scale * min(1, max_norm / (||g|| * inv_scale + 1e-6)) read from the device-side norm,
//                then the AdamW update of torch.optim.AdamW (decoupled weight decay, bias correction); the whole step
//                is skipped when the gradient held an inf/nan (GradScaler semantics).  Optionally refreshes a bf16
//                shadow of the parameters in the same pass.
// Traffic per parameter: sumsq 4 B read; step 16 B read (p, g, m, v) + 12 B write (p, m, v) [+ 2 B shadow].
#include <cstdint>

#include "common.h"
#include "ops.h"

namespace srb {

namespace {

constexpr int kRedBlocks = 148 * 8;  // partial sums (fixed grid -> fixed summation order)

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(256)
grad_sumsq_kernel(const float* __restrict__ g, size_t n, float* __restrict__ partial, int* __restrict__ bad) {
    __shared__ float sh[8];
    __shared__ int shbad;
    if (threadIdx.x == 0) shbad = 0;
    float acc = 0.f;
    bool nonfinite = false;
    const size_t n4 = n / 4;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = __ldg((const float4*)g + i);
        acc = fmaf(a.x, a.x, acc);
        acc = fmaf(a.y, a.y, acc);
        acc = fmaf(a.z, a.z, acc);
        acc = fmaf(a.w, a.w, acc);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const float a = g[n4 * 4 + threadIdx.x];
        acc = fmaf(a, a, acc);
    }
    nonfinite = !(acc <= 3.0e38f);  // inf or nan anywhere in this thread's share poisons its partial sum
    acc = warp_sum_f(acc);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    if (nonfinite) shbad = 1;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < 8 ? sh[threadIdx.x] : 0.f;
        v = warp_sum_f(v);
        if (threadIdx.x == 0) {
            partial[blockIdx.x] = v;
            bad[blockIdx.x] = shbad;
        }
    }
}

// info[0] = sum of squares (fp32), info[1] = 1.0 if any non-finite value was seen else 0.0
__global__ void __launch_bounds__(256)
grad_sumsq_final_kernel(const float* __restrict__ partial, const int* __restrict__ bad, int nb, float* __restrict__ info) {
    __shared__ double sh[256];
    __shared__ int shbad;
    if (threadIdx.x == 0) shbad = 0;
    __syncthreads();
    double a = 0.0;
    int b = 0;
    for (int i = threadIdx.x; i < nb; i += 256) { a += (double)partial[i]; b |= bad[i]; }
    sh[threadIdx.x] = a;
    if (b) shbad = 1;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float tot = (float)sh[0];
        info[0] = tot;
        info[1] = (shbad || !(tot <= 3.0e38f)) ? 1.f : 0.f;
    }
}

struct AdamArgs {
    float lr, beta1, beta2, eps, weight_decay;
    float bias_c1, bias_c2_sqrt;  // 1 - beta1^t, sqrt(1 - beta2^t)
    float inv_scale, max_norm;    // gradient pre-scale (1/loss_scale/world) and clip threshold (<= 0: no clipping)
};

__device__ __forceinline__ void adam1(float& p, float g, float& m, float& v, const AdamArgs& a, float coef) {
    g *= coef;
    p *= 1.f - a.lr * a.weight_decay;
    m = fmaf(1.f - a.beta1, g - m, m);                    // exp_avg.lerp_(grad, 1 - beta1)
    v = fmaf(1.f - a.beta2, g * g, a.beta2 * v);          // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    const float denom = sqrtf(v) / a.bias_c2_sqrt + a.eps;
    p = fmaf(-(a.lr / a.bias_c1), m / denom, p);          // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(256)
adamw_step_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                  size_t n, AdamArgs a, const float* __restrict__ info, __nv_bfloat16* __restrict__ shadow) {
    float coef = a.inv_scale;
    if (info) {
        if (info[1] != 0.f) return;  // inf/nan gradient: skip the step (uniform across the grid)
        if (a.max_norm > 0.f) {
            const float norm = sqrtf(info[0]) * a.inv_scale;
            coef *= fminf(1.f, a.max_norm / (norm + 1e-6f));  // clip_grad_norm_: coef clamped to 1
        }
    }
    const size_t n4 = n / 4;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 pp = ((float4*)p)[i], mm = ((float4*)m)[i], vv = ((float4*)v)[i];
        const float4 gg = __ldg((const float4*)g + i);
        adam1(pp.x, gg.x, mm.x, vv.x, a, coef);
        adam1(pp.y, gg.y, mm.y, vv.y, a, coef);
        adam1(pp.z, gg.z, mm.z, vv.z, a, coef);
        adam1(pp.w, gg.w, mm.w, vv.w, a, coef);
        ((float4*)p)[i] = pp;
        ((float4*)m)[i] = mm;
        ((float4*)v)[i] = vv;
        if (shadow) {
            __nv_bfloat162* s2 = (__nv_bfloat162*)shadow + 2 * i;
            s2[0] = __floats2bfloat162_rn(pp.x, pp.y);
            s2[1] = __floats2bfloat162_rn(pp.z, pp.w);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const size_t i = n4 * 4 + threadIdx.x;
        float pp = p[i], mm = m[i], vv = v[i];
        adam1(pp, g[i], mm, vv, a, coef);
        p[i] = pp; m[i] = mm; v[i] = vv;
        if (shadow) shadow[i] = __float2bfloat16_rn(pp);
    }
}

}  // namespace

size_t grad_norm_scratch_bytes() { return (size_t)kRedBlocks * (sizeof(float) + sizeof(int)); }

int grad_sumsq(const float* g, size_t n, float* info, void* scratch, cudaStream_t s) {
    if (((uintptr_t)g & 15) || ((uintptr_t)scratch & 3)) { set_error("grad_sumsq: gradient must be 16-byte aligned"); return SRB_E_ARG; }
    float* partial = (float*)scratch;
    int* bad = (int*)(partial + kRedBlocks);
    int grid = (int)((n / 4 + 255) / 256);
    if (grid > kRedBlocks) grid = kRedBlocks;
    if (grid < 1) grid = 1;
    grad_sumsq_kernel<<<grid, 256, 0, s>>>(g, n, partial, bad);
    SRB_CUDA(cudaGetLastError());
    grad_sumsq_final_kernel<<<1, 256, 0, s>>>(partial, bad, grid, info);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int adamw_step(float* p, const float* g, float* m, float* v, size_t n, float lr, float beta1, float beta2, float eps,
               float weight_decay, int step, float inv_scale, float max_norm, const float* info, void* shadow_bf16,
               cudaStream_t s) {
    if (step < 1) { set_error("adamw_step: step counts from 1"); return SRB_E_ARG; }
    if (((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15 || ((uintptr_t)shadow_bf16 & 7)) {
        set_error("adamw_step: buffers must be 16-byte aligned");
        return SRB_E_ARG;
    }
    AdamArgs a;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
    a.bias_c1 = (float)(1.0 - pow((double)beta1, (double)step));
    a.bias_c2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    a.inv_scale = inv_scale;
    a.max_norm = max_norm;
    int grid = (int)((n / 4 + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    if (grid < 1) grid = 1;
    adamw_step_kernel<<<grid, 256, 0, s>>>(p, g, m, v, n, a, info, (__nv_bfloat16*)shadow_bf16);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
