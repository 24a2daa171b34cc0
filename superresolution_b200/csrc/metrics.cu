// HBM-bound neighbours of the model forward:
//   weighted_l1      — CombinedLoss.forward, reference src/losses.py:9-21 (+ analytic gradient for the train step)
//   psnr_ssim        — Metrics.update / ssim_torch, reference src/metrics.py:5-32
//   log_normalize    — Dataset_step4 inner loop, reference scripts/Dataset_step4_normalization.py:161-179
//   blend / finalize — halo-blend stitch of overlapping tiles for whole-field inference (ours; §8e)
// All reductions: warp shuffle -> shared -> per-block partial -> fixed-order final pass (deterministic, no atomics).
#include <cfloat>

#include "common.h"
#include "ops.h"

namespace srb {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// block-wide sum; result valid in thread 0.  blockDim.x multiple of 32, <= 1024.
__device__ __forceinline__ float block_sum(float v, float* sh /*[32]*/) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = l < (blockDim.x >> 5) ? sh[l] : 0.f;
        v = warp_sum(v);
    }
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// weighted L1: loss = mean(|p - t| * (t > 0.02 ? 500 : 1));  grad = sign(p - t) * w / N  (optional)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
weighted_l1_kernel(const float* __restrict__ p, const float* __restrict__ t, float* __restrict__ grad, size_t n,
                   float inv_n, float* __restrict__ partial) {
    __shared__ float sh[32];
    float acc = 0.f;
    const size_t n4 = n / 4;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = ((const float4*)p)[i], b = ((const float4*)t)[i];
        const float pa[4] = {a.x, a.y, a.z, a.w}, tb[4] = {b.x, b.y, b.z, b.w};
        float g[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float d = pa[k] - tb[k];
            const float w = tb[k] > 0.02f ? 500.f : 1.f;
            acc += fabsf(d) * w;
            g[k] = (d > 0.f ? w : (d < 0.f ? -w : 0.f)) * inv_n;
        }
        if (grad) ((float4*)grad)[i] = make_float4(g[0], g[1], g[2], g[3]);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // tail
        const size_t i = n4 * 4 + threadIdx.x;
        const float d = p[i] - t[i];
        const float w = t[i] > 0.02f ? 500.f : 1.f;
        acc += fabsf(d) * w;
        if (grad) grad[i] = (d > 0.f ? w : (d < 0.f ? -w : 0.f)) * inv_n;
    }
    const float s = block_sum(acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

__global__ void __launch_bounds__(256)
final_mean_kernel(const float* __restrict__ partial, int nb, double inv_n, float* __restrict__ out) {
    __shared__ double sh[256];
    double a = 0.0;
    for (int i = threadIdx.x; i < nb; i += 256) a += (double)partial[i];
    sh[threadIdx.x] = a;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = (float)(sh[0] * inv_n);
}

size_t reduce_scratch_bytes() { return 148 * 8 * sizeof(float) * 4; }

int weighted_l1(const float* pred, const float* target, size_t n, float* loss, float* grad, float* scratch,
                cudaStream_t s) {
    if (((uintptr_t)pred | (uintptr_t)target | (uintptr_t)grad) & 15) { set_error("weighted_l1: pointers must be 16-byte aligned"); return SRB_E_ARG; }
    int grid = (int)((n / 4 + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    if (grid < 1) grid = 1;
    weighted_l1_kernel<<<grid, 256, 0, s>>>(pred, target, grad, n, 1.f / (float)n, scratch);
    SRB_CUDA(cudaGetLastError());
    final_mean_kernel<<<1, 256, 0, s>>>(scratch, grid, 1.0 / (double)n, loss);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// PSNR + SSIM.  One block = 32x32 tile: per-image squared-error partial over the tile's own pixels, and the SSIM
// map (11x11 Gaussian sigma 1.5, *valid* convolution, C1=0.01^2, C2=0.03^2) over the tile's valid outputs, computed
// separably from a 42x42 clamped input patch in shared memory.
// ---------------------------------------------------------------------------------------------------------------
struct GaussWin { float g[11]; };

__global__ void __launch_bounds__(256)
psnr_ssim_tile_kernel(const float* __restrict__ p, const float* __restrict__ t, int H, int W, GaussWin win,
                      float* __restrict__ part_se, float* __restrict__ part_ssim) {
    __shared__ float sp[42][43], st[42][43];
    __shared__ float hz[5][42][32];
    __shared__ float sh[32];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32, b = blockIdx.z;
    const float* pb = p + (size_t)b * H * W;
    const float* tb = t + (size_t)b * H * W;
    float se = 0.f;
    for (int i = threadIdx.x; i < 42 * 42; i += 256) {
        const int ly = i / 42, lx = i % 42;
        const int y = by + ly, x = bx + lx;
        float a = 0.f, c = 0.f;
        if (y < H && x < W) {
            a = fminf(fmaxf(pb[(size_t)y * W + x], 0.f), 1.f);
            c = fminf(fmaxf(tb[(size_t)y * W + x], 0.f), 1.f);
            if (ly < 32 && lx < 32) se += (a - c) * (a - c);
        }
        sp[ly][lx] = a;
        st[ly][lx] = c;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 42 * 32; i += 256) {
        const int ly = i / 32, ox = i % 32;
        float m1 = 0.f, m2 = 0.f, s11 = 0.f, s22 = 0.f, s12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float a = sp[ly][ox + k], c = st[ly][ox + k], g = win.g[k];
            m1 = fmaf(g, a, m1);
            m2 = fmaf(g, c, m2);
            s11 = fmaf(g, a * a, s11);
            s22 = fmaf(g, c * c, s22);
            s12 = fmaf(g, a * c, s12);
        }
        hz[0][ly][ox] = m1; hz[1][ly][ox] = m2; hz[2][ly][ox] = s11; hz[3][ly][ox] = s22; hz[4][ly][ox] = s12;
    }
    __syncthreads();
    float ss = 0.f;
    const int OH = H - 10, OW = W - 10;
    for (int i = threadIdx.x; i < 32 * 32; i += 256) {
        const int oy = i / 32, ox = i % 32;
        if (by + oy >= OH || bx + ox >= OW) continue;
        float m1 = 0.f, m2 = 0.f, s11 = 0.f, s22 = 0.f, s12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float g = win.g[k];
            m1 = fmaf(g, hz[0][oy + k][ox], m1);
            m2 = fmaf(g, hz[1][oy + k][ox], m2);
            s11 = fmaf(g, hz[2][oy + k][ox], s11);
            s22 = fmaf(g, hz[3][oy + k][ox], s22);
            s12 = fmaf(g, hz[4][oy + k][ox], s12);
        }
        const float v1 = s11 - m1 * m1, v2 = s22 - m2 * m2, cv = s12 - m1 * m2;
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        ss += ((2.f * m1 * m2 + C1) * (2.f * cv + C2)) / ((m1 * m1 + m2 * m2 + C1) * (v1 + v2 + C2));
    }
    const float se_b = block_sum(se, sh);
    const float ss_b = block_sum(ss, sh);
    if (threadIdx.x == 0) {
        const size_t blk = ((size_t)b * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        part_se[blk] = se_b;
        part_ssim[blk] = ss_b;
    }
}

// out[0] = sum_b 10*log10(1/(mse_b + 1e-8)) ; out[1] = (sum of all SSIM map values / (B*OH*OW)) * B
__global__ void __launch_bounds__(256)
psnr_ssim_final_kernel(const float* __restrict__ part_se, const float* __restrict__ part_ssim, int B, int tiles,
                       double inv_hw, double inv_ohw, float* __restrict__ out) {
    __shared__ double sh[256];
    double psnr = 0.0, ssim = 0.0;
    for (int b = 0; b < B; ++b) {
        double a = 0.0, c = 0.0;
        for (int i = threadIdx.x; i < tiles; i += 256) {
            a += (double)part_se[(size_t)b * tiles + i];
            c += (double)part_ssim[(size_t)b * tiles + i];
        }
        sh[threadIdx.x] = a;
        __syncthreads();
        for (int s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s]; __syncthreads(); }
        const double se = sh[0];
        __syncthreads();
        sh[threadIdx.x] = c;
        __syncthreads();
        for (int s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s]; __syncthreads(); }
        const double sv = sh[0];
        __syncthreads();
        const float mse = (float)(se * inv_hw);
        psnr += (double)(10.f * log10f(1.0f / (mse + 1e-8f)));
        ssim += sv;
    }
    if (threadIdx.x == 0) {
        out[0] = (float)psnr;
        out[1] = (float)(ssim * inv_ohw);  // == mean over [B,1,OH,OW] times B
    }
}

size_t psnr_ssim_scratch_bytes(int B, int H, int W) {
    return (size_t)B * ((H + 31) / 32) * ((W + 31) / 32) * 2 * sizeof(float);
}

int psnr_ssim(const float* pred, const float* target, int B, int H, int W, float* out2, float* scratch,
              cudaStream_t s) {
    if (H < 11 || W < 11) { set_error("psnr_ssim: image smaller than the 11x11 SSIM window"); return SRB_E_SHAPE; }
    GaussWin win;
    {   // reference: gauss = exp(-(x-5)^2 / (2*1.5^2)), normalised in fp32 (src/metrics.py:11-12)
        float g[11], sum = 0.f;
        for (int i = 0; i < 11; ++i) { g[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); sum += g[i]; }
        for (int i = 0; i < 11; ++i) win.g[i] = g[i] / sum;
    }
    dim3 grid((W + 31) / 32, (H + 31) / 32, B);
    const int tiles = grid.x * grid.y;
    float* part_se = scratch;
    float* part_ssim = scratch + (size_t)B * tiles;
    psnr_ssim_tile_kernel<<<grid, 256, 0, s>>>(pred, target, H, W, win, part_se, part_ssim);
    SRB_CUDA(cudaGetLastError());
    psnr_ssim_final_kernel<<<1, 256, 0, s>>>(part_se, part_ssim, B, tiles, 1.0 / ((double)H * W),
                                             1.0 / ((double)(H - 10) * (W - 10)), out2);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// LOG normalisation: u16( clip((log1p(max(nan_to_num(x),0)) - gmin) / denom, 0, 1) * 65535 )   (truncating cast)
// denom = float32(gmax - gmin + 1e-8) is formed by the caller exactly as numpy does.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint16_t lognorm1(float x, float gmin, float denom) {
    if (isnan(x)) x = 0.f;
    else if (isinf(x)) x = x > 0.f ? FLT_MAX : -FLT_MAX;
    float d = log1pf(fmaxf(x, 0.f));
    d = __fdiv_rn(d - gmin, denom);
    d = fminf(fmaxf(d, 0.f), 1.f);
    return (uint16_t)(d * 65535.f);
}

__global__ void __launch_bounds__(256)
log_normalize_kernel(const float* __restrict__ x, uint16_t* __restrict__ out, size_t n, float gmin, float denom) {
    const size_t n8 = n / 8;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = __ldcs((const float4*)x + 2 * i), b = __ldcs((const float4*)x + 2 * i + 1);
        uint4 o;
        o.x = lognorm1(a.x, gmin, denom) | ((uint32_t)lognorm1(a.y, gmin, denom) << 16);
        o.y = lognorm1(a.z, gmin, denom) | ((uint32_t)lognorm1(a.w, gmin, denom) << 16);
        o.z = lognorm1(b.x, gmin, denom) | ((uint32_t)lognorm1(b.y, gmin, denom) << 16);
        o.w = lognorm1(b.z, gmin, denom) | ((uint32_t)lognorm1(b.w, gmin, denom) << 16);
        __stcs((uint4*)out + i, o);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
        const size_t i = n8 * 8 + threadIdx.x;
        out[i] = lognorm1(x[i], gmin, denom);
    }
}

int log_normalize_u16(const float* x, uint16_t* out, size_t n, float gmin, float denom, cudaStream_t s) {
    if (((uintptr_t)x & 15) || ((uintptr_t)out & 15)) { set_error("log_normalize: pointers must be 16-byte aligned"); return SRB_E_ARG; }
    int grid = (int)((n / 8 + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    if (grid < 1) grid = 1;
    log_normalize_kernel<<<grid, 256, 0, s>>>(x, out, n, gmin, denom);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
