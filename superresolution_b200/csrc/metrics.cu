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
    auto one = [&](const float4 a, const float4 b, size_t i) {
        const float pa[4] = {a.x, a.y, a.z, a.w}, tb[4] = {b.x, b.y, b.z, b.w};
        float g[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float d = pa[k] - tb[k];
            const float w = tb[k] > 0.02f ? 500.f : 1.f;
            acc += fabsf(d) * w;
            g[k] = (d > 0.f ? w : (d < 0.f ? -w : 0.f)) * inv_n;
        }
        if (grad) __stcs((float4*)grad + i, make_float4(g[0], g[1], g[2], g[3]));  // streaming: written once, read later
    };
    // four independent 16-byte load pairs in flight per thread (a single pair per iteration left the kernel latency-bound
    // at 19 % of the HBM peak); the summation order stays fixed (deterministic)
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n4; i += 4 * stride) {
        float4 a[4], b[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { a[u] = __ldcs((const float4*)p + i + u * stride); b[u] = __ldcs((const float4*)t + i + u * stride); }
#pragma unroll
        for (int u = 0; u < 4; ++u) one(a[u], b[u], i + u * stride);
    }
    for (; i < n4; i += stride) one(__ldcs((const float4*)p + i), __ldcs((const float4*)t + i), i);
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

// ---------------------------------------------------------------------------------------------------------------
// Stale-variant loss terms (reference Backup/src/losses.py:6-15,37-48), one pass:
//   char = mean(sqrt(d^2 + 1e-6)),  l1_raw = mean(|d|),  astro = mean(sqrt(d^2 + 1e-6) * (1 + 5 t)),  d = p - t
//   grad (optional) = d(l1_w * char + astro_w * astro)/dp = (l1_w + astro_w (1 + 5 t)) * d / sqrt(d^2 + 1e-6) / n
// (the VGG19 perceptual term of that file needs downloaded ImageNet weights and is not built)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
charbonnier_astro_kernel(const float* __restrict__ p, const float* __restrict__ t, float* __restrict__ grad, size_t n,
                         float inv_n, float l1_w, float astro_w, float* __restrict__ partial /*[3][gridDim.x]*/) {
    __shared__ float sh[32];
    float a_char = 0.f, a_l1 = 0.f, a_astro = 0.f;
    auto one = [&](float pv, float tv) -> float {
        const float d = pv - tv;
        const float c = sqrtf(fmaf(d, d, 1e-6f));
        const float w = fmaf(5.f, tv, 1.f);
        a_char += c;
        a_l1 += fabsf(d);
        a_astro = fmaf(c, w, a_astro);
        return (l1_w + astro_w * w) * __fdividef(d, c) * inv_n;
    };
    const size_t n4 = n / 4, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 a = __ldcs((const float4*)p + i), b = __ldcs((const float4*)t + i);
        const float4 g = make_float4(one(a.x, b.x), one(a.y, b.y), one(a.z, b.z), one(a.w, b.w));
        if (grad) __stcs((float4*)grad + i, g);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // tail
        const size_t i = n4 * 4 + threadIdx.x;
        const float g = one(p[i], t[i]);
        if (grad) grad[i] = g;
    }
    const float s0 = block_sum(a_char, sh), s1 = block_sum(a_l1, sh), s2 = block_sum(a_astro, sh);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = s0;
        partial[gridDim.x + blockIdx.x] = s1;
        partial[2 * gridDim.x + blockIdx.x] = s2;
    }
}

int charbonnier_astro(const float* pred, const float* target, size_t n, float l1_w, float astro_w, float* out3, float* grad,
                      float* scratch, cudaStream_t s) {
    if (((uintptr_t)pred | (uintptr_t)target | (uintptr_t)grad) & 15) { set_error("charbonnier_astro: pointers must be 16-byte aligned"); return SRB_E_ARG; }
    int grid = (int)((n / 4 + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    if (grid < 1) grid = 1;
    charbonnier_astro_kernel<<<grid, 256, 0, s>>>(pred, target, grad, n, 1.f / (float)n, l1_w, astro_w, scratch);
    SRB_CUDA(cudaGetLastError());
    for (int k = 0; k < 3; ++k) {  // fixed-order final sums (deterministic)
        final_mean_kernel<<<1, 256, 0, s>>>(scratch + (size_t)k * grid, grid, 1.0 / (double)n, out3 + k);
        SRB_CUDA(cudaGetLastError());
    }
    return SRB_OK;
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
// PSNR + SSIM (11x11 Gaussian sigma 1.5, *valid* convolution, C1 = 0.01^2, C2 = 0.03^2; inputs clamped to [0,1]).
// One block = 64 x 32 output tile.  Pass A (horizontal): each thread filters 4 adjacent outputs of one input row from
// 16 clamped inputs read straight from global (float4, L1 absorbs the overlap), forms a^2, c^2, ac once per input and
// leaves the 4 filtered rows (a, c, a^2 + c^2, ac: SSIM needs the two variances only as a sum) in shared memory; the squared error of the tile's own pixels rides along.  Pass B
// (vertical): each thread owns one column and 8 consecutive output rows, so every shared-memory value feeds up to 8
// accumulators from registers (statically unrolled taps), then evaluates the SSIM map entry.  FP32-FMA bound
// (~95 FMA per pixel), not HBM bound.
// ---------------------------------------------------------------------------------------------------------------
struct GaussWin { float g[11]; };

constexpr int kSsimTX = 64, kSsimTY = 32, kSsimRows = kSsimTY + 10;
constexpr int kSsimQ = 4;  // filtered quantities: a, c, a^2 + c^2, a c (SSIM needs var(a) + var(c) only as a sum)
constexpr int kSsimSmem = kSsimQ * kSsimRows * kSsimTX * (int)sizeof(float);

// four clamped pixels of one row starting at column x (x % 4 == 0); zeros beyond the image
__device__ __forceinline__ float4 load4_sat(const float* __restrict__ row, int x, int W, bool row_ok, bool vec_ok) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row_ok) {
        if (vec_ok && x + 3 < W) {
            v = __ldg((const float4*)(row + x));
        } else {
            if (x < W) v.x = row[x];
            if (x + 1 < W) v.y = row[x + 1];
            if (x + 2 < W) v.z = row[x + 2];
            if (x + 3 < W) v.w = row[x + 3];
        }
    }
    return make_float4(__saturatef(v.x), __saturatef(v.y), __saturatef(v.z), __saturatef(v.w));
}

__global__ void __launch_bounds__(256, 4)
psnr_ssim_tile_kernel(const float* __restrict__ p, const float* __restrict__ t, int H, int W, GaussWin win,
                      float* __restrict__ part_se, float* __restrict__ part_ssim) {
    extern __shared__ float hz[];  // [kSsimQ][kSsimRows][kSsimTX]
    __shared__ float sh[32];
    const int bx = blockIdx.x * kSsimTX, by = blockIdx.y * kSsimTY, b = blockIdx.z;
    const float* pb = p + (size_t)b * H * W;
    const float* tb = t + (size_t)b * H * W;
    const bool vec_ok = (W & 3) == 0 && (((uintptr_t)p | (uintptr_t)t) & 15) == 0;
    float se = 0.f;
    // ---- pass A: 42 rows x 16 groups of 4 outputs
    for (int i = threadIdx.x; i < kSsimRows * (kSsimTX / 4); i += 256) {
        const int ly = i >> 4, gx = (i & 15) * 4;
        const int y = by + ly, x = bx + gx;
        const bool row_ok = y < H;
        const float* pr = pb + (size_t)y * W;
        const float* tr = tb + (size_t)y * W;
        float a[16], c[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 va = load4_sat(pr, x + 4 * j, W, row_ok, vec_ok);
            const float4 vc = load4_sat(tr, x + 4 * j, W, row_ok, vec_ok);
            a[4 * j] = va.x; a[4 * j + 1] = va.y; a[4 * j + 2] = va.z; a[4 * j + 3] = va.w;
            c[4 * j] = vc.x; c[4 * j + 1] = vc.y; c[4 * j + 2] = vc.z; c[4 * j + 3] = vc.w;
        }
        if (ly < kSsimTY) {  // squared error over the tile's own pixels (out-of-image pixels are 0 - 0)
#pragma unroll
            for (int j = 0; j < 4; ++j) se = fmaf(a[j] - c[j], a[j] - c[j], se);
        }
        float o[kSsimQ][4];
#pragma unroll
        for (int q = 0; q < kSsimQ; ++q)
#pragma unroll
            for (int j = 0; j < 4; ++j) o[q][j] = 0.f;
#pragma unroll
        for (int k = 0; k < 14; ++k) {
            const float sq = fmaf(a[k], a[k], c[k] * c[k]), ac = a[k] * c[k];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (k - j >= 0 && k - j < 11) {
                    const float g = win.g[k - j];
                    o[0][j] = fmaf(g, a[k], o[0][j]);
                    o[1][j] = fmaf(g, c[k], o[1][j]);
                    o[2][j] = fmaf(g, sq, o[2][j]);
                    o[3][j] = fmaf(g, ac, o[3][j]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < kSsimQ; ++q)
            *(float4*)(hz + (q * kSsimRows + ly) * kSsimTX + gx) = make_float4(o[q][0], o[q][1], o[q][2], o[q][3]);
    }
    __syncthreads();
    // ---- pass B: column ox, output rows oy0 .. oy0 + 7
    const int ox = threadIdx.x & 63, oy0 = (threadIdx.x >> 6) * 8;
    const int OH = H - 10, OW = W - 10;
    float acc[8][kSsimQ];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int q = 0; q < kSsimQ; ++q) acc[j][q] = 0.f;
#pragma unroll
    for (int r = 0; r < 18; ++r) {
        float v[kSsimQ];
#pragma unroll
        for (int q = 0; q < kSsimQ; ++q) v[q] = hz[(q * kSsimRows + oy0 + r) * kSsimTX + ox];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (r - j >= 0 && r - j < 11) {
                const float g = win.g[r - j];
#pragma unroll
                for (int q = 0; q < kSsimQ; ++q) acc[j][q] = fmaf(g, v[q], acc[j][q]);
            }
        }
    }
    float ss = 0.f;
    if (bx + ox < OW) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (by + oy0 + j < OH) {
                const float m1 = acc[j][0], m2 = acc[j][1];
                const float mm = m1 * m1 + m2 * m2;
                const float vsum = acc[j][2] - mm, cv = acc[j][3] - m1 * m2;  // var(a) + var(c), cov(a, c)
                const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
                ss += __fdividef((2.f * m1 * m2 + C1) * (2.f * cv + C2), (mm + C1) * (vsum + C2));
            }
        }
    }
    const float se_b = block_sum(se, sh);
    const float ss_b = block_sum(ss, sh);
    if (threadIdx.x == 0) {
        const size_t blk = ((size_t)b * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        part_se[blk] = se_b;
        part_ssim[blk] = ss_b;
    }
}

// per image: fixed-order sums of its tile partials -> img[b] = {10*log10(1/(mse_b + 1e-8)), sum of its SSIM map}
__global__ void __launch_bounds__(128)
psnr_ssim_image_kernel(const float* __restrict__ part_se, const float* __restrict__ part_ssim, int tiles,
                       double inv_hw, double* __restrict__ img) {
    __shared__ double sa[128], sc[128];
    const int b = blockIdx.x;
    double a = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < tiles; i += 128) {
        a += (double)part_se[(size_t)b * tiles + i];
        c += (double)part_ssim[(size_t)b * tiles + i];
    }
    sa[threadIdx.x] = a;
    sc[threadIdx.x] = c;
    __syncthreads();
    for (int s = 64; s > 0; s >>= 1) {
        if (threadIdx.x < s) { sa[threadIdx.x] += sa[threadIdx.x + s]; sc[threadIdx.x] += sc[threadIdx.x + s]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float mse = (float)(sa[0] * inv_hw);
        img[2 * b] = (double)(10.f * log10f(1.0f / (mse + 1e-8f)));
        img[2 * b + 1] = sc[0];
    }
}

// out[0] = sum_b psnr_b ; out[1] = (sum of all SSIM map values / (B*OH*OW)) * B
__global__ void __launch_bounds__(256)
psnr_ssim_final_kernel(const double* __restrict__ img, int B, double inv_ohw, float* __restrict__ out) {
    __shared__ double sa[256], sc[256];
    double a = 0.0, c = 0.0;
    for (int b = threadIdx.x; b < B; b += 256) { a += img[2 * b]; c += img[2 * b + 1]; }
    sa[threadIdx.x] = a;
    sc[threadIdx.x] = c;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) { sa[threadIdx.x] += sa[threadIdx.x + s]; sc[threadIdx.x] += sc[threadIdx.x + s]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out[0] = (float)sa[0];
        out[1] = (float)(sc[0] * inv_ohw);  // == mean over [B,1,OH,OW] times B
    }
}

static size_t ssim_tiles(int H, int W) { return (size_t)((H + kSsimTY - 1) / kSsimTY) * ((W + kSsimTX - 1) / kSsimTX); }

size_t psnr_ssim_scratch_bytes(int B, int H, int W) {
    return (size_t)B * 2 * sizeof(double) + (size_t)B * ssim_tiles(H, W) * 2 * sizeof(float);
}

int psnr_ssim(const float* pred, const float* target, int B, int H, int W, float* out2, float* scratch,
              cudaStream_t s) {
    if (H < 11 || W < 11) { set_error("psnr_ssim: image smaller than the 11x11 SSIM window"); return SRB_E_SHAPE; }
    if ((uintptr_t)scratch & 7) { set_error("psnr_ssim: scratch must be 8-byte aligned"); return SRB_E_ARG; }
    GaussWin win;
    {   // reference: gauss = exp(-(x-5)^2 / (2*1.5^2)), normalised in fp32 (src/metrics.py:11-12)
        float g[11], sum = 0.f;
        for (int i = 0; i < 11; ++i) { g[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); sum += g[i]; }
        for (int i = 0; i < 11; ++i) win.g[i] = g[i] / sum;
    }
    // per call: the attribute is per device, and a process may drive several (DataParallel-style threads)
    SRB_CUDA(cudaFuncSetAttribute(psnr_ssim_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSsimSmem));
    dim3 grid((W + kSsimTX - 1) / kSsimTX, (H + kSsimTY - 1) / kSsimTY, B);
    const int tiles = grid.x * grid.y;
    double* img = (double*)scratch;
    float* part_se = (float*)(img + 2 * (size_t)B);
    float* part_ssim = part_se + (size_t)B * tiles;
    psnr_ssim_tile_kernel<<<grid, 256, kSsimSmem, s>>>(pred, target, H, W, win, part_se, part_ssim);
    SRB_CUDA(cudaGetLastError());
    psnr_ssim_image_kernel<<<B, 128, 0, s>>>(part_se, part_ssim, tiles, 1.0 / ((double)H * W), img);
    SRB_CUDA(cudaGetLastError());
    psnr_ssim_final_kernel<<<1, 256, 0, s>>>(img, B, 1.0 / ((double)(H - 10) * (W - 10)), out2);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// psnr_ssim without any allocation by the caller or by us (srb_psnr_ssim_noalloc; no counterpart in the reference,
// added for callers that may not call cudaMalloc: graph capture, serving loops).
// The scratch lives in a file-scope __device__ array that every GPU loading the library holds until exit.  Host code
// may not use the array's name as a pointer (it is a host-side stand-in): the device address comes from
// cudaGetSymbolAddress, looked up once per device with that device current.  Every tile / image partial the kernels
// read is rewritten by the same call, so stale contents of an earlier call never matter; two calls that overlap on
// the GPU (different streams, no dependency) share the array and race.
// ---------------------------------------------------------------------------------------------------------------
constexpr size_t kPsnrSsimStaticBytes = 256 * 1024;  // = SRB_PSNR_SSIM_NOALLOC_MAX_BYTES (include/srb.h)
static __device__ __align__(256) unsigned char g_psnr_ssim_scratch[kPsnrSsimStaticBytes];

int psnr_ssim_noalloc(const float* pred, const float* target, int B, int H, int W, float* out2, cudaStream_t s) {
    if (B <= 0 || H <= 0 || W <= 0) { set_error("psnr_ssim_noalloc: bad shape"); return SRB_E_ARG; }
    const size_t need = psnr_ssim_scratch_bytes(B, H, W);
    if (need > kPsnrSsimStaticBytes) {
        set_error("psnr_ssim_noalloc: [%d,%d,%d] needs %zu bytes of scratch, the static array holds %zu "
                  "(use srb_psnr_ssim with caller scratch)", B, H, W, need, kPsnrSsimStaticBytes);
        return SRB_E_WORKSPACE;
    }
    static std::atomic<void*> addr[64];
    int dev = 0;
    SRB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("psnr_ssim_noalloc: device index out of range"); return SRB_E_ARG; }
    void* ptr = addr[dev].load(std::memory_order_acquire);
    if (!ptr) {
        SRB_CUDA(cudaGetSymbolAddress(&ptr, g_psnr_ssim_scratch));
        addr[dev].store(ptr, std::memory_order_release);
    }
    return psnr_ssim(pred, target, B, H, W, out2, (float*)ptr, s);
}

// ---------------------------------------------------------------------------------------------------------------
// LOG normalisation: u16( clip((log1p(max(nan_to_num(x),0)) - gmin) / denom, 0, 1) * 65535 )   (truncating cast)
// denom = float32(gmax - gmin + 1e-8) is formed by the caller exactly as numpy does.
// ---------------------------------------------------------------------------------------------------------------
// x -> nan_to_num -> max(.,0): fmaxf(NaN, 0) = 0 and -inf -> 0 already; +inf -> FLT_MAX as np.nan_to_num does.
// The IEEE quotient (d - gmin) / denom is formed without the divide slow path: with rcp = RN(1/denom),
// q0 = RN(a*rcp), r = a - q0*denom (exact in one FMA), q = RN(q0 + r*rcp) is the correctly rounded a/denom
// (Markstein's theorem; no overflow/underflow at these magnitudes), so the result equals numpy's division bit for bit.
__device__ __forceinline__ uint32_t lognorm1(float x, float gmin, float denom, float rcp) {
    x = fminf(fmaxf(x, 0.f), FLT_MAX);
    const float a = log1pf(x) - gmin;
    const float q0 = a * rcp;
    const float r = fmaf(-q0, denom, a);
    float d = fmaf(r, rcp, q0);
    d = fminf(fmaxf(d, 0.f), 1.f);
    return (uint32_t)(d * 65535.f);
}

__global__ void __launch_bounds__(256)
log_normalize_kernel(const float* __restrict__ x, uint16_t* __restrict__ out, size_t n, float gmin, float denom,
                     float rcp) {
    const size_t n8 = n / 8;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = __ldcs((const float4*)x + 2 * i), b = __ldcs((const float4*)x + 2 * i + 1);
        uint4 o;
        o.x = lognorm1(a.x, gmin, denom, rcp) | (lognorm1(a.y, gmin, denom, rcp) << 16);
        o.y = lognorm1(a.z, gmin, denom, rcp) | (lognorm1(a.w, gmin, denom, rcp) << 16);
        o.z = lognorm1(b.x, gmin, denom, rcp) | (lognorm1(b.y, gmin, denom, rcp) << 16);
        o.w = lognorm1(b.z, gmin, denom, rcp) | (lognorm1(b.w, gmin, denom, rcp) << 16);
        __stcs((uint4*)out + i, o);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
        const size_t i = n8 * 8 + threadIdx.x;
        out[i] = (uint16_t)lognorm1(x[i], gmin, denom, rcp);
    }
}

int log_normalize_u16(const float* x, uint16_t* out, size_t n, float gmin, float denom, cudaStream_t s) {
    if (((uintptr_t)x & 15) || ((uintptr_t)out & 15)) { set_error("log_normalize: pointers must be 16-byte aligned"); return SRB_E_ARG; }
    int grid = (int)((n / 8 + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    if (grid < 1) grid = 1;
    const float rcp = (float)(1.0 / (double)denom);  // correctly rounded reciprocal
    log_normalize_kernel<<<grid, 256, 0, s>>>(x, out, n, gmin, denom, rcp);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
