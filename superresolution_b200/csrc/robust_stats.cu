// Robust-statistics pass of Dataset_step4 (reference scripts/Dataset_step4_normalization.py:55-91), SURVEY.md §8(f) rank 1:
// the black / white points are percentiles of log1p(max(x, 0)) over the valid pixels (log1p(..) > 1e-5).  The reference
// estimates them from an unseeded random subsample; here they are the EXACT order statistics of the whole valid
// population, found by a 3-level radix select on the raw float bit pattern (log1p is monotone, so the k-th smallest
// log1p(x) is log1p of the k-th smallest x; valid pixels are positive finite floats >= a host-derived threshold, whose
// bit patterns order like unsigned integers).  Each level is one HBM-bound read of the data (4 B / pixel):
//   level 0: histogram of bits[30:21]            (1024 bins)
//   level 1: histogram of bits[20:10] where bits[31:21] == prefix  (2048 bins, up to 4 prefixes at once)
//   level 2: histogram of bits[9:0]   where bits[31:10] == prefix  (1024 bins, up to 4 prefixes)
// Per-block shared-memory histograms (native u32 atomics); a warp whose 32 pixels fall into one bin (the heavily populated
// sky background) issues a single add of 32; block partials are added to the 64-bit global bins at the end.
#include "common.h"
#include "ops.h"

namespace srb {

constexpr int RS_BINS = 2048;
constexpr int RS_MAXP = 4;

struct RsParams {
    uint32_t thresh_bits;  // smallest valid bit pattern (positive float)
    int bin_shift;         // 21, 10, 0
    uint32_t bin_mask;     // 0x3FF, 0x7FF, 0x3FF
    int prefix_shift;      // 32 (no prefix test), 21, 10
    int nprefix;
    uint32_t prefix[RS_MAXP];
};

__global__ void __launch_bounds__(512)
radix_hist_kernel(const float* __restrict__ x, size_t n, RsParams p, unsigned long long* __restrict__ hist) {
    __shared__ uint32_t sh[RS_MAXP * RS_BINS];
    const int nh = (p.prefix_shift >= 32 ? 1 : p.nprefix) * RS_BINS;
    for (int i = threadIdx.x; i < nh; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const size_t n4 = n / 4;
    auto visit = [&](uint32_t b, bool in_range) {
        // valid: finite, positive, >= threshold (NaN / inf / negatives / zeros are dropped like nan_to_num + max(.,0) + mask)
        bool ok = in_range && b >= p.thresh_bits && b < 0x7F800000u;
        int slot = 0;
        if (ok && p.prefix_shift < 32) {
            const uint32_t pre = b >> p.prefix_shift;
            ok = false;
#pragma unroll
            for (int j = 0; j < RS_MAXP; ++j)
                if (j < p.nprefix && pre == p.prefix[j]) { ok = true; slot = j; break; }
        }
        if (!__any_sync(0xFFFFFFFFu, ok)) return;  // levels 1 / 2: most warps hold no pixel of the wanted prefixes
        const uint32_t key = ok ? (uint32_t)slot * RS_BINS + ((b >> p.bin_shift) & p.bin_mask) : 0xFFFFFFFFu;
        // sky background: the whole warp usually lands in one bin -> one add of 32; otherwise native shared-memory
        // integer atomics (few lanes collide once the keys differ)
        const uint32_t k0 = __shfl_sync(0xFFFFFFFFu, key, 0);
        if (__all_sync(0xFFFFFFFFu, key == k0)) {
            if ((threadIdx.x & 31) == 0) atomicAdd(&sh[key], 32u);
        } else if (ok) {
            atomicAdd(&sh[key], 1u);
        }
    };
    // every thread of a warp runs the same number of iterations (the warp votes need the full warp)
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t iters = (n4 + stride - 1) / stride;
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t it = 0; it < iters; ++it, i += stride) {
        const bool in = i < n4;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (in) v = __ldcs((const uint4*)x + i);
        visit(v.x, in); visit(v.y, in); visit(v.z, in); visit(v.w, in);
    }
    if (blockIdx.x == 0 && threadIdx.x < 32) {  // tail (n % 4 elements), one full warp
        const size_t t = n4 * 4 + threadIdx.x;
        const bool in = t < n;
        visit(in ? __float_as_uint(x[t]) : 0u, in);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < nh; k += blockDim.x)
        if (sh[k]) atomicAdd(&hist[k], (unsigned long long)sh[k]);
}

int robust_hist(const float* x, size_t n, float thresh, int level, const uint32_t* prefixes, int nprefix,
                unsigned long long* hist, cudaStream_t s) {
    if ((uintptr_t)x & 15) { set_error("robust_hist: x must be 16-byte aligned"); return SRB_E_ARG; }
    if (!(thresh > 0.f)) { set_error("robust_hist: threshold must be a positive float"); return SRB_E_ARG; }
    RsParams p;
    memset(&p, 0, sizeof(p));
    memcpy(&p.thresh_bits, &thresh, 4);
    if (level == 0) { p.bin_shift = 21; p.bin_mask = 0x3FF; p.prefix_shift = 32; p.nprefix = 0; }
    else if (level == 1) { p.bin_shift = 10; p.bin_mask = 0x7FF; p.prefix_shift = 21; }
    else if (level == 2) { p.bin_shift = 0; p.bin_mask = 0x3FF; p.prefix_shift = 10; }
    else { set_error("robust_hist: level must be 0, 1 or 2"); return SRB_E_ARG; }
    if (level > 0) {
        if (nprefix < 1 || nprefix > RS_MAXP || !prefixes) { set_error("robust_hist: 1..4 prefixes expected"); return SRB_E_ARG; }
        p.nprefix = nprefix;
        for (int j = 0; j < nprefix; ++j) p.prefix[j] = prefixes[j];
    }
    size_t want = (n / 4 + 511) / 512;
    int grid = (int)(want < 1 ? 1 : want > 148 * 4 ? 148 * 4 : want);
    radix_hist_kernel<<<grid, 512, 0, s>>>(x, n, p, hist);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
