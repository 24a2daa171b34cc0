// Backward building blocks of the conv / linear layers (93 % of the model's FLOPs are
// 3x3 convs and linears, so these two operators carry the same share of the backward pass).  Reference: autograd of
// nn.Conv2d(.., 3, 1, 1) / nn.Linear inside BasicSR RRDBNet and HAT, reached from scripts/Modello_supporto.py:133.
//
//   data gradient    dX = conv3x3(dY, W') with W'[ci][co][ky][kx] = W[co][ci][2-ky][2-kx]  (stride 1, pad 1): the forward
//                    tcgen05 kernel (conv_tc.cu) on a flipped / transposed weight image -> pack_conv_weight_dgrad.
//                    Accumulating into a gradient slice of a dense block is the forward epilogue's `+ res1` in place.
//   weight gradient  dW[co][ci][tap] = sum_px dY[px][co] * X[px + tap][ci]: a GEMM whose contraction runs over pixels, both
//                    operands channel-contiguous (= "MN-major").  conv_wgrad_kernel: split-K over pixel chunks, one CTA
//                    tile = (co block x ci block) x all taps from one haloed X tile, mma.sync.m16n8k16 with
//                    ldmatrix.trans fragments from XOR-swizzled shared memory, fp32 partials + fixed-order
//                    reduction (deterministic, no atomics).
//   bias gradient    db[co] = sum_px dY[px][co]: per-split column sums taken from the dY tiles the weight-gradient kernel
//                    has staged in shared memory anyway, then a fixed-order sum over the splits.
#include <cuda_bf16.h>

#include "common.h"
#include "ops.h"
#include "ptx.cuh"

namespace srb {

namespace {

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// 16-byte async copy; !valid -> the destination is zero-filled (src-size 0), nothing is read
__device__ __forceinline__ void cp_async16_zfill(void* dst, const void* src, bool valid) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)),
                 "l"(src), "r"(valid ? 16 : 0)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}

// fp32 [Cout][Cin][taps] -> bf16 [taps][Cin][Cout] with the taps reversed: the weight image of the data-gradient conv
__global__ void pack_conv_weight_dgrad_kernel(const float* __restrict__ w, bf16* __restrict__ o, int Cout, int Cin,
                                              int taps) {
    const int n = Cout * Cin * taps;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int co = i % Cout, ci = (i / Cout) % Cin, t = i / (Cin * Cout);
        o[i] = __float2bfloat16_rn(w[((size_t)co * Cin + ci) * taps + (taps - 1 - t)]);
    }
}

constexpr int WG_KC = 32;        // pixels per k-chunk (two m16n8k16 k-steps): one row segment of the image
constexpr int WG_THREADS = 256;  // 8 warps, warp tile 32 (co) x 16 (ci)
constexpr int WG_STAGES = 3;

struct WgradArgs {
    const bf16* x;   // [B,H,W,x_pitch]  layer input (channels [0,Cin))
    const bf16* dy;  // [B,H,W,dy_pitch] gradient of the layer output (channels [0,Cout))
    float* partial;  // [ksplit][taps][Cout][Cin]
    float* bpartial; // [ksplit][Cout] column sums of dY per split (bias gradient), or nullptr
    int x_pitch, dy_pitch, B, H, W, Cin, Cout, taps;
    int nchunks, chunks_per_split;
};

// CTA tile: BM = 32 * WM output channels x BN = 8 * (8 / WM) input channels, ALL taps at once: the dY tile of a k-chunk
// is loaded once and the haloed X tile (3 rows x 34 pixels) once, the nine taps are nine shifted ldmatrix views of it
// (per-tap CTAs re-read both operands nine times and were L2-bandwidth bound).  WM = 1 (32 x 64) serves the Cout = 32
// layers (4 of the 5 convs of every dense block) without multiplying zero rows; WM = 2 (64 x 32) everything else.
// Warp tile 32 (co) x 8 (ci) x TAPS -> 72 fp32 accumulators per thread.
template <int WM, int TAPS>
struct WgTile {
    static constexpr int BM = 32 * WM, BN = 8 * (8 / WM);
    static constexpr int UM = BM / 8, UN = BN / 8;  // 16-byte units per pixel row
    static constexpr int HALO = TAPS == 9 ? 1 : 0;
    static constexpr int XR = 1 + 2 * HALO, XP = WG_KC + 2 * HALO;  // rows / pixels of the X tile
    static constexpr int DY_BYTES = WG_KC * BM * 2, X_BYTES = XR * XP * BN * 2;
    static constexpr int STAGE = (DY_BYTES + X_BYTES + 127) / 128 * 128;
};

// 16-byte unit swizzle of a [px][units] tile: any 8 consecutive pixels at one logical unit hit 8 different bank groups
template <int UNITS>
__device__ __forceinline__ int wg_swz(int px, int u) {
    if (UNITS == 4) return u ^ ((px >> 1) & 3);  // 64-byte rows: two rows per 128 bytes
    return u ^ (px & 7);                          // 128-byte rows
}

// One k-chunk = WG_KC consecutive pixels of one image row: dY tile [32 px][BM co], X tile [XR rows][XP px][BN ci]
// (zero-filled outside the image / beyond the channel counts).
template <int WM, int TAPS>
__device__ __forceinline__ void wg_load_chunk(const WgradArgs& a, int chunk, int co0, int ci0, uint8_t* stage, int tid) {
    using T = WgTile<WM, TAPS>;
    const int segs = a.W / WG_KC;
    const int row = chunk / segs, x0 = (chunk % segs) * WG_KC;
    const int b = row / a.H, y = row % a.H;
    for (int i = tid; i < WG_KC * T::UM; i += WG_THREADS) {
        const int px = i / T::UM, u = i % T::UM;
        const bool ok = co0 + u * 8 < a.Cout;
        const bf16* src = a.dy + (((size_t)b * a.H + y) * a.W + x0 + px) * a.dy_pitch + (ok ? co0 + u * 8 : 0);
        cp_async16_zfill(stage + (px * T::UM + wg_swz<T::UM>(px, u)) * 16, src, ok);
    }
    for (int i = tid; i < T::XR * T::XP * T::UN; i += WG_THREADS) {
        const int u = i % T::UN, p = (i / T::UN) % T::XP, r = i / (T::UN * T::XP);
        const int yy = y + r - T::HALO, xx = x0 + p - T::HALO;
        const bool ok = (ci0 + u * 8 < a.Cin) && yy >= 0 && yy < a.H && xx >= 0 && xx < a.W;
        const bf16* src = a.x + (((size_t)b * a.H + (ok ? yy : 0)) * a.W + (ok ? xx : 0)) * a.x_pitch + (ok ? ci0 + u * 8 : 0);
        cp_async16_zfill(stage + T::DY_BYTES + ((r * T::XP + p) * T::UN + wg_swz<T::UN>(p, u)) * 16, src, ok);
    }
}

template <int WM, int TAPS>
__global__ void __launch_bounds__(WG_THREADS, 2)
conv_wgrad_kernel(const WgradArgs a) {
    using T = WgTile<WM, TAPS>;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm = warp / (8 / WM), wn = warp % (8 / WM);  // warp tile: co [32 wm, +32), ci [8 wn, +8)
    const int nci = (a.Cin + T::BN - 1) / T::BN;
    const int co0 = (blockIdx.x / nci) * T::BM, ci0 = (blockIdx.x % nci) * T::BN;
    const int c_begin = blockIdx.y * a.chunks_per_split;
    const int c_end = min(a.nchunks, c_begin + a.chunks_per_split);

    // bias gradient = column sums of dY: the dY tile is already in shared memory, so the CTAs of the first ci tile add
    // their 32 pixels per chunk here (thread = one output channel) instead of a separate pass over dY
    // thread = (16-byte unit of 8 channels, pixel slot): one 16-byte load per chunk, 8 register sums kept over all chunks,
    // folded over the pixel slots once after the loop.
    constexpr int BSLOTS = WG_THREADS / T::UM < WG_KC ? WG_THREADS / T::UM : WG_KC;  // 32 pixel slots (UM 4 or 8)
    const int bu = tid % T::UM, bslot = tid / T::UM;
    const bool do_bias = a.bpartial != nullptr && ci0 == 0 && bslot < BSLOTS;
    float bsum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float acc[TAPS][2][4];
#pragma unroll
    for (int t = 0; t < TAPS; ++t)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[t][j][k] = 0.f;

    // prologue: WG_STAGES - 1 chunks in flight
#pragma unroll
    for (int s = 0; s < WG_STAGES - 1; ++s) {
        if (c_begin + s < c_end) wg_load_chunk<WM, TAPS>(a, c_begin + s, co0, ci0, smem + s * T::STAGE, tid);
        cp_async_commit();
    }
    for (int c = c_begin; c < c_end; ++c) {
        const int st = (c - c_begin) % WG_STAGES;
        cp_async_wait<WG_STAGES - 2>();
        __syncthreads();  // chunk c landed for every thread; the stage refilled below was consumed in iteration c-1
        {
            const int cn = c + WG_STAGES - 1;
            const int sn = (cn - c_begin) % WG_STAGES;
            if (cn < c_end) wg_load_chunk<WM, TAPS>(a, cn, co0, ci0, smem + sn * T::STAGE, tid);
            cp_async_commit();
        }
        if (do_bias) {
            const uint8_t* sd = smem + st * T::STAGE;
#pragma unroll
            for (int px = bslot; px < WG_KC; px += BSLOTS) {
                const uint4 raw = *(const uint4*)(sd + (px * T::UM + wg_swz<T::UM>(px, bu)) * 16);
                const __nv_bfloat162* h2 = (const __nv_bfloat162*)&raw;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = __bfloat1622float2(h2[k]);
                    bsum[2 * k] += f.x;
                    bsum[2 * k + 1] += f.y;
                }
            }
        }
        const uint32_t s_dy = smem_u32(smem + st * T::STAGE), s_x = s_dy + T::DY_BYTES;
        // A = dY^T [co 16][px 16] per (k-step, m-tile): memory is [px][co] -> ldmatrix.trans; matrices ordered
        // (co 0-7, px 0-7), (co 8-15, px 0-7), (co 0-7, px 8-15), (co 8-15, px 8-15) = a0..a3.  Shared by all taps.
        uint32_t af[2][2][4];
#pragma unroll
        for (int ks = 0; ks < 2; ++ks)
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                const int px = ks * 16 + (lane & 7) + ((lane >> 4) << 3);
                const int unit = ((wm * 32 + mt * 16) >> 3) + ((lane >> 3) & 1);
                ldmatrix_x4_trans(s_dy + (px * T::UM + wg_swz<T::UM>(px, unit)) * 16, af[ks][mt][0], af[ks][mt][1],
                                  af[ks][mt][2], af[ks][mt][3]);
            }
#pragma unroll
        for (int t = 0; t < TAPS; ++t) {
            // B = X [px 32][ci 8] of tap t: one x4.trans = {b0, b1} of k-step 0 and {b0, b1} of k-step 1 (lane = pixel)
            const int r = TAPS == 9 ? t / 3 : 0, p = lane + (TAPS == 9 ? t % 3 : 0);
            uint32_t bfr[4];
            ldmatrix_x4_trans(s_x + ((r * T::XP + p) * T::UN + wg_swz<T::UN>(p, wn)) * 16, bfr[0], bfr[1], bfr[2], bfr[3]);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                mma_bf16_16816(acc[t][mt], af[0][mt], bfr[0], bfr[1]);
                mma_bf16_16816(acc[t][mt], af[1][mt], bfr[2], bfr[3]);
            }
        }
    }
    cp_async_wait<0>();
    if (a.bpartial != nullptr && ci0 == 0) {  // fold the pixel slots (CTA-uniform branch): smem [slot][BM] floats
        __syncthreads();
        float* red = (float*)smem;
        if (do_bias) {
#pragma unroll
            for (int k = 0; k < 8; ++k) red[bslot * T::BM + bu * 8 + k] = bsum[k];
        }
        __syncthreads();
        if (tid < T::BM && co0 + tid < a.Cout) {
            float v = 0.f;
            for (int sl = 0; sl < BSLOTS; ++sl) v += red[sl * T::BM + tid];
            a.bpartial[(size_t)blockIdx.y * a.Cout + co0 + tid] = v;
        }
    }
    // partial[split][tap][co][ci]
    const int g = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int t = 0; t < TAPS; ++t) {
        float* dst = a.partial + ((size_t)blockIdx.y * TAPS + t) * a.Cout * a.Cin;
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int co = co0 + wm * 32 + mt * 16 + g + 8 * h;
                const int ci = ci0 + wn * 8 + 2 * tq;
                if (co < a.Cout && ci < a.Cin)
                    *(float2*)(dst + (size_t)co * a.Cin + ci) = make_float2(acc[t][mt][2 * h], acc[t][mt][2 * h + 1]);
            }
    }
}

template <int WM, int TAPS>
int launch_wgrad(const WgradArgs& a, int tiles, int ksplit, cudaStream_t s) {
    using T = WgTile<WM, TAPS>;
    const int smem = WG_STAGES * T::STAGE;
    SRB_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<WM, TAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    conv_wgrad_kernel<WM, TAPS><<<dim3(tiles, ksplit), WG_THREADS, smem, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---- Linear layers (taps == 1) with Cout = 32 MT, Cin = 32 NT: ONE CTA tile covers the whole [Cout x Cin] gradient, so
// every pixel of dY and X is read exactly once (the 64 x 32 tiles above re-read dY Cin/32 times and X Cout/64 times:
// 700 MB of L2 traffic per qkv linear at 4 patches against 200 MB of operands).  8 warps = 2 (co) x 4 (ci), warp tile
// 16 MT x 8 NT, fp32 accumulators in registers (108 for qkv 288 x 96), same split-K partial layout and reduce kernel.
template <int MT, int NT>
struct WgFull {
    static constexpr int UM = 4 * MT, UN = 4 * NT;  // 16-byte units per pixel row of dY / X
    static constexpr int DY_BYTES = WG_KC * UM * 16, X_BYTES = WG_KC * UN * 16;
    static constexpr int STAGE = DY_BYTES + X_BYTES;
};
template <int UNITS>
__device__ __forceinline__ int wgf_swz(int px, int u) {
    if (UNITS % 8 == 0) return u ^ (px & 7);  // row = whole 128-byte lines
    return u ^ ((px >> 1) & 3);               // row = odd number of 64-byte halves: two rows alternate halves
}
template <int MT, int NT>
__device__ __forceinline__ void wgf_load_chunk(const WgradArgs& a, int chunk, uint8_t* stage, int tid) {
    using T = WgFull<MT, NT>;
    const size_t px0 = (size_t)chunk * WG_KC;  // chunks tile the flattened pixel index (W % 32 == 0)
    for (int i = tid; i < WG_KC * T::UM; i += WG_THREADS) {
        const int px = i / T::UM, u = i % T::UM;
        cp_async16_zfill(stage + (px * T::UM + wgf_swz<T::UM>(px, u)) * 16, a.dy + (px0 + px) * a.dy_pitch + u * 8, true);
    }
    for (int i = tid; i < WG_KC * T::UN; i += WG_THREADS) {
        const int px = i / T::UN, u = i % T::UN;
        cp_async16_zfill(stage + T::DY_BYTES + (px * T::UN + wgf_swz<T::UN>(px, u)) * 16, a.x + (px0 + px) * a.x_pitch + u * 8,
                         true);
    }
}

template <int MT, int NT>
__global__ void __launch_bounds__(WG_THREADS, (MT * NT > 18) ? 1 : 2)
linear_wgrad_full_kernel(const WgradArgs a) {
    using T = WgFull<MT, NT>;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm = warp >> 2, wn = warp & 3;  // warp tile: co [16 MT wm, +16 MT), ci [8 NT wn, +8 NT)
    const int c_begin = blockIdx.x * a.chunks_per_split;
    const int c_end = min(a.nchunks, c_begin + a.chunks_per_split);
    // bias gradient: thread = (16-byte unit of 8 channels, pixel slot), 8 register sums over all chunks, folded at the end
    constexpr int BSLOTS = WG_THREADS / T::UM;  // 7 (qkv), 10 or 21 pixel slots
    const int bu = tid % T::UM, bslot = tid / T::UM;
    const bool do_bias = a.bpartial != nullptr && bslot < BSLOTS;
    float bsum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float acc[MT][NT][4];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = acc[i][j][2] = acc[i][j][3] = 0.f;
#pragma unroll
    for (int s = 0; s < WG_STAGES - 1; ++s) {
        if (c_begin + s < c_end) wgf_load_chunk<MT, NT>(a, c_begin + s, smem + s * T::STAGE, tid);
        cp_async_commit();
    }
    for (int c = c_begin; c < c_end; ++c) {
        const int st = (c - c_begin) % WG_STAGES;
        cp_async_wait<WG_STAGES - 2>();
        __syncthreads();
        {
            const int cn = c + WG_STAGES - 1;
            if (cn < c_end) wgf_load_chunk<MT, NT>(a, cn, smem + ((cn - c_begin) % WG_STAGES) * T::STAGE, tid);
            cp_async_commit();
        }
        if (do_bias) {
            const uint8_t* sd = smem + st * T::STAGE;
            for (int px = bslot; px < WG_KC; px += BSLOTS) {
                const uint4 raw = *(const uint4*)(sd + (px * T::UM + wgf_swz<T::UM>(px, bu)) * 16);
                const __nv_bfloat162* h2 = (const __nv_bfloat162*)&raw;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = __bfloat1622float2(h2[k]);
                    bsum[2 * k] += f.x;
                    bsum[2 * k + 1] += f.y;
                }
            }
        }
        const uint32_t s_dy = smem_u32(smem + st * T::STAGE), s_x = s_dy + T::DY_BYTES;
        uint32_t bfr[NT][4];  // {b0, b1} of k-step 0 and of k-step 1 per n-tile (lane = pixel)
#pragma unroll
        for (int j = 0; j < NT; ++j)
            ldmatrix_x4_trans(s_x + (lane * T::UN + wgf_swz<T::UN>(lane, wn * NT + j)) * 16, bfr[j][0], bfr[j][1], bfr[j][2],
                              bfr[j][3]);
#pragma unroll
        for (int i = 0; i < MT; ++i) {
            uint32_t af[2][4];
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                const int px = ks * 16 + (lane & 7) + ((lane >> 4) << 3);
                const int unit = (wm * MT + i) * 2 + ((lane >> 3) & 1);
                ldmatrix_x4_trans(s_dy + (px * T::UM + wgf_swz<T::UM>(px, unit)) * 16, af[ks][0], af[ks][1], af[ks][2], af[ks][3]);
            }
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                mma_bf16_16816(acc[i][j], af[0], bfr[j][0], bfr[j][1]);
                mma_bf16_16816(acc[i][j], af[1], bfr[j][2], bfr[j][3]);
            }
        }
    }
    cp_async_wait<0>();
    if (a.bpartial) {  // fold the pixel slots: smem [slot][Cout] floats (<= 21 x 96 or 7 x 288)
        __syncthreads();
        float* red = (float*)smem;
        if (do_bias) {
#pragma unroll
            for (int k = 0; k < 8; ++k) red[bslot * (32 * MT) + bu * 8 + k] = bsum[k];
        }
        __syncthreads();
        for (int co = tid; co < 32 * MT; co += WG_THREADS) {
            float v = 0.f;
            for (int sl = 0; sl < BSLOTS; ++sl) v += red[sl * (32 * MT) + co];
            a.bpartial[(size_t)blockIdx.x * a.Cout + co] = v;
        }
    }
    const int g = lane >> 2, tq = lane & 3;
    float* dst = a.partial + (size_t)blockIdx.x * a.Cout * a.Cin;
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int co = (wm * MT + i) * 16 + g + 8 * h, ci = (wn * NT + j) * 8 + 2 * tq;
                *(float2*)(dst + (size_t)co * a.Cin + ci) = make_float2(acc[i][j][2 * h], acc[i][j][2 * h + 1]);
            }
}

template <int MT, int NT>
int launch_wgrad_full(const WgradArgs& a, int ksplit, cudaStream_t s) {
    using T = WgFull<MT, NT>;
    const int smem = WG_STAGES * T::STAGE;
    SRB_CUDA(cudaFuncSetAttribute(linear_wgrad_full_kernel<MT, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    linear_wgrad_full_kernel<MT, NT><<<ksplit, WG_THREADS, smem, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// the linear shapes of HAT: qkv 96->288, fc1 96->192, fc2 192->96, proj 96->96
bool wgrad_full_shape(int Cin, int Cout, int taps) {
    return taps == 1 && ((Cout == 288 && Cin == 96) || (Cout == 192 && Cin == 96) || (Cout == 96 && Cin == 192) ||
                         (Cout == 96 && Cin == 96));
}

// dW[co][ci][tap] (PyTorch layout) = (accumulate ? dW : 0) + sum_split partial[split][tap][co][ci];
// db[c] = (accumulate ? db[c] : 0) + sum_split bpartial[split][c].
// Both reductions of a weight-gradient launch in ONE kernel (the step makes ~1000 of these pairs): blocks [0, gw) sum the
// weight partials, 64 outputs x 4 split slices per block (the serial walk over 148 splits by one thread per output was
// latency-bound: 9 us per call), the rest the bias partials.  Fixed order: deterministic.
__global__ void __launch_bounds__(256)
wgrad_finish_kernel(const float* __restrict__ partial, float* __restrict__ dw, const float* __restrict__ bpartial,
                    float* __restrict__ db, int ksplit, int taps, int Cout, int Cin, int accumulate, int gw) {
    __shared__ double sh[4][64];
    const int o = threadIdx.x & 63, sl = threadIdx.x >> 6;
    if ((int)blockIdx.x < gw) {
        const int n = taps * Cout * Cin, i = blockIdx.x * 64 + o;
        float s = 0.f;
        if (i < n)
            for (int k = sl; k < ksplit; k += 4) s += partial[(size_t)k * n + i];
        sh[sl][o] = (double)s;
        __syncthreads();
        if (sl == 0 && i < n) {
            const float tot = (float)sh[0][o] + (float)sh[1][o] + (float)sh[2][o] + (float)sh[3][o];
            const int ci = i % Cin, co = (i / Cin) % Cout, tap = i / (Cin * Cout);
            float* d = dw + ((size_t)co * Cin + ci) * taps + tap;
            *d = accumulate ? *d + tot : tot;
        }
    } else {
        const int c = (blockIdx.x - gw) * 64 + o;
        double s = 0.0;
        if (c < Cout)
            for (int k = sl; k < ksplit; k += 4) s += (double)bpartial[(size_t)k * Cout + c];
        sh[sl][o] = s;
        __syncthreads();
        if (sl == 0 && c < Cout) {
            const float tot = (float)(sh[0][o] + sh[1][o] + sh[2][o] + sh[3][o]);
            db[c] = accumulate ? db[c] + tot : tot;
        }
    }
}

int wgrad_finish(const float* partial, float* dw, const float* bpartial, float* db, int ksplit, int taps, int Cout, int Cin,
                 int accumulate, cudaStream_t s) {
    const int gw = (taps * Cout * Cin + 63) / 64, gb = db ? (Cout + 63) / 64 : 0;
    wgrad_finish_kernel<<<gw + gb, 256, 0, s>>>(partial, dw, bpartial, db, ksplit, taps, Cout, Cin, accumulate, gw);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

constexpr int BG_ROWS = 256;  // (scratch sizing only: the column sums now come from the weight-gradient kernels)

int wgrad_tiles(int Cin, int Cout) {
    const int bm = Cout <= 32 ? 32 : 64, bn = Cout <= 32 ? 64 : 32;
    return ((Cout + bm - 1) / bm) * ((Cin + bn - 1) / bn);
}

int wgrad_ksplit(int Cin, int Cout, int taps, int nchunks) {
    if (wgrad_full_shape(Cin, Cout, taps)) {  // one CTA per split, 1 (qkv: 150 registers) or 2 CTAs per SM
        const int ks = num_sms() * (Cout == 288 ? 1 : 2);
        return ks < nchunks ? ks : (nchunks < 1 ? 1 : nchunks);
    }
    const int tiles = wgrad_tiles(Cin, Cout);
    int ks = (num_sms() * 2 + tiles - 1) / tiles;  // one wave: 2 CTAs per SM are resident (128 registers x 256 threads)
    if (ks > nchunks) ks = nchunks;
    if (ks < 1) ks = 1;
    return ks;
}

}  // namespace

int pack_conv_weight_dgrad(const float* w, bf16* o, int Cout, int Cin, int taps, cudaStream_t s) {
    const int n = Cout * Cin * taps;
    pack_conv_weight_dgrad_kernel<<<(n + 255) / 256, 256, 0, s>>>(w, o, Cout, Cin, taps);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

size_t conv_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout, int taps) {
    if (const int ks = wgrad_tc_splits(B, H, W, Cin, Cout, taps))
        return (size_t)ks * ((size_t)taps * Cout * Cin + Cout) * sizeof(float);
    if (W % WG_KC) return 0;
    const int nchunks = B * H * (W / WG_KC);
    const size_t npx = (size_t)B * H * W;
    const size_t wpart = (size_t)wgrad_ksplit(Cin, Cout, taps, nchunks) * taps * Cout * Cin * sizeof(float);
    const size_t nb = (npx + BG_ROWS - 1) / BG_ROWS;
    const size_t ks = (size_t)wgrad_ksplit(Cin, Cout, taps, nchunks);
    const size_t bpart = (nb > ks ? nb : ks) * Cout * sizeof(float);
    return wpart + bpart;
}

int conv_wgrad(const bf16* x, int x_pitch, const bf16* dy, int dy_pitch, float* dw, float* db, void* scratch, int B,
               int H, int W, int Cin, int Cout, int taps, int accumulate, cudaStream_t s) {
    if (taps != 9 && taps != 1) { set_error("conv_wgrad: taps must be 9 or 1"); return SRB_E_ARG; }
    if (Cin % 8 || Cout % 8 || Cout > 512) { set_error("conv_wgrad: Cin=%d Cout=%d must be multiples of 8 (Cout <= 512)", Cin, Cout); return SRB_E_SHAPE; }
    if (x_pitch % 8 || dy_pitch % 8 || (((uintptr_t)x | (uintptr_t)dy) & 15) || ((uintptr_t)scratch & 15)) {
        set_error("conv_wgrad: pitches / pointers must be 16-byte aligned");
        return SRB_E_ARG;
    }
    if (x_pitch < Cin || dy_pitch < Cout) { set_error("conv_wgrad: pitches (%d, %d) smaller than the channel counts (%d, %d)", x_pitch, dy_pitch, Cin, Cout); return SRB_E_ARG; }
    if (const int ks = wgrad_tc_splits(B, H, W, Cin, Cout, taps)) {
        // tcgen05 path (wgrad_tc.cu): same partial layouts, same fixed-order reductions
        const int n = taps * Cout * Cin;
        float* partial = (float*)scratch;
        float* bpartial = db ? partial + (size_t)ks * n : nullptr;
        SRB_TRY(wgrad_tc_launch(x, x_pitch, dy, dy_pitch, partial, bpartial, ks, B, H, W, Cin, Cout, taps, s));
        return wgrad_finish(partial, dw, bpartial, db, ks, taps, Cout, Cin, accumulate, s);
    }
    if (W % WG_KC) { set_error("conv_wgrad: W=%d must be a multiple of %d", W, WG_KC); return SRB_E_SHAPE; }
    WgradArgs a;
    a.x = x; a.dy = dy; a.partial = (float*)scratch; a.bpartial = nullptr;
    a.x_pitch = x_pitch; a.dy_pitch = dy_pitch; a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.Cout = Cout; a.taps = taps;
    a.nchunks = B * H * (W / WG_KC);
    const int ksplit = wgrad_ksplit(Cin, Cout, taps, a.nchunks);
    a.chunks_per_split = (a.nchunks + ksplit - 1) / ksplit;
    const int tiles = wgrad_tiles(Cin, Cout);
    const int n = taps * Cout * Cin;
    if (db) a.bpartial = a.partial + (size_t)ksplit * n;  // [ksplit][Cout], written by the weight-gradient kernel
    int rc;
    if (wgrad_full_shape(Cin, Cout, taps)) {
        rc = Cout == 288 ? launch_wgrad_full<9, 3>(a, ksplit, s)
           : Cout == 192 ? launch_wgrad_full<6, 3>(a, ksplit, s)
           : Cin == 192  ? launch_wgrad_full<3, 6>(a, ksplit, s)
                         : launch_wgrad_full<3, 3>(a, ksplit, s);
    } else if (taps == 9) rc = Cout <= 32 ? launch_wgrad<1, 9>(a, tiles, ksplit, s) : launch_wgrad<2, 9>(a, tiles, ksplit, s);
    else rc = Cout <= 32 ? launch_wgrad<1, 1>(a, tiles, ksplit, s) : launch_wgrad<2, 1>(a, tiles, ksplit, s);
    if (rc != SRB_OK) return rc;
    return wgrad_finish(a.partial, dw, a.bpartial, db, ksplit, taps, Cout, Cin, accumulate, s);
}

}  // namespace srb
