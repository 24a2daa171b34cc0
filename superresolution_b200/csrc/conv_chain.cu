// A chain of dependent 3x3 convolutions (one RRDB = 15 dense-block convs) as ONE persistent
// tcgen05 kernel.  Replaces the per-conv launches of RRDBNet.body (upstream basicsr/archs/rrdbnet_arch.py, call site
// src/architecture.py:75): at B = 16 a 64x64 trunk is only 128 tiles, so a single conv is 1-3 us of tensor work
// behind ~20 us of launch / fill / drain.  Here every CTA keeps its spatial tile for the whole chain and the layers
// are separated by a software grid barrier (release/acquire counter in global memory) instead of a kernel boundary:
//   * per layer the kernel is the patch-resident implicit GEMM of conv_tc.cu (one 4-D TMA halo load per 64-channel
//     chunk, nine taps = nine descriptor views, TMEM accumulators, fused epilogue, TMA store into the channel slice)
//   * the weight producer is not ordered by the barrier: it streams the next layer's weights while the grid waits
//   * only the activation patch loads cross CTAs (1-pixel halo); residual reads hit the CTA's own earlier writes
// All CTAs are co-resident (grid <= SM count, 1 CTA / SM), which the barrier requires.
#include <cuda.h>
#include <cstring>

#include "common.h"
#include "conv_epi.cuh"
#include "ptx.cuh"
#include "tmap.h"

namespace srb {

namespace {

constexpr int kMaxChain = 15;
constexpr int MT = 4, TH = 16, TW = 8 * MT, PH = TH + 2, PW = TW + 2;
constexpr int PATCH_BYTES = PH * PW * 128;
constexpr int PATCH_STRIDE = (PATCH_BYTES + 1023) / 1024 * 1024;
constexpr int W_SLOT = 64 * 128;  // one [N <= 64][64] bf16 weight tile
constexpr int W_RING = 4;
constexpr int PS = 2;             // patch ring stages
constexpr int ACC_STRIDE = MT * 64;
constexpr int kChainThreads = 384;
constexpr int kStage = 8192;      // one 128 x 32 bf16 block, 64B-swizzled
constexpr int OFF_W = PS * PATCH_STRIDE;
constexpr int OFF_STG = OFF_W + W_RING * W_SLOT;
constexpr int OFF_BAR = OFF_STG + 4 * kStage;
constexpr int OFF_BIAS = OFF_BAR + 256;  // [kMaxChain][64] fp32: every layer's bias, staged once
constexpr int CHAIN_SMEM = OFF_BIAS + kMaxChain * 64 * 4 + 1024;
static_assert(CHAIN_SMEM <= 227 * 1024, "shared memory budget");

struct alignas(64) ChainLayer {
    CUtensorMap tmA, tmW, tmO;
    ConvArgs a;
};

struct ChainParams {
    ChainLayer L[kMaxChain];
    int nlayers;
    int dep_chunk[kMaxChain];  // first 64-channel input chunk of layer l that holds outputs of layer l - 1 (see the producer)
    int tiles_x, tiles_y, n_spatial;
    unsigned int* counters;  // [nlayers] zeroed before the launch
    long long* dbg;          // optional [grid][16] cycle totals per role / phase (development aid)
};

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

__global__ void __launch_bounds__(kChainThreads, 1)
conv_chain_kernel(const __grid_constant__ ChainParams P) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* patch0 = smem;
    uint8_t* wt0 = smem + OFF_W;
    uint8_t* stg0 = smem + OFF_STG;
    uint64_t* bars = (uint64_t*)(smem + OFF_BAR);
    uint64_t* full_p = bars;              // [PS]
    uint64_t* empty_p = full_p + PS;      // [PS]
    uint64_t* full_w = empty_p + PS;      // [W_RING]
    uint64_t* empty_w = full_w + W_RING;  // [W_RING]
    uint64_t* acc_full = empty_w + W_RING;  // [2]
    uint64_t* acc_empty = acc_full + 2;     // [2]
    uint32_t* tmem_slot = (uint32_t*)(acc_empty + 2);
    float* sbias = (float*)(smem + OFF_BIAS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    for (int i = threadIdx.x; i < P.nlayers * 64; i += kChainThreads) {  // biases are constants: no dependency wait
        const ConvArgs& la = P.L[i >> 6].a;
        sbias[i] = (la.bias && (i & 63) < la.Cout) ? __ldg(la.bias + (i & 63)) : 0.f;
    }

    if (threadIdx.x == 0) {
        for (int i = 0; i < PS; ++i) { mbar_init(&full_p[i], 1); mbar_init(&empty_p[i], 1); }
        for (int i = 0; i < W_RING; ++i) { mbar_init(&full_w[i], 1); mbar_init(&empty_w[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 256); }
        fence_barrier_init();
    }
    if (warp == 3) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) pdl_launch_dependents();

    const int per_img = P.tiles_x * P.tiles_y;
    const unsigned int ncta = gridDim.x;

    if (warp == 0) {
        // ===================== patch TMA producer: the only role ordered by the grid barrier =====================
        pdl_wait();
        uint32_t it = 0;
        long long dbg_bar = 0;
        const long long tk0 = clock64();
        for (int l = 0; l < P.nlayers; ++l) {
            const ChainLayer& Ly = P.L[l];
            if (lane == 0) tma_prefetch_desc(&Ly.tmA);
            // Dense-block inputs are [x | x1 | ... | x_{k-1}]: only the chunk that holds the previous layer's 32 new channels
            // (and every chunk of an RDB's conv1) depends on the grid barrier.  The older chunks were made visible by
            // earlier barriers, so their patch loads - and the MMAs that consume them - run ahead of the barrier, under the
            // previous layer's epilogue: 60 % of an RDB's MMA work leaves the serial load -> MMA -> epilogue -> barrier chain.
            const int dep = (l > 0) ? P.dep_chunk[l] : 0;
            bool synced = (l == 0);
            const int nchunks = (Ly.a.Cin + 63) / 64;
            for (int st = blockIdx.x; st < P.n_spatial; st += gridDim.x) {
                const int b = st / per_img, r = st % per_img;
                const int ty = r / P.tiles_x, tx = r % P.tiles_x;
                for (int ch = 0; ch < nchunks; ++ch, ++it) {
                    if (!synced && ch >= dep) {
                        if (lane == 0) {
                            const long long tb0 = clock64();
                            uint32_t spins = 0;
                            while (ld_acquire_gpu(&P.counters[l - 1]) < ncta) {
                                if (++spins > (1u << 24)) __trap();  // a protocol bug fails the launch instead of hanging
                            }
                            fence_proxy_async_global();  // order the acquire before this thread's async-proxy (TMA) reads
                            dbg_bar += clock64() - tb0;
                        }
                        __syncwarp();
                        synced = true;
                    }
                    const uint32_t s = it % PS;
                    mbar_wait(&empty_p[s], ((it / PS) & 1) ^ 1);
                    if (lane == 0) {
                        mbar_arrive_expect_tx(&full_p[s], PATCH_BYTES);
                        tma_load_4d(patch0 + s * PATCH_STRIDE, &Ly.tmA, &full_p[s], ch * 64, tx * TW - 1, ty * TH - 1, b);
                    }
                    __syncwarp();
                }
            }
        }
        if (P.dbg && lane == 0) { P.dbg[blockIdx.x * 16 + 0] = dbg_bar; P.dbg[blockIdx.x * 16 + 1] = clock64() - tk0; }
    } else if (warp == 2) {
        // ===================== weight TMA producer (constants: runs ahead of the grid barrier) =====================
        uint32_t it = 0;
        for (int l = 0; l < P.nlayers; ++l) {
            const ChainLayer& Ly = P.L[l];
            if (lane == 0) { tma_prefetch_desc(&Ly.tmW); tma_prefetch_desc(&Ly.tmO); }
            const int nchunks = (Ly.a.Cin + 63) / 64;
            const uint32_t wbytes = (uint32_t)Ly.a.Cout * 128;
            for (int st = blockIdx.x; st < P.n_spatial; st += gridDim.x)
                for (int ch = 0; ch < nchunks; ++ch)
                    for (int tap = 0; tap < 9; ++tap, ++it) {
                        const uint32_t s = it % W_RING;
                        mbar_wait(&empty_w[s], ((it / W_RING) & 1) ^ 1);
                        if (lane == 0) {
                            mbar_arrive_expect_tx(&full_w[s], wbytes);
                            tma_load_3d(wt0 + s * W_SLOT, &Ly.tmW, &full_w[s], ch * 64, 0, tap);
                        }
                        __syncwarp();
                    }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const uint32_t p_u32 = smem_u32(patch0), w_u32 = smem_u32(wt0);
        uint32_t it_p = 0, it_w = 0, it_a = 0;
        long long d_wp = 0, d_ww = 0, d_wa = 0;
        const long long tm0 = clock64();
        for (int l = 0; l < P.nlayers; ++l) {
            const ChainLayer& Ly = P.L[l];
            const int Cin = Ly.a.Cin, nchunks = (Cin + 63) / 64;
            const uint32_t idesc = umma_idesc_bf16_f32(128, (uint32_t)Ly.a.Cout);
            const uint32_t ncol = (uint32_t)Ly.a.Cout;  // accumulator columns per m-tile
            for (int st = blockIdx.x; st < P.n_spatial; st += gridDim.x, ++it_a) {
                const uint32_t as = it_a & 1;
                long long tq = clock64();
                mbar_wait(&acc_empty[as], ((it_a >> 1) & 1) ^ 1);
                d_wa += clock64() - tq;
                tc_fence_after();
                const uint32_t d_base = tmem_base + as * ACC_STRIDE;
                for (int ch = 0; ch < nchunks; ++ch, ++it_p) {
                    const uint32_t ps = it_p % PS;
                    tq = clock64();
                    mbar_wait(&full_p[ps], (it_p / PS) & 1);
                    d_wp += clock64() - tq;
                    tc_fence_after();
                    const uint64_t ad0 = umma_smem_desc(p_u32 + ps * PATCH_STRIDE, 16, PW * 128, UMMA_SW_128B);
                    const bool full_chunk = (Cin - ch * 64) >= 64;  // else 32 valid channels (2 k-steps)
                    const uint32_t acc0 = (ch != 0);
#pragma unroll
                    for (int tap = 0; tap < 9; ++tap, ++it_w) {
                        const uint32_t ws = it_w % W_RING;
                        tq = clock64();
                        mbar_wait(&full_w[ws], (it_w / W_RING) & 1);
                        d_ww += clock64() - tq;
                        tc_fence_after();
                        const uint64_t bd0 = umma_smem_desc(w_u32 + ws * W_SLOT, 16, 1024, UMMA_SW_128B);
                        constexpr int kRowStep = 128 >> 4;
                        const int arow = (tap / 3) * PW + (tap % 3);
                        const uint64_t ad_tap = ad0 + (uint64_t)(arow * kRowStep);
                        const bool leader = elect_one();
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) {
                                if (leader && (ks < 2 || full_chunk))
                                    umma_bf16_ss(d_base + mt * ncol, ad_tap + (uint64_t)(mt * 8 * kRowStep + ks * 2),
                                                 bd0 + (uint64_t)(ks * 2), idesc, (tap == 0 && ks == 0) ? acc0 : 1u);
                            }
                        }
                        if (leader) umma_commit(&empty_w[ws]);
                        __syncwarp();
                    }
                    if (elect_one()) umma_commit(&empty_p[ps]);
                    __syncwarp();
                }
                if (elect_one()) umma_commit(&acc_full[as]);
                __syncwarp();
            }
        }
        if (P.dbg && lane == 0) {
            P.dbg[blockIdx.x * 16 + 2] = d_wp; P.dbg[blockIdx.x * 16 + 3] = d_ww; P.dbg[blockIdx.x * 16 + 4] = d_wa;
            P.dbg[blockIdx.x * 16 + 5] = clock64() - tm0;
        }
    } else if (warp >= 4) {
        // ===================== epilogue: 2 groups x 4 warps =====================
        pdl_wait();
        long long d_acc = 0, d_epi = 0, d_fin = 0;
        long long te = clock64();
        const int eg = (warp - 4) >> 2;
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const bool issuer = (q == 0 && lane == 0);
        uint8_t* stg = stg0 + eg * 2 * kStage;
        uint32_t it_a = 0, it_s = 0;
        for (int l = 0; l < P.nlayers; ++l) {
            const ChainLayer& Ly = P.L[l];
            const ConvArgs& a = Ly.a;
            const int nsub = a.Cout / 32;       // 32-channel blocks per m-tile
            const int units = MT * nsub;
            for (int st = blockIdx.x; st < P.n_spatial; st += gridDim.x, ++it_a) {
                const int b = st / per_img, r = st % per_img;
                const int ty = r / P.tiles_x, tx = r % P.tiles_x;
                const uint32_t as = it_a & 1;
                mbar_wait(&acc_full[as], (it_a >> 1) & 1);
                tc_fence_after();
                { const long long tn = clock64(); d_acc += tn - te; te = tn; }
#pragma unroll 1
                for (int u = eg; u < units; u += 2) {
                    const int mt = u / nsub, c0 = (u % nsub) * 32;
                    const int y = ty * TH + (row >> 3);
                    const int x = tx * TW + 8 * mt + (row & 7);
                    const size_t pix = ((size_t)b * a.H + y) * a.W + x;
                    uint32_t acc[32];
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + as * ACC_STRIDE + mt * a.Cout + c0, acc);
                    tmem_ld_wait();
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
                    epilogue_math(v, a, pix, c0, b, sbias + l * 64);
                    if (a.out32) {
                        float4* op = (float4*)(a.out32 + pix * a.out32_pitch + c0);
#pragma unroll
                        for (int j = 0; j < 8; ++j) op[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    uint8_t* sb = stg + (it_s & 1) * kStage;
                    if (issuer) bulk_wait_group_read<1>();
                    named_bar_sync(1 + eg, 128);
                    const uint32_t sw = (row >> 1) & 3;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint4 o;
                        __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
                        for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[8 * j + 2 * k], v[8 * j + 2 * k + 1]);
                        *(uint4*)(sb + row * 64 + ((j ^ sw) << 4)) = o;
                    }
                    fence_proxy_async_smem();
                    named_bar_sync(1 + eg, 128);
                    if (issuer) {
                        tma_store_4d(&Ly.tmO, sb, c0, tx * TW + 8 * mt, ty * TH, b);
                        bulk_commit_group();
                    }
                    ++it_s;
                }
                tc_fence_before();
                mbar_arrive(&acc_empty[as]);
                { const long long tn = clock64(); d_epi += tn - te; te = tn; }
            }
            if (l + 1 < P.nlayers) {
                // this CTA's layer-l outputs are complete and visible -> arrive on the grid barrier
                if (issuer) { bulk_wait_group<0>(); fence_proxy_async_global(); }
                __threadfence();
                named_bar_sync(3, 256);
                if (warp == 4 && lane == 0) red_release_gpu_add(&P.counters[l], 1u);
                { const long long tn = clock64(); d_fin += tn - te; te = tn; }
            }
        }
        if (issuer) bulk_wait_group<0>();
        if (P.dbg && warp == 4 && lane == 0) {
            P.dbg[blockIdx.x * 16 + 6] = d_acc; P.dbg[blockIdx.x * 16 + 7] = d_epi; P.dbg[blockIdx.x * 16 + 8] = d_fin;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 3) tmem_dealloc(tmem_base, 512);
}

}  // namespace

extern long long* g_fa_dbg;  // attn.cu (srb_debug_set_phase_buffer)

// layers[i]: taps = 9, Cout in {32, 64}, bf16 output through `out` (16-byte aligned, pitch % 8 == 0); all layers share
// B, H, W (H % 16 == 0, W % 32 == 0).  counters: device buffer of >= n uint32 that the caller has zeroed (stream-ordered)
// and that no other launch in flight uses.
int conv_chain_launch(const ConvArgs* layers, int n, unsigned int* counters, cudaStream_t stream) {
    if (n <= 0 || n > kMaxChain) { set_error("conv_chain: 1..%d layers (got %d)", kMaxChain, n); return SRB_E_ARG; }
    const int B = layers[0].B, H = layers[0].H, W = layers[0].W;
    if (H % TH || W % TW) { set_error("conv_chain: H=%d W=%d not divisible by tile %dx%d", H, W, TH, TW); return SRB_E_SHAPE; }
    ChainParams P;
    memset(&P, 0, sizeof(P));
    P.nlayers = n;
    P.tiles_x = W / TW; P.tiles_y = H / TH; P.n_spatial = B * P.tiles_x * P.tiles_y;
    P.counters = counters;
    P.dbg = g_fa_dbg;
    for (int i = 0; i < n; ++i) {
        const ConvArgs& a = layers[i];
        if (a.taps != 9 || (a.Cout != 32 && a.Cout != 64) || a.B != B || a.H != H || a.W != W || a.Cin % 32 || !a.out ||
            a.pixel_shuffle || ((uintptr_t)a.out & 15) || a.out_pitch % 8 || a.in_pitch % 8 || ((uintptr_t)a.in & 15)) {
            set_error("conv_chain: layer %d not chainable (Cin=%d Cout=%d taps=%d)", i, a.Cin, a.Cout, a.taps);
            return SRB_E_ARG;
        }
        ChainLayer& L = P.L[i];
        L.a = a;
        // first input chunk that overlaps what layer i - 1 wrote (bf16 `out` slice or, conservatively, everything when the
        // layers do not share a channel pitch); chunks before it may be loaded ahead of the grid barrier
        P.dep_chunk[i] = 0;
        if (i > 0) {
            const ConvArgs& pv = layers[i - 1];
            const int nchunks = (a.Cin + 63) / 64;
            int dep = nchunks;  // no overlap at all: still synchronise before the last chunk (keeps the layers ordered)
            if (pv.out_pitch != a.in_pitch) dep = 0;
            else {
                const ptrdiff_t d = pv.out - a.in;  // elements; same buffer <=> |d| < pitch
                for (int c = nchunks - 1; c >= 0; --c) {
                    const ptrdiff_t lo = 64 * c, hi = (64 * c + 64 < a.Cin) ? 64 * c + 64 : a.Cin;
                    if (d < hi && d + pv.Cout > lo && d > -(ptrdiff_t)a.in_pitch && d < (ptrdiff_t)a.in_pitch) dep = c;
                }
            }
            P.dep_chunk[i] = dep < nchunks ? dep : nchunks - 1;
        }
        {
            cuuint64_t dims[4] = {(cuuint64_t)a.Cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
            cuuint64_t str[3] = {(cuuint64_t)a.in_pitch * 2, (cuuint64_t)W * a.in_pitch * 2, (cuuint64_t)H * W * a.in_pitch * 2};
            cuuint32_t box[4] = {64, (cuuint32_t)PW, (cuuint32_t)PH, 1};
            SRB_TRY(encode_bf16_map(&L.tmA, a.in, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B));
        }
        {
            cuuint64_t dims[3] = {(cuuint64_t)a.Cin, (cuuint64_t)a.Cout, 9};
            cuuint64_t str[2] = {(cuuint64_t)a.Cin * 2, (cuuint64_t)a.Cout * a.Cin * 2};
            cuuint32_t box[3] = {64, (cuuint32_t)a.Cout, 1};
            SRB_TRY(encode_bf16_map(&L.tmW, a.w, 3, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B));
        }
        {
            cuuint64_t dims[4] = {(cuuint64_t)a.Cout, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
            cuuint64_t str[3] = {(cuuint64_t)a.out_pitch * 2, (cuuint64_t)W * a.out_pitch * 2, (cuuint64_t)H * W * a.out_pitch * 2};
            cuuint32_t box[4] = {32, 8, 16, 1};
            SRB_TRY(encode_bf16_map(&L.tmO, a.out, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_64B));
        }
    }
    static std::atomic<unsigned long long> attr{0};
    SRB_TRY(opt_in_dynamic_smem(conv_chain_kernel, CHAIN_SMEM, attr));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    const int sms = num_sms();
    cfg.gridDim = dim3(P.n_spatial < sms ? P.n_spatial : sms);
    cfg.blockDim = dim3(kChainThreads);
    cfg.dynamicSmemBytes = CHAIN_SMEM;
    cfg.stream = stream;
    static const bool use_pdl = env_flag("SRB_PDL", true);
    // The software grid barrier needs every CTA co-resident.  grid <= SM count and 1 CTA / SM make that true on a whole
    // GPU; under MPS / green-context SM limits or MIG it may not be, so the launch first asks the occupancy calculator
    // (which honours those limits through the device's reported SM count) and refuses instead of spinning into the trap.
    {
        static std::atomic<int> cached[64];  // per device, 0 = not asked yet
        int dev = 0;
        SRB_CUDA(cudaGetDevice(&dev));
        int per_sm = cached[dev & 63].load(std::memory_order_relaxed);
        if (per_sm == 0) {
            SRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, conv_chain_kernel, kChainThreads, CHAIN_SMEM));
            cached[dev & 63].store(per_sm > 0 ? per_sm : -1, std::memory_order_relaxed);
        }
        if (per_sm < 1 || (int)cfg.gridDim.x > per_sm * sms) {
            set_error("conv_chain: %d CTAs cannot be co-resident (%d per SM x %d SMs)", (int)cfg.gridDim.x, per_sm, sms);
            return SRB_E_SHAPE;
        }
    }
    cudaLaunchAttribute attrs[1];
    attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = use_pdl ? 1 : 0;
    SRB_CUDA(cudaLaunchKernelEx(&cfg, conv_chain_kernel, P));
    return SRB_OK;
}

int conv_chain_max_layers() { return kMaxChain; }

}  // namespace srb
