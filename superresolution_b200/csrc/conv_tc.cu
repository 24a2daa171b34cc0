// Patch-resident implicit-GEMM convolution (3x3 / 1x1) for NHWC bf16 activations on tcgen05 (sm_100a).
//
// Replaces every nn.Conv2d(...,3,1,1) and nn.Linear call site of RRDBNet / HAT (SURVEY.md §2.2 K3-K7, K11, K13-K17;
// reference call sites src/architecture.py:75,78 into the un-vendored BasicSR / HAT modules).
//
// Design (B200-first):
//  * One CTA tile = 16 x (8*MT) output pixels.  The input patch with its 1-pixel halo is fetched ONCE per 64-channel
//    chunk by a single 4-D TMA box load (OOB = zero padding) into 128B-swizzled shared memory: one 128-byte row per
//    pixel.  The nine taps are nine *descriptor views* of that same patch: a UMMA M=128 tile is 16 groups of 8
//    consecutive pixels (one 8-pixel row segment per group) with stride-byte-offset = patch row pitch, and the tap
//    shift is a 128-byte-granular start-address offset (validated on B200, profiles/r01_umma_probe.log).
//    L2->SMEM activation traffic is therefore ~1.2x the tensor instead of 9x.
//  * Weights: 3x3 -> [N_TILE x 64] K-major tiles per (chunk, tap) through a 4-deep TMA ring; 1x1 (linears) -> all
//    tiles resident in shared memory for the whole kernel.  When Cout spans several N tiles the patch stays resident
//    across them (activations are read from L2 once per layer).
//  * Accumulators live in TMEM (MT tiles x N_TILE fp32 columns, double buffered when they fit) so the epilogue of
//    tile i overlaps the MMAs of tile i+1.  Persistent CTAs, warp-specialised (384 threads):
//      warp 0 patch TMA | warp 1 MMA issue | warp 2 weight TMA | warp 3 TMEM alloc | warps 4+ epilogue (2 or 3 groups
//      of 4 warps, each group covering the 128 TMEM lanes and taking every other 128x32 accumulator block).
//  * Epilogue (TMEM -> registers): bias, LeakyReLU / exact GELU, two scaled residuals (fp32 or bf16), then bf16 tiles
//    are staged in 64B-swizzled shared memory and written with TMA stores into a channel slice of a wider NHWC buffer
//    (dense-block concat is free, stores are fully coalesced); fp32 trunk / PixelShuffle(2) outputs use direct stores.
//  * Programmatic dependent launch: barrier init, TMEM allocation, tensor-map prefetch and the weight loads of layer
//    i+1 overlap the tail of layer i (griddepcontrol.wait sits in front of the first activation access).
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "common.h"
#include "ptx.cuh"
#include "conv_epi.cuh"
#include "tmap.h"

namespace srb {

// TRAIN_: the epilogue also honours ConvArgs::pre_out / gelu_grad_of (training path).  A compile-time flag: the two
// extra run-time tests in the inference epilogue cost 0.4 ms per batch-16 forward when they were unconditional.
// KC_: channels per K chunk = bytes / 2 of one patch row (pixel) and one weight row in shared memory: 64 (128-byte rows,
// SWIZZLE_128B) or 32 (64-byte rows, SWIZZLE_64B; for Cin = 32 layers: cab.2).  With KC = 32 the layer's whole 3x3 weight
// image (9 x N_TILE x 64 B) stays RESIDENT in shared memory: re-streaming 108 KB of (zero-padded) weight tiles per 256-px
// tile was cab.2's bottleneck (the MMA warp waited on the weight ring a third of the time).
template <int N_TILE_, int MT_, int TAPS_, bool TRAIN_ = false, int KC_ = 64>
struct ConvCfg {
    static constexpr int N_TILE = N_TILE_, MT = MT_, TAPS = TAPS_, KC = KC_;
    static constexpr bool TRAIN = TRAIN_;
    static constexpr bool W_RESIDENT = (TAPS == 9 && KC == 32);  // all tap tiles of the layer loaded once per CTA
    static constexpr int ROW_BYTES = KC * 2;
    static constexpr uint32_t SWZ = (KC == 64) ? UMMA_SW_128B : UMMA_SW_64B;
    static constexpr int HALO = (TAPS == 9) ? 1 : 0;
    static constexpr int TH = 16, TW = 8 * MT;
    static constexpr int PH = TH + 2 * HALO, PW = TW + 2 * HALO;
    static constexpr int PATCH_BYTES = PH * PW * ROW_BYTES;
    static constexpr int PATCH_STRIDE = (PATCH_BYTES + 1023) / 1024 * 1024;
    static constexpr int W_BYTES = N_TILE * ROW_BYTES;
    static_assert(KC == 64 || KC == 32, "KC");
    static constexpr int W_RING = 4;
    static constexpr int ACC_STAGES = (2 * MT * N_TILE <= 512) ? 2 : 1;
    static constexpr int ACC_COLS = ACC_STAGES * MT * N_TILE;
    static constexpr int TMEM_COLS = ACC_COLS <= 32 ? 32 : ACC_COLS <= 64 ? 64 : ACC_COLS <= 128 ? 128
                                     : ACC_COLS <= 256 ? 256 : 512;
    static constexpr int UNITS = MT * (N_TILE / 32);  // 128x32 accumulator blocks per tile
    // epilogue groups of 4 warps (each group covers the 128 TMEM lanes).  The N = 96 layers (cab.2, the 1x1 linears) are
    // epilogue-bound and have 3 n-blocks per m-tile: three groups; the others keep two (their patches need the smem).
    static constexpr int EG = (N_TILE % 96 == 0) ? 3 : 2;
    static constexpr int THREADS = 128 + EG * 128;
    // output staging: ConvLaunch::nstg blocks of 128 x 32 bf16 per epilogue group (2; 3 when an operand is staged in them)
    static_assert(ACC_COLS <= 512, "accumulators exceed TMEM");
    static_assert(N_TILE % 32 == 0 && N_TILE >= 32 && N_TILE <= 256, "N_TILE");
    static_assert(W_BYTES % 1024 == 0, "weight tile must keep 1024B alignment");
};


constexpr int kStageBytes = 8192;               // one 128 x 32 bf16 block, 64B-swizzled
constexpr int kBarrierBytes = 512;
constexpr int kBiasBytes = 2048;   // the layer's bias in shared memory (Cout <= 512 floats)
constexpr int kColsumBytes = 2 * 16 * 128 * 4;  // [2 tiles][MT * 4 row groups <= 16][N_TILE <= 128] fp32 partial channel sums
constexpr int kSmemLimit = 227 * 1024;
constexpr int kMaxPS = 4;

struct ConvLaunch {
    int tiles_x, tiles_y, n_tiles_n, n_spatial, nchunks;
    int ps;         // patch ring stages
    int wslots;     // weight slots (ring depth for 3x3, resident tile count for 1x1)
    int reuse;      // patch chunks stay resident across the N tiles of one spatial tile
    int tma_store;  // bf16 output through smem staging + TMA store
    int stage_g;    // tiles of one epilogue operand arrive by TMA in the output staging blocks (tmG):
                    // 1 = gelu_grad_of (TRAIN), 2 = res1 (bf16), 0 = none;
                    // 3 (TRAIN) = no operand, but pre_out LEAVES through the staging blocks as a second TMA store (tmG)
    int nstg;       // staging blocks per epilogue group: 2, or 3 when stage_g
    int colsum_bytes;  // shared-memory bytes of the per-tile channel-sum scratch (0 unless a.colsum)
    long long* dbg;    // development aid (-DSRB_CONV_PHASES): [grid][16] cycle totals of the MMA / epilogue warps
};

// Sum of v[c] over the 32 lanes of the warp, returned in lane c (recursive halving: 31 shuffles). Destroys v.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
            const float send = up ? v[i] : v[i + off];
            const float keep = up ? v[i + off] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return v[0];
}

template <class Cfg>
__global__ void __launch_bounds__(Cfg::THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
               const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmG, const ConvArgs a,
               const ConvLaunch t) {
    constexpr int N_TILE = Cfg::N_TILE, MT = Cfg::MT, TAPS = Cfg::TAPS, AS = Cfg::ACC_STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* patch0 = smem;
    uint8_t* wt0 = patch0 + t.ps * Cfg::PATCH_STRIDE;
    uint8_t* stg0 = wt0 + t.wslots * Cfg::W_BYTES;
    const int staging = Cfg::EG * t.nstg * kStageBytes;
    float* scol = (float*)(stg0 + staging);  // only carved when a.colsum (t.colsum_bytes)
    float* sbias = (float*)(stg0 + staging + t.colsum_bytes);
    uint64_t* bars = (uint64_t*)(stg0 + staging + t.colsum_bytes + kBiasBytes);
    uint64_t* full_p = bars;                 // [kMaxPS]
    uint64_t* empty_p = full_p + kMaxPS;     // [kMaxPS]
    uint64_t* full_w = empty_p + kMaxPS;     // [W_RING]  (3x3: ring; 1x1: full_w[0] = all resident tiles landed)
    uint64_t* empty_w = full_w + Cfg::W_RING;  // [W_RING]
    uint64_t* acc_full = empty_w + Cfg::W_RING;  // [2]
    uint64_t* acc_empty = acc_full + 2;          // [2]
    uint32_t* tmem_slot = (uint32_t*)(acc_empty + 2);
    uint64_t* g_bar = acc_empty + 3;             // [EG][3] staged epilogue operand landed (TRAIN)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kMaxPS; ++i) { mbar_init(&full_p[i], 1); mbar_init(&empty_p[i], 1); }
        for (int i = 0; i < Cfg::W_RING; ++i) { mbar_init(&full_w[i], 1); mbar_init(&empty_w[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], Cfg::EG * 128); }
        for (int i = 0; i < Cfg::EG * 3; ++i) mbar_init(&g_bar[i], 1);
        fence_barrier_init();
    }
    if (a.dyn_weights) pdl_wait();  // weights / bias written by the preceding kernel: no prologue overlap
    for (int i = threadIdx.x; i < a.Cout; i += Cfg::THREADS) sbias[i] = a.bias ? __ldg(a.bias + i) : 0.f;  // constants
    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmO); }
    if (warp == 3) { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) pdl_launch_dependents();  // the next layer may begin its own prologue

    const int per_img = t.tiles_x * t.tiles_y;
    const int nreps = t.reuse ? 1 : t.n_tiles_n;  // patch (re)loads per spatial tile

    if (warp == 0) {
        // ===================== patch TMA producer (warp-uniform control flow, one elected lane issues) ==========
        pdl_wait();  // activations come from the previous kernel in the stream
        uint32_t it = 0;
        for (int st = blockIdx.x; st < t.n_spatial; st += gridDim.x) {
            const int b = st / per_img, r = st % per_img;
            const int ty = r / t.tiles_x, tx = r % t.tiles_x;
            for (int rep = 0; rep < nreps; ++rep) {
                for (int ch = 0; ch < t.nchunks; ++ch, ++it) {
                    const uint32_t s = it % t.ps;
                    mbar_wait(&empty_p[s], ((it / t.ps) & 1) ^ 1);
                    if (elect_one()) {
                        mbar_arrive_expect_tx(&full_p[s], Cfg::PATCH_BYTES);
                        tma_load_4d(patch0 + s * Cfg::PATCH_STRIDE, &tmA, &full_p[s], ch * Cfg::KC,
                                    tx * Cfg::TW - Cfg::HALO, ty * Cfg::TH - Cfg::HALO, b);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp == 2) {
        // ===================== weight TMA producer (weights are constants: no dependency wait) =====================
        if (TAPS == 1 || Cfg::W_RESIDENT) {  // every weight tile of the layer, once: slot (nt * nchunks + ch) * TAPS + tap
            if (elect_one()) {
                mbar_arrive_expect_tx(&full_w[0], (uint32_t)(t.wslots * Cfg::W_BYTES));
                for (int nt = 0; nt < t.n_tiles_n; ++nt)
                    for (int ch = 0; ch < t.nchunks; ++ch)
                        for (int tap = 0; tap < TAPS; ++tap)
                            tma_load_3d(wt0 + ((nt * t.nchunks + ch) * TAPS + tap) * Cfg::W_BYTES, &tmW, &full_w[0],
                                        ch * Cfg::KC, nt * N_TILE, tap);
            }
            __syncwarp();
        } else {
            uint32_t it = 0;
            for (int st = blockIdx.x; st < t.n_spatial; st += gridDim.x)
                for (int nt = 0; nt < t.n_tiles_n; ++nt)
                    for (int ch = 0; ch < t.nchunks; ++ch)
                        for (int tap = 0; tap < TAPS; ++tap, ++it) {
                            const uint32_t s = it % Cfg::W_RING;
                            mbar_wait(&empty_w[s], ((it / Cfg::W_RING) & 1) ^ 1);
                            if (elect_one()) {
                                mbar_arrive_expect_tx(&full_w[s], Cfg::W_BYTES);
                                tma_load_3d(wt0 + s * Cfg::W_BYTES, &tmW, &full_w[s], ch * Cfg::KC, nt * N_TILE, tap);
                            }
                            __syncwarp();
                        }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: warp-uniform loops, one elected lane issues =====================
        // Everything that feeds a descriptor is computed in uniform control flow so that it lives in uniform
        // registers; an N=32 MMA is only 16-40 tensor cycles, so the issue path must be a handful of instructions.
        const uint32_t idesc = umma_idesc_bf16_f32(128, N_TILE);
        const uint32_t p_u32 = smem_u32(patch0), w_u32 = smem_u32(wt0);
        uint32_t it_p = 0, it_w = 0, it_a = 0;
#ifdef SRB_CONV_PHASES
        long long ph_acc = 0, ph_patch = 0, ph_w = 0, ph_t0 = clock64(), ph_c;
#define PH_BEGIN() ph_c = clock64()
#define PH_END(x) x += clock64() - ph_c
#else
#define PH_BEGIN()
#define PH_END(x)
#endif
        if (TAPS == 1 || Cfg::W_RESIDENT) { mbar_wait(&full_w[0], 0); tc_fence_after(); }
        for (int st = blockIdx.x; st < t.n_spatial; st += gridDim.x) {
            for (int nt = 0; nt < t.n_tiles_n; ++nt, ++it_a) {
                const uint32_t as = it_a % AS;
                PH_BEGIN();
                mbar_wait(&acc_empty[as], ((it_a / AS) & 1) ^ 1);
                PH_END(ph_acc);
                tc_fence_after();
                const uint32_t d_base = tmem_base + as * (MT * N_TILE);
                for (int ch = 0; ch < t.nchunks; ++ch) {
                    const uint32_t pi = it_p + ch;
                    const uint32_t ps = pi % t.ps;
                    PH_BEGIN();
                    mbar_wait(&full_p[ps], (pi / t.ps) & 1);  // immediate when the chunk is already resident
                    PH_END(ph_patch);
                    tc_fence_after();
                    const uint64_t ad0 = umma_smem_desc(p_u32 + ps * Cfg::PATCH_STRIDE, 16, Cfg::PW * Cfg::ROW_BYTES, Cfg::SWZ);
                    const bool full_chunk = Cfg::KC == 64 && (a.Cin - ch * 64) >= 64;  // else 32 valid channels (2 k-steps)
                    const uint32_t acc0 = (ch != 0);
#pragma unroll
                    for (int tap = 0; tap < TAPS; ++tap) {
                        uint32_t wbase;
                        uint32_t ws = 0;
                        if (TAPS == 1 || Cfg::W_RESIDENT) {
                            wbase = w_u32 + ((nt * t.nchunks + ch) * TAPS + tap) * Cfg::W_BYTES;
                        } else {
                            ws = it_w % Cfg::W_RING;
                            PH_BEGIN();
                            mbar_wait(&full_w[ws], (it_w / Cfg::W_RING) & 1);
                            PH_END(ph_w);
                            tc_fence_after();
                            wbase = w_u32 + ws * Cfg::W_BYTES;
                        }
                        const uint64_t bd0 = umma_smem_desc(wbase, 16, 8 * Cfg::ROW_BYTES, Cfg::SWZ);
                        constexpr int kRowStep = Cfg::ROW_BYTES >> 4;  // one patch row (pixel) in descriptor address units
                        const int arow = (TAPS == 9) ? ((tap / 3) * Cfg::PW + (tap % 3)) : 0;
                        const uint64_t ad_tap = ad0 + (uint64_t)(arow * kRowStep);
                        const bool leader = elect_one();
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) {
                                if (leader && (ks < 2 || full_chunk))
                                    umma_bf16_ss(d_base + mt * N_TILE, ad_tap + (uint64_t)(mt * 8 * kRowStep + ks * 2),
                                                 bd0 + (uint64_t)(ks * 2), idesc, (tap == 0 && ks == 0) ? acc0 : 1u);
                            }
                        }
                        if (TAPS != 1 && !Cfg::W_RESIDENT) {
                            if (leader) umma_commit(&empty_w[ws]);
                            ++it_w;
                        }
                        __syncwarp();
                    }
                    if (!t.reuse || nt == t.n_tiles_n - 1) {
                        if (elect_one()) umma_commit(&empty_p[ps]);
                        __syncwarp();
                    }
                }
                if (elect_one()) umma_commit(&acc_full[as]);
                __syncwarp();
                if (!t.reuse) it_p += t.nchunks;
            }
            if (t.reuse) it_p += t.nchunks;
        }
#ifdef SRB_CONV_PHASES
        if (t.dbg && lane == 0) {
            long long* d = t.dbg + blockIdx.x * 16;
            d[0] = clock64() - ph_t0; d[1] = ph_acc; d[2] = ph_patch; d[3] = ph_w;
        }
#endif
    } else if (warp >= 4) {
        // ===================== epilogue: EG groups x 4 warps, each group = all 128 TMEM lanes =====================
        pdl_wait();  // residual reads / output writes must not overtake the previous kernel
        const int eg = (warp - 4) >> 2;
        const int q = warp & 3;
        const int row = q * 32 + lane;  // M row inside the 128-row UMMA tile
        const bool issuer = (q == 0 && lane == 0);
        uint8_t* stg = stg0 + eg * t.nstg * kStageBytes;
        uint32_t it_a = 0, it_s = 0;
#ifdef SRB_CONV_PHASES
        long long ph_full = 0, ph_e0 = clock64(), ph_c, ph_ld = 0, ph_math = 0, ph_b1 = 0, ph_st = 0, ph_b2 = 0;
#endif
        // TRAIN, GELU' epilogue: the saved pre-activation tile of a unit (128 pixels x 32 channels) is fetched by TMA
        // into the staging block the unit's output will leave from, one unit ahead; each thread then multiplies its
        // 64-byte row in place.  Read by the threads themselves (LDG or cp.async: 16 bytes per lane at a 384-byte
        // stride = one L1 wavefront per lane) those operands cost ~1.4k cycles of LSU time per unit whether or not their
        // latency was hidden; the TMA tile costs none.
        const bool stage_g = t.stage_g == 1 || t.stage_g == 2;
        const bool pre_tma = Cfg::TRAIN && t.stage_g == 3;
        const uint32_t sw_row = (row >> 1) & 3;
        uint64_t* gb = g_bar + eg * 3;
        auto issue_g = [&](int st_, int nt_, int u_, uint32_t buf) {  // issuer thread only
            const int b_ = st_ / per_img, r_ = st_ % per_img;
            const int ty_ = r_ / t.tiles_x, tx_ = r_ % t.tiles_x;
            const int mt_ = u_ / (N_TILE / 32), c0_ = (u_ % (N_TILE / 32)) * 32;
            mbar_arrive_expect_tx(&gb[buf], kStageBytes);
            tma_load_4d(stg + buf * kStageBytes, &tmG, &gb[buf], nt_ * N_TILE + c0_, tx_ * Cfg::TW + 8 * mt_, ty_ * Cfg::TH, b_);
        };
        if (stage_g && issuer && (int)blockIdx.x < t.n_spatial) issue_g(blockIdx.x, 0, eg, 0);
        // stores of one staging block: wait for the store `depth` stores back to have read it, fill, hand to TMA
        auto stage_store = [&](const CUtensorMap* tm, const float (&vv)[32], uint8_t* sb, int cb_, int x_, int y_, int b_) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint4 o;
                __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
                for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(vv[8 * j + 2 * k], vv[8 * j + 2 * k + 1]);
                *(uint4*)(sb + row * 64 + ((j ^ sw_row) << 4)) = o;
            }
            fence_proxy_async_smem();
            named_bar_sync(2 + eg, 128);
            if (issuer) {
                tma_store_4d(tm, sb, cb_, x_, y_, b_);
                bulk_commit_group();
            }
        };
        for (int st = blockIdx.x; st < t.n_spatial; st += gridDim.x) {
            const int b = st / per_img, r = st % per_img;
            const int ty = r / t.tiles_x, tx = r % t.tiles_x;
            for (int nt = 0; nt < t.n_tiles_n; ++nt, ++it_a) {
                const uint32_t as = it_a % AS;
                PH_BEGIN();
                mbar_wait(&acc_full[as], (it_a / AS) & 1);
                PH_END(ph_full);
                tc_fence_after();
#pragma unroll 1
                for (int u = eg; u < Cfg::UNITS; u += Cfg::EG) {
                    const int mt = u / (N_TILE / 32), c0 = (u % (N_TILE / 32)) * 32;
                    const int y = ty * Cfg::TH + (row >> 3);
                    const int x = tx * Cfg::TW + 8 * mt + (row & 7);
                    // ragged last tile row / column (H, W not multiples of the tile): rows beyond the image compute on
                    // zero-filled input, read their residuals from a clamped (valid) pixel and are never stored
                    const bool valid = (y < a.H) && (x < a.W);
                    const size_t pix = ((size_t)b * a.H + (y < a.H ? y : a.H - 1)) * a.W + (x < a.W ? x : a.W - 1);
                    const int cb = nt * N_TILE + c0;  // first output channel of this 32-wide block
                    uint32_t acc[32];
                    PH_BEGIN();
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + as * (MT * N_TILE) + mt * N_TILE + c0,
                                       acc);
                    tmem_ld_wait();
                    PH_END(ph_ld);
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
                    const uint8_t* g_row = nullptr;  // this thread's row of the staged operand
                    if (stage_g) {
                        // three blocks in rotation: the block the next unit's operand goes to was last stored from two
                        // units ago (waiting for the store issued a moment ago would expose its shared-memory read)
                        if (issuer) {
                            bulk_wait_group_read<1>();
                            int st2 = st, nt2 = nt, u2 = u + Cfg::EG;
                            bool more = true;
                            if (u2 >= Cfg::UNITS) {
                                u2 = eg;
                                if (++nt2 >= t.n_tiles_n) { nt2 = 0; st2 += gridDim.x; more = st2 < t.n_spatial; }
                            }
                            if (more) issue_g(st2, nt2, u2, (it_s + 1) % 3);
                        }
                        mbar_wait(&gb[it_s % 3], (it_s / 3) & 1);  // this unit's tile (requested one unit ago) has landed
                        g_row = stg + (it_s % 3) * kStageBytes + row * 64;
                    }
                    if (Cfg::TRAIN && pre_tma) {
                        // the bf16 pre-activation (acc + bias) leaves as a TMA tile of its own: blocks rotate per STORE, the
                        // block of store s was last read by store s - 3
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            const float4 bb = *(const float4*)(sbias + cb + j);
                            v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
                        }
                        if (issuer) bulk_wait_group_read<2>();
                        named_bar_sync(2 + eg, 128);
                        stage_store(&tmG, v, stg + (it_s % 3) * kStageBytes, cb, tx * Cfg::TW + 8 * mt, ty * Cfg::TH, b);
                        ++it_s;
                    }
                    PH_BEGIN();
                    epilogue_math<Cfg::TRAIN>(v, a, pix, cb, b, sbias, (Cfg::TRAIN && t.stage_g == 1) ? g_row : nullptr,
                                              t.stage_g == 2 ? g_row : nullptr, sw_row, Cfg::TRAIN && pre_tma);
                    PH_END(ph_math);
                    if (a.out32 && valid) {
                        float4* op = (float4*)(a.out32 + pix * a.out32_pitch + cb);
#pragma unroll
                        for (int j = 0; j < 8; ++j) op[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    if (a.out) {
                        if (t.tma_store) {
                            // stage the 128 x 32 bf16 block (64-byte rows, SWIZZLE_64B) and hand it to TMA
                            uint8_t* sb = stg + ((stage_g || pre_tma) ? it_s % 3 : (it_s & 1)) * kStageBytes;
                            PH_BEGIN();
                            if (!stage_g) {  // (the staged-operand path has waited above)
                                if (issuer) {  // the store that last used this buffer has read it
                                    if (pre_tma) bulk_wait_group_read<2>(); else bulk_wait_group_read<1>();
                                }
                                named_bar_sync(2 + eg, 128);
                            }
                            PH_END(ph_b1);
                            PH_BEGIN();
                            const uint32_t sw = (row >> 1) & 3;
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                uint4 o;
                                __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
                                for (int k = 0; k < 4; ++k)
                                    h[k] = __floats2bfloat162_rn(v[8 * j + 2 * k], v[8 * j + 2 * k + 1]);
                                *(uint4*)(sb + row * 64 + ((j ^ sw) << 4)) = o;
                            }
                            fence_proxy_async_smem();
                            PH_END(ph_st);
                            PH_BEGIN();
                            named_bar_sync(2 + eg, 128);
                            PH_END(ph_b2);
                            if (issuer) {
                                tma_store_4d(&tmO, sb, cb, tx * Cfg::TW + 8 * mt, ty * Cfg::TH, b);
                                bulk_commit_group();
                            }
                            ++it_s;
                        } else if (!valid) {
                            // nothing to store
                        } else if (!a.pixel_shuffle) {
                            uint4* op = (uint4*)(a.out + pix * a.out_pitch + cb);
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                uint4 o;
                                __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
                                for (int k = 0; k < 4; ++k)
                                    h[k] = __floats2bfloat162_rn(v[8 * j + 2 * k], v[8 * j + 2 * k + 1]);
                                op[j] = o;
                            }
                        } else {
                            // PixelShuffle(2): out[b, 2y+i, 2x+j, c] = conv[b, y, x, 4c + 2i + j]
#pragma unroll
                            for (int ij = 0; ij < 4; ++ij) {
                                const size_t opix = ((size_t)b * 2 * a.H + 2 * y + (ij >> 1)) * (2 * a.W) + 2 * x + (ij & 1);
                                uint4 o;
                                __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
                                for (int k = 0; k < 4; ++k)
                                    h[k] = __floats2bfloat162_rn(v[4 * (2 * k) + ij], v[4 * (2 * k + 1) + ij]);
                                *(uint4*)(a.out + opix * a.out_pitch + (cb >> 2)) = o;
                            }
                        }
                    }
                    if (a.colsum) {  // channel sums of this unit over the warp's 32 pixels (v is dead from here on)
                        if (!valid) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = 0.f;
                        }
                        const float cs_ = warp_colsum32(v, lane);
                        scol[((it_a & 1) * 16 + mt * 4 + q) * N_TILE + c0 + lane] = cs_;
                    }
                }
                tc_fence_before();
                mbar_arrive(&acc_empty[as]);
                if (a.colsum) {  // all epilogue groups: fixed-order sum of the MT*4 row-group partials -> colsum[st][c]
                    named_bar_sync(1, Cfg::EG * 128);
                    const int c = threadIdx.x - 128;
                    if (c < N_TILE) {
                        float tot = 0.f;
#pragma unroll
                        for (int g8 = 0; g8 < MT * 4; ++g8) tot += scol[((it_a & 1) * 16 + g8) * N_TILE + c];
                        a.colsum[(size_t)st * a.Cout + c] = tot;
                    }
                }
            }
        }
        if (issuer) bulk_wait_group<0>();  // all TMA stores of this group have completed
#ifdef SRB_CONV_PHASES
        if (t.dbg && warp == 4 && lane == 0) {
            long long* d = t.dbg + blockIdx.x * 16;
            d[4] = clock64() - ph_e0; d[5] = ph_full; d[6] = ph_ld; d[7] = ph_math; d[8] = ph_b1; d[9] = ph_st; d[10] = ph_b2;
        }
#endif
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 3) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

extern long long* g_fa_dbg;  // attn.cu (srb_debug_set_phase_buffer)
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

static EncodeTiledFn get_encode() {
    std::call_once(g_encode_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            g_encode = (EncodeTiledFn)fn;
    });
    return g_encode;
}

int encode_bf16_map(CUtensorMap* m, const void* ptr, int rank, const cuuint64_t* dims,
                           const cuuint64_t* strides_bytes, const cuuint32_t* box, CUtensorMapSwizzle swz) {
    EncodeTiledFn enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
        return SRB_E_CUDA;
    }
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), dims,
                     strides_bytes, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]",
                  (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1],
                  (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0],
                  box[1], rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
        return SRB_E_CUDA;
    }
    return SRB_OK;
}

int num_sms() {
    static std::atomic<int> cache[64];  // per device (zero-initialised)
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    std::atomic<int>& c = cache[dev & 63];
    n = c.load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        c.store(n, std::memory_order_relaxed);
    }
    return n;
}

bool env_flag(const char* name, bool dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    return v[0] != '0';
}

template <class Cfg>
static int launch_cfg(const ConvArgs& a, cudaStream_t stream) {
    // Inference layers take any H, W: the last tile row / column is ragged (TMA zero-fills the loads and clips the
    // stores, the epilogue masks the rest).  The training epilogue (pre_out) keeps whole tiles.
    if (Cfg::TRAIN && (a.H % Cfg::TH || a.W % Cfg::TW)) {
        set_error("conv_tc(train): H=%d W=%d not divisible by tile %dx%d", a.H, a.W, Cfg::TH, Cfg::TW);
        return SRB_E_SHAPE;
    }
    ConvLaunch t;
    t.tiles_x = (a.W + Cfg::TW - 1) / Cfg::TW;
    t.tiles_y = (a.H + Cfg::TH - 1) / Cfg::TH;
    t.n_tiles_n = a.Cout / Cfg::N_TILE;
    t.n_spatial = a.B * t.tiles_x * t.tiles_y;
    t.nchunks = (a.Cin + Cfg::KC - 1) / Cfg::KC;
    t.wslots = (Cfg::TAPS == 1 || Cfg::W_RESIDENT) ? t.n_tiles_n * t.nchunks * Cfg::TAPS : Cfg::W_RING;
    t.colsum_bytes = a.colsum ? kColsumBytes : 0;
    if (a.colsum && (t.n_tiles_n != 1 || Cfg::MT * 4 > 16 || Cfg::N_TILE > 128)) {
        set_error("conv_tc: colsum needs Cout == N tile (Cout=%d)", a.Cout);
        return SRB_E_ARG;
    }
    static const bool use_tma_store = env_flag("SRB_TMA_STORE", true);
    static const bool use_pdl = env_flag("SRB_PDL", true);
    static const bool use_stage = env_flag("SRB_STAGE_OPERAND", true);
    t.tma_store = (a.out && !a.pixel_shuffle && use_tma_store && (((uintptr_t)a.out & 15) == 0) &&
                   (a.out_pitch % 8 == 0)) ? 1 : 0;
    // one per-pixel epilogue operand can ride the staging blocks (TMA tile in, multiply / add in place, TMA tile out)
    const bf16* st_op = nullptr;
    int st_pitch = 0;
    t.stage_g = 0;
    if (t.tma_store && use_stage) {
        if (Cfg::TRAIN && a.gelu_grad_of) { st_op = a.gelu_grad_of; st_pitch = a.gelu_grad_pitch; t.stage_g = 1; }
        else if (a.res1 && !a.res1_f32) { st_op = (const bf16*)a.res1; st_pitch = a.res1_pitch; t.stage_g = 2; }
        else if (Cfg::TRAIN && a.pre_out) { st_op = a.pre_out; st_pitch = a.pre_pitch; t.stage_g = 3; }
        if (st_op && ((((uintptr_t)st_op & 15) != 0) || st_pitch % 8 != 0)) t.stage_g = 0;
    }
    t.nstg = t.stage_g ? 3 : 2;
    int fixed = t.wslots * Cfg::W_BYTES + Cfg::EG * t.nstg * kStageBytes + t.colsum_bytes + kBiasBytes + kBarrierBytes + 1024;
    if (t.stage_g && (kSmemLimit - fixed) / Cfg::PATCH_STRIDE < (t.nchunks < 2 ? 2 : t.nchunks)) {
        t.stage_g = 0;  // the third block would cost a patch stage the layer needs
        t.nstg = 2;
        fixed -= Cfg::EG * kStageBytes;
    }
    t.ps = (kSmemLimit - fixed) / Cfg::PATCH_STRIDE;
    if (t.ps > kMaxPS) t.ps = kMaxPS;
    if (t.ps < 2) {
        set_error("conv_tc: layer does not fit shared memory (Cin=%d Cout=%d taps=%d)", a.Cin, a.Cout, Cfg::TAPS);
        return SRB_E_SHAPE;
    }
    t.reuse = (t.n_tiles_n > 1 && t.nchunks <= t.ps) ? 1 : 0;
    const int smem = t.ps * Cfg::PATCH_STRIDE + fixed;
    t.dbg = g_fa_dbg;

    CUtensorMap tmA, tmW, tmO;
    {
        cuuint64_t dims[4] = {(cuuint64_t)a.Cin, (cuuint64_t)a.W, (cuuint64_t)a.H, (cuuint64_t)a.B};
        cuuint64_t str[3] = {(cuuint64_t)a.in_pitch * 2, (cuuint64_t)a.W * a.in_pitch * 2,
                             (cuuint64_t)a.H * a.W * a.in_pitch * 2};
        cuuint32_t box[4] = {(cuuint32_t)Cfg::KC, (cuuint32_t)Cfg::PW, (cuuint32_t)Cfg::PH, 1};
        SRB_TRY(encode_bf16_map(&tmA, a.in, 4, dims, str, box, Cfg::KC == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B));
    }
    {
        cuuint64_t dims[3] = {(cuuint64_t)a.Cin, (cuuint64_t)a.Cout, (cuuint64_t)Cfg::TAPS};
        cuuint64_t str[2] = {(cuuint64_t)a.Cin * 2, (cuuint64_t)a.Cout * a.Cin * 2};
        cuuint32_t box[3] = {(cuuint32_t)Cfg::KC, (cuuint32_t)Cfg::N_TILE, 1};
        SRB_TRY(encode_bf16_map(&tmW, a.w, 3, dims, str, box, Cfg::KC == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B));
    }
    if (t.tma_store) {
        cuuint64_t dims[4] = {(cuuint64_t)a.Cout, (cuuint64_t)a.W, (cuuint64_t)a.H, (cuuint64_t)a.B};
        cuuint64_t str[3] = {(cuuint64_t)a.out_pitch * 2, (cuuint64_t)a.W * a.out_pitch * 2,
                             (cuuint64_t)a.H * a.W * a.out_pitch * 2};
        cuuint32_t box[4] = {32, 8, 16, 1};
        SRB_TRY(encode_bf16_map(&tmO, a.out, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_64B));
    } else {
        tmO = tmA;  // unused
    }
    CUtensorMap tmG = tmA;  // unused unless an epilogue operand is staged
    if (t.stage_g) {
        cuuint64_t dims[4] = {(cuuint64_t)a.Cout, (cuuint64_t)a.W, (cuuint64_t)a.H, (cuuint64_t)a.B};
        cuuint64_t str[3] = {(cuuint64_t)st_pitch * 2, (cuuint64_t)a.W * st_pitch * 2, (cuuint64_t)a.H * a.W * st_pitch * 2};
        cuuint32_t box[4] = {32, 8, 16, 1};
        SRB_TRY(encode_bf16_map(&tmG, st_op, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_64B));
    }
    static std::atomic<unsigned long long> smem_set{0};  // per template instantiation, one bit per device
    SRB_TRY(opt_in_dynamic_smem(conv_tc_kernel<Cfg>, kSmemLimit, smem_set));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(t.n_spatial < num_sms() ? t.n_spatial : num_sms());
    cfg.blockDim = dim3(Cfg::THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = use_pdl ? 1 : 0;
    SRB_CUDA(cudaLaunchKernelEx(&cfg, conv_tc_kernel<Cfg>, tmA, tmW, tmO, tmG, a, t));
    return SRB_OK;
}

int conv_tc_launch(const ConvArgs& a_in, cudaStream_t stream) {
    ConvArgs a = a_in;
    if (a.taps != 9 && a.taps != 1) { set_error("conv_tc: taps must be 9 or 1"); return SRB_E_ARG; }
    if (a.Cin % 16 || a.Cin < 16) { set_error("conv_tc: Cin=%d must be a multiple of 16", a.Cin); return SRB_E_SHAPE; }
    if (a.Cout % 32 || a.Cout > kBiasBytes / 4) { set_error("conv_tc: Cout=%d must be a multiple of 32, at most %d", a.Cout, kBiasBytes / 4); return SRB_E_SHAPE; }
    if (a.in_pitch % 8 || ((uintptr_t)a.in & 15) || ((uintptr_t)a.w & 15)) {
        set_error("conv_tc: input pitch / pointers must be 16-byte aligned");
        return SRB_E_ARG;
    }
    if (a.pixel_shuffle && (a.Cout % 128 || !a.out)) { set_error("conv_tc: pixel_shuffle needs Cout%%128==0"); return SRB_E_ARG; }
    if (a.pre_out || a.gelu_grad_of || a.alpha1_samples) {  // training-path epilogue: instantiated for the two shapes that use it
        if (a.taps == 1 && a.Cout % 96 == 0) return launch_cfg<ConvCfg<96, 2, 1, true>>(a, stream);
        if (a.taps == 9 && a.Cout == 32 && a.Cin == 96 && a.W % 32 == 0) return launch_cfg<ConvCfg<32, 4, 9, true, 32>>(a, stream);  // cab.0 / cab.2 data gradient: resident weights
        if (a.taps == 9 && a.Cout == 32 && a.W % 32 == 0) return launch_cfg<ConvCfg<32, 4, 9, true>>(a, stream);
        set_error("conv_tc: pre_out / gelu_grad_of / alpha1_samples need a linear with Cout %% 96 == 0 or a 3x3 conv with Cout == 32, W %% 32 == 0");
        return SRB_E_SHAPE;
    }
    if (a.taps == 9) {
        if (a.Cout == 32 && a.Cin == 96 && a.W % 32 == 0) return launch_cfg<ConvCfg<32, 4, 9, false, 32>>(a, stream);  // cab.0
        if (a.Cout == 32) return (a.W % 32 == 0) ? launch_cfg<ConvCfg<32, 4, 9>>(a, stream) : launch_cfg<ConvCfg<32, 2, 9>>(a, stream);
        if (a.Cout == 64) return (a.W % 32 == 0) ? launch_cfg<ConvCfg<64, 4, 9>>(a, stream) : launch_cfg<ConvCfg<64, 2, 9>>(a, stream);
        if (a.Cout % 128 == 0) return launch_cfg<ConvCfg<128, 2, 9>>(a, stream);
        if (a.Cout % 96 == 0 && a.Cin == 32 && a.Cout == 96) return launch_cfg<ConvCfg<96, 2, 9, false, 32>>(a, stream);  // cab.2
        if (a.Cout % 96 == 0) return launch_cfg<ConvCfg<96, 2, 9>>(a, stream);
        // any other multiple of 32 (e.g. the 160-channel data gradient of a dense-block conv4): N tiles of 32
        return (a.W % 32 == 0) ? launch_cfg<ConvCfg<32, 4, 9>>(a, stream) : launch_cfg<ConvCfg<32, 2, 9>>(a, stream);
    } else {
        if (a.Cout % 96 == 0) return launch_cfg<ConvCfg<96, 2, 1>>(a, stream);
        if (a.Cout % 64 == 0) return launch_cfg<ConvCfg<64, 2, 1>>(a, stream);
    }
    set_error("conv_tc: no kernel instance for Cout=%d taps=%d", a.Cout, a.taps);
    return SRB_E_SHAPE;
}

}  // namespace srb
