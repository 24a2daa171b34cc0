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
//  * Weights stream per (chunk, tap) as [N_TILE x 64] K-major tiles through their own TMA ring.
//  * Accumulators live in TMEM (MT tiles x N_TILE fp32 columns, double buffered when they fit) so the epilogue of
//    tile i overlaps the MMAs of tile i+1.  Persistent CTAs, warp-specialised:
//      warp 0 patch TMA | warp 1 MMA issue | warp 2 weight TMA | warp 3 TMEM alloc | warps 4-7 epilogue.
//  * Epilogue (TMEM -> registers): bias, LeakyReLU / exact GELU, two scaled residuals (fp32 or bf16), bf16 and/or
//    fp32 stores into a channel slice of a wider NHWC buffer (dense-block concat is free), optional PixelShuffle(2).
#include <cuda.h>
#include <mutex>
#include <cstdio>

#include "common.h"
#include "ptx.cuh"

namespace srb {

template <int N_TILE_, int MT_, int TAPS_>
struct ConvCfg {
    static constexpr int N_TILE = N_TILE_, MT = MT_, TAPS = TAPS_;
    static constexpr int HALO = (TAPS == 9) ? 1 : 0;
    static constexpr int TH = 16, TW = 8 * MT;
    static constexpr int PH = TH + 2 * HALO, PW = TW + 2 * HALO;
    static constexpr int PATCH_BYTES = PH * PW * 128;
    static constexpr int PATCH_STRIDE = (PATCH_BYTES + 1023) / 1024 * 1024;
    static constexpr int W_BYTES = N_TILE * 128;
    static constexpr int WS = 4;
    static constexpr int PS_MAX = (200 * 1024 - WS * W_BYTES) / PATCH_STRIDE;
    static constexpr int PS = PS_MAX >= 4 ? 4 : (PS_MAX >= 3 ? 3 : 2);
    static constexpr int ACC_STAGES = (2 * MT * N_TILE <= 512) ? 2 : 1;
    static constexpr int ACC_COLS = ACC_STAGES * MT * N_TILE;
    static constexpr int TMEM_COLS = ACC_COLS <= 32 ? 32 : ACC_COLS <= 64 ? 64 : ACC_COLS <= 128 ? 128
                                     : ACC_COLS <= 256 ? 256 : 512;
    static constexpr int SMEM_BYTES = PS * PATCH_STRIDE + WS * W_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
    static_assert(ACC_COLS <= 512, "accumulators exceed TMEM");
    static_assert(N_TILE % 32 == 0 && N_TILE >= 32 && N_TILE <= 256, "N_TILE");
    static_assert(W_BYTES % 1024 == 0, "weight tile must keep 1024B alignment");
    static_assert(PS >= 2 && SMEM_BYTES <= 227 * 1024, "smem budget");
};

struct ConvTiles {
    int tiles_x, tiles_y, n_tiles_n, total, nchunks;
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }

template <class Cfg>
__global__ void __launch_bounds__(256, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const ConvArgs a,
               const ConvTiles t) {
    constexpr int N_TILE = Cfg::N_TILE, MT = Cfg::MT, TAPS = Cfg::TAPS;
    constexpr int PS = Cfg::PS, WS = Cfg::WS, AS = Cfg::ACC_STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* patch0 = smem;
    uint8_t* wt0 = smem + PS * Cfg::PATCH_STRIDE;
    uint64_t* bars = (uint64_t*)(wt0 + WS * Cfg::W_BYTES);
    uint64_t* full_p = bars;            // [PS]
    uint64_t* empty_p = full_p + PS;    // [PS]
    uint64_t* full_w = empty_p + PS;    // [WS]
    uint64_t* empty_w = full_w + WS;    // [WS]
    uint64_t* acc_full = empty_w + WS;  // [AS]
    uint64_t* acc_empty = acc_full + AS;  // [AS]
    uint32_t* tmem_slot = (uint32_t*)(acc_empty + AS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < PS; ++i) { mbar_init(&full_p[i], 1); mbar_init(&empty_p[i], 1); }
        for (int i = 0; i < WS; ++i) { mbar_init(&full_w[i], 1); mbar_init(&empty_w[i], 1); }
        for (int i = 0; i < AS; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
        fence_barrier_init();
    }
    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW); }
    if (warp == 3) { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int per_img = t.tiles_x * t.tiles_y;

    if (warp == 0) {
        // ===================== patch TMA producer =====================
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < t.total; tile += gridDim.x) {
                const int s = tile / t.n_tiles_n;
                const int b = s / per_img, r = s % per_img;
                const int ty = r / t.tiles_x, tx = r % t.tiles_x;
                for (int ch = 0; ch < t.nchunks; ++ch, ++it) {
                    const uint32_t st = it % PS;
                    mbar_wait(&empty_p[st], ((it / PS) & 1) ^ 1);
                    mbar_arrive_expect_tx(&full_p[st], Cfg::PATCH_BYTES);
                    tma_load_4d(patch0 + st * Cfg::PATCH_STRIDE, &tmA, &full_p[st], ch * 64, tx * Cfg::TW - Cfg::HALO,
                                ty * Cfg::TH - Cfg::HALO, b);
                }
            }
        }
        __syncwarp();
    } else if (warp == 2) {
        // ===================== weight TMA producer =====================
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < t.total; tile += gridDim.x) {
                const int nt = tile % t.n_tiles_n;
                for (int ch = 0; ch < t.nchunks; ++ch) {
                    for (int tap = 0; tap < TAPS; ++tap, ++it) {
                        const uint32_t st = it % WS;
                        mbar_wait(&empty_w[st], ((it / WS) & 1) ^ 1);
                        mbar_arrive_expect_tx(&full_w[st], Cfg::W_BYTES);
                        tma_load_3d(wt0 + st * Cfg::W_BYTES, &tmW, &full_w[st], ch * 64, nt * N_TILE, tap);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer (single thread) =====================
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16_f32(128, N_TILE);
            const uint32_t p_u32 = smem_u32(patch0), w_u32 = smem_u32(wt0);
            uint32_t it_p = 0, it_w = 0, it_a = 0;
            for (int tile = blockIdx.x; tile < t.total; tile += gridDim.x, ++it_a) {
                const uint32_t as = it_a % AS;
                mbar_wait(&acc_empty[as], ((it_a / AS) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_base = tmem_base + as * (MT * N_TILE);
                for (int ch = 0; ch < t.nchunks; ++ch, ++it_p) {
                    const uint32_t ps = it_p % PS;
                    mbar_wait(&full_p[ps], (it_p / PS) & 1);
                    const uint32_t pbase = p_u32 + ps * Cfg::PATCH_STRIDE;
                    int kv = (a.Cin - ch * 64) / 16;
                    kv = kv > 4 ? 4 : kv;
                    for (int tap = 0; tap < TAPS; ++tap, ++it_w) {
                        const uint32_t ws = it_w % WS;
                        mbar_wait(&full_w[ws], (it_w / WS) & 1);
                        tc_fence_after();
                        const uint32_t wbase = w_u32 + ws * Cfg::W_BYTES;
                        const uint32_t arow = (TAPS == 9) ? (uint32_t)((tap / 3) * Cfg::PW + (tap % 3)) : 0u;
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {
                            for (int ks = 0; ks < kv; ++ks) {
                                const uint64_t ad = umma_smem_desc(pbase + (arow + 8 * mt) * 128 + ks * 32, 16,
                                                                   Cfg::PW * 128, UMMA_SW_128B);
                                const uint64_t bd = umma_smem_desc(wbase + ks * 32, 16, 1024, UMMA_SW_128B);
                                umma_bf16_ss(d_base + mt * N_TILE, ad, bd, idesc, (uint32_t)((ch | tap | ks) != 0));
                            }
                        }
                        umma_commit(&empty_w[ws]);
                    }
                    umma_commit(&empty_p[ps]);
                }
                umma_commit(&acc_full[as]);
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ===================== epilogue (4 warps, one TMEM lane quarter each) =====================
        const int q = warp & 3;
        const int row = q * 32 + lane;  // M row inside the 128-row UMMA tile
        uint32_t it_a = 0;
        for (int tile = blockIdx.x; tile < t.total; tile += gridDim.x, ++it_a) {
            const int nt = tile % t.n_tiles_n;
            const int s = tile / t.n_tiles_n;
            const int b = s / per_img, r = s % per_img;
            const int ty = r / t.tiles_x, tx = r % t.tiles_x;
            const uint32_t as = it_a % AS;
            mbar_wait(&acc_full[as], (it_a / AS) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int mt = 0; mt < MT; ++mt) {
                const int y = ty * Cfg::TH + (row >> 3);
                const int x = tx * Cfg::TW + 8 * mt + (row & 7);
                const size_t pix = ((size_t)b * a.H + y) * a.W + x;
#pragma unroll 1
                for (int c0 = 0; c0 < N_TILE; c0 += 32) {
                    uint32_t acc[32];
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + as * (MT * N_TILE) + mt * N_TILE + c0,
                                       acc);
                    tmem_ld_wait();
                    const int cb = nt * N_TILE + c0;  // first output channel of this 32-wide chunk
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
                    if (a.bias) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            const float4 bb = __ldg((const float4*)(a.bias + cb + j));
                            v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
                        }
                    }
                    if (a.act == ACT_LRELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : v[j] * a.slope;
                    } else if (a.act == ACT_GELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
                    }
                    if (a.res1) {
                        if (a.res1_f32) {
                            const float4* rp = (const float4*)((const float*)a.res1 + pix * a.res1_pitch + cb);
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 rr = rp[j];
                                v[4 * j] = v[4 * j] * a.alpha1 + rr.x;
                                v[4 * j + 1] = v[4 * j + 1] * a.alpha1 + rr.y;
                                v[4 * j + 2] = v[4 * j + 2] * a.alpha1 + rr.z;
                                v[4 * j + 3] = v[4 * j + 3] * a.alpha1 + rr.w;
                            }
                        } else {
                            const uint4* rp = (const uint4*)((const bf16*)a.res1 + pix * a.res1_pitch + cb);
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const uint4 rr = rp[j];
                                const __nv_bfloat162* h = (const __nv_bfloat162*)&rr;
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const float2 f = __bfloat1622float2(h[k]);
                                    v[8 * j + 2 * k] = v[8 * j + 2 * k] * a.alpha1 + f.x;
                                    v[8 * j + 2 * k + 1] = v[8 * j + 2 * k + 1] * a.alpha1 + f.y;
                                }
                            }
                        }
                    }
                    if (a.res2) {
                        float cs[32];
                        if (a.res2_cs) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4) {
                                const float4 c4 = __ldg((const float4*)(a.res2_cs + (size_t)b * a.Cout + cb + j));
                                cs[j] = c4.x; cs[j + 1] = c4.y; cs[j + 2] = c4.z; cs[j + 3] = c4.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) cs[j] = 1.f;
                        }
                        if (a.res2_f32) {
                            const float4* rp = (const float4*)((const float*)a.res2 + pix * a.res2_pitch + cb);
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 rr = rp[j];
                                v[4 * j] = v[4 * j] * a.alpha2 + rr.x * cs[4 * j];
                                v[4 * j + 1] = v[4 * j + 1] * a.alpha2 + rr.y * cs[4 * j + 1];
                                v[4 * j + 2] = v[4 * j + 2] * a.alpha2 + rr.z * cs[4 * j + 2];
                                v[4 * j + 3] = v[4 * j + 3] * a.alpha2 + rr.w * cs[4 * j + 3];
                            }
                        } else {
                            const uint4* rp = (const uint4*)((const bf16*)a.res2 + pix * a.res2_pitch + cb);
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const uint4 rr = rp[j];
                                const __nv_bfloat162* h = (const __nv_bfloat162*)&rr;
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const float2 f = __bfloat1622float2(h[k]);
                                    v[8 * j + 2 * k] = v[8 * j + 2 * k] * a.alpha2 + f.x * cs[8 * j + 2 * k];
                                    v[8 * j + 2 * k + 1] = v[8 * j + 2 * k + 1] * a.alpha2 + f.y * cs[8 * j + 2 * k + 1];
                                }
                            }
                        }
                    }
                    if (a.out32) {
                        float4* op = (float4*)(a.out32 + pix * a.out32_pitch + cb);
#pragma unroll
                        for (int j = 0; j < 8; ++j) op[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    if (a.out) {
                        if (!a.pixel_shuffle) {
                            uint4* op = (uint4*)(a.out + pix * a.out_pitch + cb);
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                uint4 o;
                                __nv_bfloat162* h = (__nv_bfloat162*)&o;
#pragma unroll
                                for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[8 * j + 2 * k], v[8 * j + 2 * k + 1]);
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
                }
            }
            tc_fence_before();
            mbar_arrive(&acc_empty[as]);
        }
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

static int encode_bf16_map(CUtensorMap* m, const void* ptr, int rank, const cuuint64_t* dims,
                           const cuuint64_t* strides_bytes, const cuuint32_t* box) {
    EncodeTiledFn enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
        return SRB_E_CUDA;
    }
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), dims,
                     strides_bytes, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]",
                  (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1],
                  (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0],
                  box[1], rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
        return SRB_E_CUDA;
    }
    return SRB_OK;
}

static int g_num_sms = 0;
static int num_sms() {
    if (g_num_sms == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
            g_num_sms = n;
        else
            g_num_sms = 148;
    }
    return g_num_sms;
}

template <class Cfg>
static int launch_cfg(const ConvArgs& a, cudaStream_t stream) {
    if (a.H % Cfg::TH || a.W % Cfg::TW) {
        set_error("conv_tc: H=%d W=%d not divisible by tile %dx%d", a.H, a.W, Cfg::TH, Cfg::TW);
        return SRB_E_SHAPE;
    }
    CUtensorMap tmA, tmW;
    {
        cuuint64_t dims[4] = {(cuuint64_t)a.Cin, (cuuint64_t)a.W, (cuuint64_t)a.H, (cuuint64_t)a.B};
        cuuint64_t str[3] = {(cuuint64_t)a.in_pitch * 2, (cuuint64_t)a.W * a.in_pitch * 2,
                             (cuuint64_t)a.H * a.W * a.in_pitch * 2};
        cuuint32_t box[4] = {64, (cuuint32_t)Cfg::PW, (cuuint32_t)Cfg::PH, 1};
        SRB_TRY(encode_bf16_map(&tmA, a.in, 4, dims, str, box));
    }
    {
        cuuint64_t dims[3] = {(cuuint64_t)a.Cin, (cuuint64_t)a.Cout, (cuuint64_t)Cfg::TAPS};
        cuuint64_t str[2] = {(cuuint64_t)a.Cin * 2, (cuuint64_t)a.Cout * a.Cin * 2};
        cuuint32_t box[3] = {64, (cuuint32_t)Cfg::N_TILE, 1};
        SRB_TRY(encode_bf16_map(&tmW, a.w, 3, dims, str, box));
    }
    ConvTiles t;
    t.tiles_x = a.W / Cfg::TW;
    t.tiles_y = a.H / Cfg::TH;
    t.n_tiles_n = a.Cout / Cfg::N_TILE;
    t.total = a.B * t.tiles_x * t.tiles_y * t.n_tiles_n;
    t.nchunks = (a.Cin + 63) / 64;
    static bool attr_set = false;  // per template instantiation
    if (!attr_set) {
        SRB_CUDA(cudaFuncSetAttribute(conv_tc_kernel<Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      Cfg::SMEM_BYTES));
        attr_set = true;
    }
    int grid = t.total < num_sms() ? t.total : num_sms();
    conv_tc_kernel<Cfg><<<grid, 256, Cfg::SMEM_BYTES, stream>>>(tmA, tmW, a, t);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int conv_tc_launch(const ConvArgs& a, cudaStream_t stream) {
    if (a.taps != 9 && a.taps != 1) { set_error("conv_tc: taps must be 9 or 1"); return SRB_E_ARG; }
    if (a.Cin % 16 || a.Cin < 16) { set_error("conv_tc: Cin=%d must be a multiple of 16", a.Cin); return SRB_E_SHAPE; }
    if (a.Cout % 32) { set_error("conv_tc: Cout=%d must be a multiple of 32", a.Cout); return SRB_E_SHAPE; }
    if (a.in_pitch % 8 || ((uintptr_t)a.in & 15) || ((uintptr_t)a.w & 15)) {
        set_error("conv_tc: input pitch / pointers must be 16-byte aligned");
        return SRB_E_ARG;
    }
    if (a.pixel_shuffle && (a.Cout % 128 || !a.out)) { set_error("conv_tc: pixel_shuffle needs Cout%%128==0"); return SRB_E_ARG; }
    if (a.taps == 9) {
        if (a.Cout == 32) return (a.W % 32 == 0) ? launch_cfg<ConvCfg<32, 4, 9>>(a, stream) : launch_cfg<ConvCfg<32, 2, 9>>(a, stream);
        if (a.Cout == 64) return (a.W % 32 == 0) ? launch_cfg<ConvCfg<64, 4, 9>>(a, stream) : launch_cfg<ConvCfg<64, 2, 9>>(a, stream);
        if (a.Cout % 128 == 0) return launch_cfg<ConvCfg<128, 2, 9>>(a, stream);
        if (a.Cout % 96 == 0) return launch_cfg<ConvCfg<96, 2, 9>>(a, stream);
    } else {
        if (a.Cout % 96 == 0) return launch_cfg<ConvCfg<96, 2, 1>>(a, stream);
        if (a.Cout % 64 == 0) return launch_cfg<ConvCfg<64, 2, 1>>(a, stream);
    }
    set_error("conv_tc: no kernel instance for Cout=%d taps=%d", a.Cout, a.taps);
    return SRB_E_SHAPE;
}

}  // namespace srb
