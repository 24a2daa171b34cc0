// Fused token-wise MLP branch of HAB / OCAB on tcgen05 (sm_100a):
//     y0 = t + cab * cs[b,:]          (HAB: attention branch + CAB * sigmoid(...) * conv_scale; OCAB: y0 = t)
//     y  = y0 + fc2( GELU( fc1( LayerNorm(y0) ) ) )
// Replaces norm2 / mlp.fc1 / act / mlp.fc2 / residual of the un-vendored HAT blocks (call site src/architecture.py:78).
//
// One persistent CTA per SM walks 128-token tiles.  Both GEMMs run on the 5th-gen tensor cores with the accumulators
// in TMEM; the 192-wide hidden activation never leaves the SM.  16 ALU warps + 1 MMA warp + 1 TMEM-allocator warp:
//   LN(i+1):  4 lanes per token (coalesced 64-byte row segments), t (+cab*cs) -> fp32 statistics by quad shuffles ->
//             (v - mean) * rstd as the bf16 A operand of fc1 (the LayerNorm affine is folded into fc1 at pack time:
//             W1' = W1 diag(gamma), b1' = b1 + W1 beta), y0 (bf16) parked in the output staging buffer
//   GELU(i):  D1 (TMEM 128x192 fp32) -> GELU -> bf16 A operand of fc2; thread = TMEM lane (row), 48 columns
//   EPI(i):   D2 (TMEM 128x96 fp32) + y0 (smem, in place) -> one 24 KB bulk store per tile
//   biases ride in the GEMMs: the A operands carry two constant 1.0 columns in an extra k-step and the weight planes
//   hold bias_hi / bias_lo (bf16 split, ~16 mantissa bits) there, so no epilogue touches a bias vector.
// ALU order GELU(i) -> LN(i+1) -> EPI(i) hides both MMAs behind ALU work and makes every buffer hazard implicit.
// Operand layout: un-swizzled K-major "planes" [k/8][row][8 bf16] (core matrix = 8 rows x 16 B contiguous; LBO = plane
// pitch, SBO = 128 B; validated on B200 by tools/probe_umma.cu "nosw planes"): 16-byte row writes are bank-conflict
// free and K = 96 / 192 need no padding.  Weights (fc1 192x96, fc2 96x192) stay resident in smem.
#include <cstring>

#include "act.cuh"
#include "common.h"
#include "ops.h"
#include "ptx.cuh"

namespace srb {

namespace {

constexpr int kAluWarps = 16;
constexpr int kAluThreads = kAluWarps * 32;
constexpr int kMlpThreads = kAluThreads + 64;
constexpr int A1_PITCH = 128 * 16 + 32;  // plane pitch; the 32-byte skew makes the 4-lanes-per-row LN stores conflict free
constexpr int A1_PLANES = 12 + 2;        // + one k-step carrying the constant-one columns (bias)
constexpr int A2_PITCH = 128 * 16;
constexpr int A2_PLANES = 24 + 2;
constexpr int W1_PITCH = 192 * 16;
constexpr int W2_PITCH = 96 * 16;
constexpr int YP_PITCH = 128 * 16 + 32;  // y0 / output tile as k-group planes (same skew as A1)
constexpr int YP_BYTES = 12 * YP_PITCH;
constexpr int kMaxImg = 64;              // images per launch (bf16 channel-attention table in smem)
constexpr int OFF_A1 = 0;
constexpr int OFF_A2 = OFF_A1 + A1_PLANES * A1_PITCH;
constexpr int OFF_W1 = OFF_A2 + A2_PLANES * A2_PITCH;
constexpr int OFF_W2 = OFF_W1 + A1_PLANES * W1_PITCH;
constexpr int OFF_Y = OFF_W2 + A2_PLANES * W2_PITCH;
constexpr int OFF_CS = OFF_Y + 2 * YP_BYTES;
constexpr int OFF_BAR = OFF_CS + kMaxImg * 192;
constexpr int OFF_LN2 = OFF_BAR + 64;  // [2][96] fp32: affine of the optional second LayerNorm output
constexpr int MLP_TC_SMEM = OFF_LN2 + 2 * 96 * 4;
constexpr uint32_t D1_COL = 0, D2_COL = 192, MLP_TMEM_COLS = 512;
static_assert(MLP_TC_SMEM <= 227 * 1024, "shared memory budget");
static_assert(OFF_A2 % 16 == 0 && OFF_W1 % 16 == 0 && OFF_W2 % 16 == 0 && OFF_Y % 16 == 0 && OFF_BAR % 8 == 0, "align");

struct MlpArgs {
    const bf16* t;
    const bf16* cab;   // nullable
    const float* cs;   // [B][96], nullable with cab
    const bf16* w1;    // [192][96]  LayerNorm-folded fc1 weight
    const float* b1;   // [192]      LayerNorm-folded fc1 bias
    const bf16* w2;    // [96][192]
    const float* b2;
    bf16* y;
    const float* ln2_g;  // optional second output xn2 = LayerNorm(y) * ln2_g + ln2_b (the next block's norm1,
    const float* ln2_b;  // fused into the copy-out: OCAB's norm1 after the last HAB of a group)
    bf16* xn2;
    uint32_t ntiles, tiles_per_img, tok_per_img;
    uint32_t nimg;  // images covered by this launch (cs table rows)
    long long* dbg;  // optional [grid][8] per-phase cycle totals of thread 0 (development aid)
};

__device__ __forceinline__ uint32_t packbf(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *(uint32_t*)&h;
}
__device__ __forceinline__ float2 unpackbf(uint32_t v) { return __bfloat1622float2(*(const __nv_bfloat162*)&v); }
// {bias_hi, bias_lo, 0 x 6} as one 16-byte k-group
__device__ __forceinline__ uint4 bias_split(float b) {
    const bf16 hi = __float2bfloat16_rn(b);
    const bf16 lo = __float2bfloat16_rn(b - __bfloat162float(hi));
    return make_uint4((uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(lo) << 16), 0u, 0u, 0u);
}

// Development aid (-DSRB_MLP_TRACE): clock64 stamps of CTA 0 on its tiles 3 and 4, [tile][warp 0..17][16 slots] behind
// the phase totals (tools/mlp_trace.py prints the timeline).
#ifdef SRB_MLP_TRACE
#define TR(k) do { if (a.dbg && blockIdx.x == 0 && (i == 3 || i == 4) && (threadIdx.x & 31) == 0) \
    a.dbg[148 * 8 + ((i - 3) * 18 + (threadIdx.x >> 5)) * 16 + (k)] = clock64(); } while (0)
#else
#define TR(k) do { } while (0)
#endif

template <bool LN2>
__global__ void __launch_bounds__(kMlpThreads, 1)
mlp_tc_kernel(const MlpArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA1 = smem + OFF_A1;
    uint8_t* sA2 = smem + OFF_A2;
    uint8_t* sW1 = smem + OFF_W1;
    uint8_t* sW2 = smem + OFF_W2;
    uint8_t* sYP = smem + OFF_Y;
    uint8_t* sCS = smem + OFF_CS;
    uint64_t* bars = (uint64_t*)(smem + OFF_BAR);
    float* sLN2 = (float*)(smem + OFF_LN2);
    uint64_t* bar_a1 = bars + 0;  // A1 (normalised tokens) written          count kAluThreads
    uint64_t* bar_a2 = bars + 1;  // A2 (GELU hidden) written                 count kAluThreads
    uint64_t* bar_d1 = bars + 2;  // fc1 accumulator complete                 count 1 (tcgen05.commit)
    uint64_t* bar_d2 = bars + 3;  // fc2 accumulator complete                 count 1
    uint32_t* tmem_slot = (uint32_t*)(bars + 4);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        mbar_init(bar_a1, kAluThreads); mbar_init(bar_a2, kAluThreads); mbar_init(bar_d1, 1); mbar_init(bar_d2, 1);
        fence_barrier_init();
    }
    if (warp == kAluWarps + 1) { tmem_alloc(tmem_slot, MLP_TMEM_COLS); tmem_relinquish(); }
    // weights are constants of the launch: stage them before waiting on the producer kernel (PDL prologue overlap)
    for (int i = tid; i < 12 * 192; i += kMlpThreads) {
        const int kg = i / 192, n = i % 192;
        *(uint4*)(sW1 + kg * W1_PITCH + n * 16) = __ldg((const uint4*)a.w1 + n * 12 + kg);
    }
    for (int i = tid; i < 24 * 96; i += kMlpThreads) {
        const int kg = i / 96, n = i % 96;
        *(uint4*)(sW2 + kg * W2_PITCH + n * 16) = __ldg((const uint4*)a.w2 + n * 24 + kg);
    }
    if (LN2 && tid < 192) sLN2[tid] = tid < 96 ? __ldg(a.ln2_g + tid) : __ldg(a.ln2_b + tid - 96);
    const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
    const uint4 ones4 = make_uint4(0x3F803F80u, 0u, 0u, 0u);  // {1.0, 1.0, 0, 0, 0, 0, 0, 0} bf16
    if (tid < 192) {
        *(uint4*)(sW1 + 12 * W1_PITCH + tid * 16) = bias_split(a.b1[tid]);
        *(uint4*)(sW1 + 13 * W1_PITCH + tid * 16) = zero4;
    }
    if (tid < 96) {
        *(uint4*)(sW2 + 24 * W2_PITCH + tid * 16) = bias_split(a.b2[tid]);
        *(uint4*)(sW2 + 25 * W2_PITCH + tid * 16) = zero4;
    }
    if (tid < 128) {
        *(uint4*)(sA1 + 12 * A1_PITCH + tid * 16) = ones4;
        *(uint4*)(sA1 + 13 * A1_PITCH + tid * 16) = zero4;
        *(uint4*)(sA2 + 24 * A2_PITCH + tid * 16) = ones4;
        *(uint4*)(sA2 + 25 * A2_PITCH + tid * 16) = zero4;
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) pdl_launch_dependents();
    pdl_wait();  // t / cab / cs come from the previous kernels in the stream

    const uint32_t first = blockIdx.x, step = gridDim.x;
    const uint32_t n_my = (a.ntiles > first) ? (a.ntiles - first + step - 1) / step : 0;

    if (warp == kAluWarps) {
        // ===================== MMA issuer =====================
        const uint32_t idesc1 = umma_idesc_bf16_f32(128, 192), idesc2 = umma_idesc_bf16_f32(128, 96);
        const uint32_t a1 = smem_u32(sA1), a2 = smem_u32(sA2), w1 = smem_u32(sW1), w2 = smem_u32(sW2);
        for (uint32_t i = 0; i < n_my; ++i) {
            mbar_wait(bar_a1, i & 1);
            tc_fence_after();
            TR(0);
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < A1_PLANES / 2; ++ks)
                    umma_bf16_ss(tmem_base + D1_COL,
                                 umma_smem_desc(a1 + ks * 2 * A1_PITCH, A1_PITCH, 128, UMMA_SW_NONE),
                                 umma_smem_desc(w1 + ks * 2 * W1_PITCH, W1_PITCH, 128, UMMA_SW_NONE), idesc1, ks > 0);
                umma_commit(bar_d1);
            }
            __syncwarp();
            TR(1);
            mbar_wait(bar_a2, i & 1);
            tc_fence_after();
            TR(2);
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < A2_PLANES / 2; ++ks)
                    umma_bf16_ss(tmem_base + D2_COL,
                                 umma_smem_desc(a2 + ks * 2 * A2_PITCH, A2_PITCH, 128, UMMA_SW_NONE),
                                 umma_smem_desc(w2 + ks * 2 * W2_PITCH, W2_PITCH, 128, UMMA_SW_NONE), idesc2, ks > 0);
                umma_commit(bar_d2);
            }
            __syncwarp();
            TR(3);
        }
    } else if (warp < kAluWarps) {
        // ===================== ALU warps =====================
        const int q = warp & 3, h4 = warp >> 2;
        const int r = q * 32 + lane;                 // GELU / EPI: token row inside the tile == TMEM lane
        const uint32_t tm_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const int ln_row = warp * 8 + (lane >> 2);   // LN / copy-out: 4 lanes per token, lane j owns 16-byte units j, j+4, j+8
        const int ln_j = lane & 3;

        uint4 pt[3], pc[3];                          // prefetched row thirds of t and cab
        uint4 csr[3];                                // channel-attention scales (packed bf16) of image cs_img
        uint32_t cs_img = 0xffffffffu;
#ifdef SRB_MLP_PHASES
        long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#endif
#ifdef SRB_MLP_PHASES
#define SRB_PHASE(k) do { if (a.dbg && tid == 0) { const long long tn = clock64(); ph[k] += tn - tprev; tprev = tn; } } while (0)
#else
#define SRB_PHASE(k) do { } while (0)
#endif

        // token row of this thread in `tile`: an image is ceil(tok_per_img / 128) tiles, the last one possibly ragged
        // (HAT token counts are multiples of 64): rows beyond the image read a clamped (valid) row and are never stored
        auto token_row = [&](uint32_t tile, bool& valid) -> size_t {
            const uint32_t img = tile / a.tiles_per_img;
            uint32_t r = (tile - img * a.tiles_per_img) * 128 + ln_row;
            valid = r < a.tok_per_img;
            if (!valid) r = a.tok_per_img - 1;
            return (size_t)img * a.tok_per_img + r;
        };
        auto issue_loads = [&](uint32_t tile) {
            bool valid;
            const size_t row = token_row(tile, valid);
            const uint4* tp = (const uint4*)(a.t + row * 96) + ln_j;
#pragma unroll
            for (int m = 0; m < 3; ++m) pt[m] = tp[4 * m];  // plain loads: y may alias t (OCAB runs in place)
            if (a.cab) {
                const uint4* cp = (const uint4*)(a.cab + row * 96) + ln_j;
#pragma unroll
                for (int m = 0; m < 3; ++m) pc[m] = __ldg(cp + 4 * m);
            }
        };

        // combine + LayerNorm (no affine) of the prefetched tile -> A1; residual y0 (bf16) -> plane buffer YP[it & 1]
        auto layer_norm = [&](uint32_t tile, uint32_t it) {
            if (a.cab) {
                // y0 = t + cab * cs in packed bf16 (one rounding: exactly the value the residual stream stores)
                const uint32_t img = tile / a.tiles_per_img;
                if (img != cs_img) {  // warp-uniform; a CTA changes image every few tiles
                    const uint4* csp = (const uint4*)(sCS + img * 192) + ln_j;
#pragma unroll
                    for (int m = 0; m < 3; ++m) csr[m] = csp[4 * m];
                    cs_img = img;
                }
#pragma unroll
                for (int m = 0; m < 3; ++m) {
                    const uint4 s4 = csr[m];
                    __nv_bfloat162* tv = (__nv_bfloat162*)&pt[m];
                    const __nv_bfloat162* cv = (const __nv_bfloat162*)&pc[m];
                    const __nv_bfloat162* sv = (const __nv_bfloat162*)&s4;
#pragma unroll
                    for (int k = 0; k < 4; ++k) tv[k] = __hfma2(cv[k], sv[k], tv[k]);
                }
            }
            uint8_t* yp = sYP + (it & 1) * YP_BYTES + ln_row * 16;
#pragma unroll
            for (int m = 0; m < 3; ++m) *(uint4*)(yp + (ln_j + 4 * m) * YP_PITCH) = pt[m];
            SRB_PHASE(2);
            float v[24];
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                const uint32_t* w = (const uint32_t*)&pt[m];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = unpackbf(w[k]);
                    v[8 * m + 2 * k] = f.x; v[8 * m + 2 * k + 1] = f.y;
                }
            }
            float s4[4] = {0.f, 0.f, 0.f, 0.f};  // four independent chains instead of one 24-deep one
#pragma unroll
            for (int k = 0; k < 24; ++k) s4[k & 3] += v[k];
            float s = (s4[0] + s4[1]) + (s4[2] + s4[3]);
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            const float mean = s * (1.f / 96.f);
            float q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 24; ++k) { v[k] -= mean; q4[k & 3] = fmaf(v[k], v[k], q4[k & 3]); }
            float m2 = (q4[0] + q4[1]) + (q4[2] + q4[3]);
            m2 += __shfl_xor_sync(0xffffffffu, m2, 1);
            m2 += __shfl_xor_sync(0xffffffffu, m2, 2);
            const float rs = rsqrtf(m2 * (1.f / 96.f) + 1e-5f);
#pragma unroll
            for (int m = 0; m < 3; ++m)
                *(uint4*)(sA1 + (ln_j + 4 * m) * A1_PITCH + ln_row * 16) =
                    make_uint4(packbf(v[8 * m] * rs, v[8 * m + 1] * rs), packbf(v[8 * m + 2] * rs, v[8 * m + 3] * rs),
                               packbf(v[8 * m + 4] * rs, v[8 * m + 5] * rs), packbf(v[8 * m + 6] * rs, v[8 * m + 7] * rs));
            SRB_PHASE(3);
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(bar_a1);
            SRB_PHASE(4);
        };

        // per-image channel-attention scales of this launch -> packed bf16 table (read by every LN)
        if (a.cab)
            for (int i = tid; i < (int)a.nimg * 96; i += kAluThreads)
                ((bf16*)sCS)[i] = __float2bfloat16_rn(a.cs[i]);
        named_bar_sync(1, kAluThreads);
        if (n_my > 0) {
            issue_loads(first);
            layer_norm(first, 0);
            if (n_my > 1) issue_loads(first + step);
        }
        for (uint32_t i = 0; i < n_my; ++i) {
            const uint32_t tile = first + i * step;
            // ---- (a) GELU(i): D1 -> A2 ----
            SRB_PHASE(7);
            TR(0);
            mbar_wait(bar_d1, i & 1);
            tc_fence_after();
            TR(1);
            SRB_PHASE(0);
            {
                uint32_t acc[48];
                {
                    uint32_t lo[32], hi[16];
                    // (h4 laundered: hoisted out of the tile loop, the two TMEM addresses are spilled in the LN2 variant and
                    // reloaded from local memory right behind the accumulator wait)
                    int h4v = h4;
                    if (LN2) asm volatile("" : "+r"(h4v));
                    tmem_ld_32x32b_x32(tm_lane + D1_COL + 48 * h4v, lo);
                    tmem_ld_32x32b_x16(tm_lane + D1_COL + 48 * h4v + 32, hi);
                    tmem_ld_wait();
#pragma unroll
                    for (int k = 0; k < 32; ++k) acc[k] = lo[k];
#pragma unroll
                    for (int k = 0; k < 16; ++k) acc[32 + k] = hi[k];
                }
#pragma unroll
                for (int u = 0; u < 6; ++u) {
                    uint4 o4;
                    uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        ow[k] = packbf(gelu_fast(__uint_as_float(acc[8 * u + 2 * k])),
                                       gelu_fast(__uint_as_float(acc[8 * u + 2 * k + 1])));
                    *(uint4*)(sA2 + (6 * h4 + u) * A2_PITCH + r * 16) = o4;
                }
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(bar_a2);
            TR(2);
            SRB_PHASE(1);
            // ---- (b) LN(i+1) -> A1 (fc1(i) has completed: bar_d1 was waited), prefetch tile i+2 ----
            // the barrier orders copy-out(i-1) (last readers of YP[(i+1)&1]) before the y0(i+1) stores
            named_bar_sync(1, kAluThreads);
            TR(3);
            SRB_PHASE(1);
            if (i + 1 < n_my) {
                layer_norm(tile + step, i + 1);
            }
            TR(4);
            SRB_PHASE(5);
            // ---- (c) EPI(i): D2 + y0 -> YP[i&1] in place (row per lane: conflict free) ----
            mbar_wait(bar_d2, i & 1);
            tc_fence_after();
            TR(5);
            {
                uint32_t lo[16], hi[8];
                int h4v = h4;
                if (LN2) asm volatile("" : "+r"(h4v));
                tmem_ld_32x32b_x16(tm_lane + D2_COL + 24 * h4v, lo);
                tmem_ld_32x32b_x8(tm_lane + D2_COL + 24 * h4v + 16, hi);
                tmem_ld_wait();
                uint8_t* yp = sYP + (i & 1) * YP_BYTES + (3 * h4) * YP_PITCH + r * 16;
#pragma unroll
                for (int u = 0; u < 3; ++u) {
                    const uint4 yy = *(const uint4*)(yp + u * YP_PITCH);
                    const uint32_t* yw = (const uint32_t*)&yy;
                    uint4 o4;
                    uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float2 f = unpackbf(yw[k]);
                        const uint32_t c0 = u < 2 ? lo[8 * u + 2 * k] : hi[2 * k];
                        const uint32_t c1 = u < 2 ? lo[8 * u + 2 * k + 1] : hi[2 * k + 1];
                        ow[k] = packbf(__uint_as_float(c0) + f.x, __uint_as_float(c1) + f.y);
                    }
                    *(uint4*)(yp + u * YP_PITCH) = o4;
                }
            }
            tc_fence_before();
            TR(6);
            SRB_PHASE(6);
            named_bar_sync(2, kAluThreads);
            TR(7);
            // ---- (d) copy-out(i): 4 lanes per token, 64-byte row segments -> fully coalesced 16-byte stores ----
            {
                const uint8_t* yp = sYP + (i & 1) * YP_BYTES + ln_row * 16;
                bool valid;
                const size_t orow = token_row(tile, valid);
                uint4* gp = (uint4*)(a.y + orow * 96) + ln_j;
                uint4 o[3];
#pragma unroll
                for (int m = 0; m < 3; ++m) {
                    o[m] = *(const uint4*)(yp + (ln_j + 4 * m) * YP_PITCH);
                    if (valid) gp[4 * m] = o[m];
                }
                if (LN2) {  // xn2 = LayerNorm(y) of the bf16 row just written (4 lanes per token: two shuffles per sum)
                    // the row stays packed (12 registers) and is unpacked in each of the three passes: holding 24 floats
                    // next to the prefetched tiles spilled, which cost this variant 110 us per launch
                    float sm = 0.f;
#pragma unroll
                    for (int m = 0; m < 3; ++m) {
                        const uint32_t* w = (const uint32_t*)&o[m];
#pragma unroll
                        for (int k = 0; k < 4; ++k) { const float2 f = unpackbf(w[k]); sm += f.x + f.y; }
                    }
                    sm += __shfl_xor_sync(0xffffffffu, sm, 1);
                    sm += __shfl_xor_sync(0xffffffffu, sm, 2);
                    const float mean = sm * (1.f / 96.f);
                    float qq = 0.f;
#pragma unroll
                    for (int m = 0; m < 3; ++m) {
                        const uint32_t* w = (const uint32_t*)&o[m];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpackbf(w[k]);
                            qq = fmaf(f.x - mean, f.x - mean, qq);
                            qq = fmaf(f.y - mean, f.y - mean, qq);
                        }
                    }
                    qq += __shfl_xor_sync(0xffffffffu, qq, 1);
                    qq += __shfl_xor_sync(0xffffffffu, qq, 2);
                    const float rstd = rsqrtf(qq * (1.f / 96.f) + 1e-5f);
                    uint4* xp = (uint4*)(a.xn2 + orow * 96) + ln_j;
#pragma unroll
                    for (int m = 0; m < 3; ++m) {
                        const float* gg = sLN2 + (ln_j + 4 * m) * 8;
                        const uint32_t* w = (const uint32_t*)&o[m];
                        uint4 x4;
                        uint32_t* xw = (uint32_t*)&x4;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpackbf(w[k]);
                            xw[k] = packbf(fmaf((f.x - mean) * rstd, gg[2 * k], gg[96 + 2 * k]),
                                           fmaf((f.y - mean) * rstd, gg[2 * k + 1], gg[96 + 2 * k + 1]));
                        }
                        if (valid) xp[4 * m] = x4;
                    }
                }
            }
            TR(8);
            if (i + 2 < n_my) issue_loads(tile + 2 * step);
            TR(9);
        }
#ifdef SRB_MLP_PHASES
        if (a.dbg && tid == 0)
            for (int k = 0; k < 8; ++k) a.dbg[blockIdx.x * 8 + k] = ph[k];
#endif
#undef SRB_PHASE
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kAluWarps + 1) tmem_dealloc(tmem_base, MLP_TMEM_COLS);
}

// W'[n][k] = W[n][k] * gamma[k] (bf16), b'[n] = b[n] + sum_k W[n][k] * beta[k]: LayerNorm affine folded into the
// following Linear (fp32 inputs, one rounding).  grid = N blocks of 128 threads.
__global__ void fold_ln_linear_kernel(const float* __restrict__ w, const float* __restrict__ b,
                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                      bf16* __restrict__ w_out, float* __restrict__ b_out, int K) {
    __shared__ float red[4];
    const int n = blockIdx.x;
    float acc = 0.f;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const float wv = w[(size_t)n * K + k];
        w_out[(size_t)n * K + k] = __float2bfloat16_rn(wv * gamma[k]);
        acc = fmaf(wv, beta[k], acc);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) b_out[n] = b[n] + red[0] + red[1] + red[2] + red[3];
}

}  // namespace

extern long long* g_fa_dbg;  // attn.cu (srb_debug_set_phase_buffer)

int fold_ln_linear(const float* w, const float* b, const float* gamma, const float* beta, bf16* w_out, float* b_out,
                   int N, int K, cudaStream_t s) {
    fold_ln_linear_kernel<<<N, 128, 0, s>>>(w, b, gamma, beta, w_out, b_out, K);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int mlp_fused(const bf16* t, const bf16* cab, const float* cs, const bf16* w1f, const float* b1f, const bf16* w2,
              const float* b2, bf16* y, size_t ntok, size_t tok_per_img, cudaStream_t s, const float* ln2_g,
              const float* ln2_b, bf16* xn2) {
    if (ntok == 0 || tok_per_img == 0 || tok_per_img >= ((size_t)1 << 31)) { set_error("mlp_fused: bad token counts"); return SRB_E_SHAPE; }
    if (((uintptr_t)t | (uintptr_t)y | (uintptr_t)cab | (uintptr_t)w1f | (uintptr_t)w2) & 15) {
        set_error("mlp_fused: pointers must be 16-byte aligned");
        return SRB_E_ARG;
    }
    if (xn2 && (!ln2_g || !ln2_b || ((uintptr_t)xn2 & 15))) { set_error("mlp_fused: xn2 needs ln2_g / ln2_b and 16-byte alignment"); return SRB_E_ARG; }
    static std::atomic<unsigned long long> attr0{0}, attr1{0};
    SRB_TRY(opt_in_dynamic_smem(mlp_tc_kernel<false>, MLP_TC_SMEM, attr0));
    SRB_TRY(opt_in_dynamic_smem(mlp_tc_kernel<true>, MLP_TC_SMEM, attr1));
    const size_t nimg_total = ntok / tok_per_img;
    if (ntok % tok_per_img) { set_error("mlp_fused: ntok must be a multiple of tok_per_img"); return SRB_E_SHAPE; }
    const int sms = num_sms();
    for (size_t img = 0; img < nimg_total; img += kMaxImg) {  // the smem cs table holds kMaxImg images per launch
        const size_t nimg = nimg_total - img < (size_t)kMaxImg ? nimg_total - img : (size_t)kMaxImg;
        const size_t tok0 = img * tok_per_img;
        MlpArgs a;
        a.t = t + tok0 * 96; a.cab = cab ? cab + tok0 * 96 : nullptr; a.cs = cs ? cs + img * 96 : nullptr;
        a.w1 = w1f; a.b1 = b1f; a.w2 = w2; a.b2 = b2; a.y = y + tok0 * 96;
        a.ln2_g = ln2_g; a.ln2_b = ln2_b; a.xn2 = xn2 ? xn2 + tok0 * 96 : nullptr;
        a.tiles_per_img = (uint32_t)((tok_per_img + 127) / 128);
        a.ntiles = (uint32_t)(nimg * a.tiles_per_img);
        a.tok_per_img = (uint32_t)tok_per_img;
        a.nimg = (uint32_t)nimg;
        a.dbg = g_fa_dbg;
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(a.ntiles < (uint32_t)sms ? a.ntiles : sms);
        cfg.blockDim = dim3(kMlpThreads);
        cfg.dynamicSmemBytes = MLP_TC_SMEM;
        cfg.stream = s;
        cudaLaunchAttribute attrs[1];
        attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attrs[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attrs;
        cfg.numAttrs = 1;
        if (xn2) SRB_CUDA(cudaLaunchKernelEx(&cfg, mlp_tc_kernel<true>, a));
        else SRB_CUDA(cudaLaunchKernelEx(&cfg, mlp_tc_kernel<false>, a));
    }
    return SRB_OK;
}

}  // namespace srb
