// Fused HAB attention branch:  xn = LN1(x),  t = x + proj( (S)W-MSA(xn) ).
// Replaces norm1 / attn.qkv / window attention / attn.proj / residual of the un-vendored HAT block (reached from
// src/architecture.py:78); torch.roll, window_partition / window_reverse are index math.
//
// One persistent CTA per SM walks tiles of 128 tokens = two 8x8 windows.  EVERY contraction of the branch runs on
// tcgen05 with TMEM accumulators (round 2: the attention core too, no mma.sync left):
//   qkv   3 x 6 MMAs  M128 N96 K16 (SS)                 -> TMEM cols [0, 288)
//   S     per (head, window): one M64 N64 K16 MMA Q K^T (SS: A = Q planes, B = K planes of that window), then per head four
//         M128 N64 TS MMAs that ADD THE RELATIVE-POSITION BIAS on the tensor pipe: A = one-hot rows held in TMEM (lane L
//         selects its query), B = the head's bias as a [key][query] operand (S += E * Bias^T): no thread ever loads or
//         adds a bias.  The M = 64
//         form puts row m of the window in lane 32*(m/16) + m%16 (+16 for the second window): both windows of the
//         pair share the 64 columns of S_h, one query row per thread -> softmax without shuffles.
//   P     exp2(S - max) (fp32 MUFU, fp32 row sum) as bf16 pairs, written back over the first half of S_h (tcgen05.st)
//   O     per head: 4 MMAs M128 N32 K16, A = P from TMEM (TS form), B = V as an MN-major operand holding
//         [window 0 dims | window 1 dims] along N: every row computes both products, each thread reads its window's
//         16 columns.  O_h occupies the second half of S_h.
//   proj  6 MMAs M128 N96 K16 (SS), A = normalised O (bf16) written over the Q planes -> TMEM cols [384, 480)
// A single thread can hand the tensor pipe one small MMA per ~45 cycles (tools/probe_issue.cu; the pipe itself takes
// 13 cycles for an N = 16 TS MMA), so three issuer warps (one per head group) feed it in parallel.
//
// 16 warps.  Warps 0-11 compute = 3 groups of 4 (TMEM lane quadrant = warp & 3): group g owns channel third g of the
// token-wise phases (LayerNorm, qkv epilogue part g = q | k | v, proj epilogue) and heads g, g+3 of the attention;
// they never issue an MMA and meet only at mbarriers (+ one named barrier inside the LayerNorm); their only global
// traffic is the proj epilogue (shortcut re-read from L2, t out).  Warp 12 is the I/O warp: x in one tile ahead and xn
// out as TMA tensor boxes (half a window = 4 rows x 8 tokens per box, straight into / out of the swizzled A1 chunks),
// plus the per-tile index bookkeeping (window coordinates, shift-mask region ids, pixel indices).  Warps 13-15 issue the
// MMAs of group 0 / 1 / 2: qkv part g, S + bias and PV of heads g, g+3 (warp 13 also proj).  Per tile i the compute
// warps run
//   qkv epilogue(i) -> Q/K/V planes | LN(i+1) in place on A1 | softmax(g), softmax(g+3) | O/rowsum -> Q planes |
//   proj epilogue(i) + x -> the group's own plane buffer -> HBM,   with S(i) + bias under LN(i+1), PV(g) under
//   softmax(g+3) and qkv(i+1) under the proj epilogue of tile i.
// Operand layouts: A1 (x -> xn, A of qkv) = the TMA-native swizzled K-major tiles (64 channels x 128 B rows SWIZZLE_128B
// + 32 channels x 64 B rows SWIZZLE_64B); Q / K / weights: un-swizzled K-major planes [k/8][row][8 bf16] (core matrix =
// 8 rows x 16 B contiguous, LBO = plane pitch, SBO = 128 B); V: [head][window][dim/8][key][8 dims] (MN-major core matrices: LBO = 128 B over keys, SBO = 1 KB
// over the four 8-dim groups); bias: [head][query/8][key][8 queries].  All forms validated on B200 by
// tools/probe_umma.cu / tools/probe_attn.cu.
#include <cuda.h>
#include <cstring>

#include "common.h"
#include "ops.h"
#include "ptx.cuh"
#include "tmap.h"

namespace srb {

namespace {

constexpr int kComputeWarps = 12;
constexpr int kComputeThreads = kComputeWarps * 32;
constexpr int kIoWarp = kComputeWarps;              // warp 12
constexpr int kIoThreads = 32;
constexpr int kIssuerWarp0 = kComputeWarps + 1;     // warps 13, 14, 15: MMA issuers of group 0, 1, 2
constexpr int kHabThreads = (kComputeWarps + 4) * 32;
constexpr int PP = 128 * 16 + 32;        // Q / K / staging plane pitch (bytes): 128 token rows x 16 B + a 32-byte skew that
                                         // makes the 4-lanes-per-token copy-out of the proj epilogue bank-conflict free
// A1 (x -> xn, the A operand of qkv) in the canonical swizzled K-major layouts the TMA writes at full speed: channels
// [0, 64) as 128 rows x 128 B (SWIZZLE_128B), channels [64, 96) as 128 rows x 64 B (SWIZZLE_64B).  One window row (8 tokens
// = 8 consecutive tile rows, 1536 contiguous bytes in HBM) = two boxes {64 ch, 8 tokens} + {32 ch, 8 tokens}.
constexpr int A1_C1 = 128 * 128;         // byte offset of the second channel chunk
__device__ __forceinline__ uint32_t a1_off(int row, int kg) {  // 16-byte unit kg (8 channels) of token row `row`
    return kg < 8 ? (uint32_t)(row * 128 + ((kg ^ (row & 7)) << 4))
                  : (uint32_t)(A1_C1 + row * 64 + (((kg - 8) ^ ((row >> 1) & 3)) << 4));
}
constexpr int WQ_PITCH = 288 * 16;       // qkv weight plane pitch
constexpr int WP_PITCH = 96 * 16;        // proj weight plane pitch
constexpr int VG = 64 * 16;              // V: one (head, window, 8-dim group) block = 64 keys x 16 B
constexpr int OFF_WQ = 0;
constexpr int OFF_WP = OFF_WQ + 12 * WQ_PITCH;
constexpr int OFF_A1 = OFF_WP + 12 * WP_PITCH;
constexpr int OFF_Q = OFF_A1 + 128 * 192;   // A1: 128 rows x 96 channels (two swizzled chunks)
constexpr int OFF_K = OFF_Q + 12 * PP;
constexpr int OFF_V = OFF_K + 12 * PP;
constexpr int OFF_BIAS = OFF_V + 24 * VG;               // bf16 [6][8][64][8] * log2(e): B operand planes [query/8][key][8 queries]
constexpr int OFF_PAR = OFF_BIAS + 6 * 64 * 64 * 2;     // bq[96] (+192 unused) | bproj'[96] | gamma[96] | beta[96]  (fp32)
constexpr int OFF_EX = OFF_PAR + (288 + 96 + 96 + 96) * 4;  // float2 [3][128] LayerNorm partials
constexpr int OFF_RID = OFF_EX + 3 * 128 * 8;           // int [2][128] shift-mask region ids (double buffered)
constexpr int OFF_GP = OFF_RID + 2 * 128 * 4;           // uint32 [3][128] pixel index of every tile row: written when x(j) is
                                                        // fetched (during tile j-1), read by the proj epilogue of tile j, by
                                                        // which time x(j+1) has been fetched -> three buffers
constexpr int OFF_BAR = OFF_GP + 3 * 128 * 4;
constexpr int HAB_TC_SMEM = OFF_BAR + 32 * 8;
constexpr uint32_t COL_T = 384, COL_E = 480, HAB_TMEM_COLS = 512;  // qkv: cols [0, 288); S/P/O of head h: cols [64h, 64h + 64);
                                                                   // E: 64 x 64 identity (bf16 pairs), cols [480, 512)
static_assert(HAB_TC_SMEM <= 227 * 1024, "shared memory budget");
static_assert(OFF_A1 % 1024 == 0 && A1_C1 % 1024 == 0, "the swizzled A1 chunks need 1024-byte alignment");
static_assert(OFF_A1 % 16 == 0 && OFF_Q % 16 == 0 && OFF_K % 16 == 0 && OFF_V % 16 == 0 && OFF_BIAS % 16 == 0 &&
              OFF_BAR % 8 == 0, "alignment");

struct HabArgs {
    const bf16* x;        // [B,H,W,96] token stream (shortcut)
    const float *ln_g, *ln_b;
    const bf16* wqkv;     // [288][96]
    const float* bqkv;    // [288]
    const bf16* bias;     // [6][8][64][8] bf16: relative-position bias * log2(e) as [head][query/8][key][query%8]
                          // (expand_rel_bias_bf16, NK = 64)
    const bf16* wproj;    // [96][96]
    const float* bproj;   // [96]
    bf16* xn;             // [B,H,W,96] LN1(x)
    bf16* t;              // [B,H,W,96] x + proj(attn)
    int B, H, W, shift;
    int ntiles;           // window pairs
    long long* dbg;       // optional [grid][8] per-phase cycle totals of thread 0 (development aid)
};

// Development aid (-DSRB_HAB_TRACE): clock64 time stamps of CTA 0 on its tiles 3 and 4, [tile][warp][32 slots] behind the
// phase totals (tools/hab_trace.py prints the timeline).
#ifdef SRB_HAB_TRACE
#define TR(k) do { if (a.dbg && blockIdx.x == 0 && (i == 3 || i == 4) && (threadIdx.x & 31) == 0) \
    a.dbg[2 * 148 * 8 + ((i - 3) * 16 + (threadIdx.x >> 5)) * 32 + (k)] = clock64(); } while (0)
#else
#define TR(k) do { } while (0)
#endif

__device__ __forceinline__ uint32_t packbf(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *(uint32_t*)&h;
}
__device__ __forceinline__ float2 unpackbf(uint32_t v) { return __bfloat1622float2(*(const __nv_bfloat162*)&v); }
constexpr float kLog2e = 1.4426950408889634f;
__device__ __forceinline__ float ex2f_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Softmax numerator of one query row of head h (thread = TMEM lane): the S row (64 fp32 columns at `tm_s`, bias
// already accumulated by the tensor pipe) (+ shift mask) -> P = exp2(. - max) as 32 bf16 pairs stored over S's first
// 32 columns; returns the fp32 row sum.
__device__ __forceinline__ float softmax_row(uint32_t tm_s, bool masked, const int* rid_w, int m) {
    float s[64];
    {
        uint32_t a[32], b[32];
        tmem_ld_32x32b_x32(tm_s, a);
        tmem_ld_32x32b_x32(tm_s + 32, b);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) { s[j] = __uint_as_float(a[j]); s[32 + j] = __uint_as_float(b[j]); }
    }
    if (masked) {  // shift mask (-100 across region borders): only windows touching the last window row / column
        const int rid_me = rid_w[m];
#pragma unroll
        for (int j = 0; j < 64; ++j)
            if (rid_w[j] != rid_me) s[j] -= 100.f * kLog2e;
    }
    float mx[4] = {s[0], s[1], s[2], s[3]};
#pragma unroll
    for (int j = 4; j < 64; j += 4) {
        mx[0] = fmaxf(mx[0], s[j]); mx[1] = fmaxf(mx[1], s[j + 1]);
        mx[2] = fmaxf(mx[2], s[j + 2]); mx[3] = fmaxf(mx[3], s[j + 3]);
    }
    const float mrow = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
    uint32_t p[32];
    float sum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const float e0 = ex2f_approx(s[2 * j] - mrow), e1 = ex2f_approx(s[2 * j + 1] - mrow);
        sum[j & 1] += e0;
        sum[2 + (j & 1)] += e1;
        p[j] = packbf(e0, e1);
    }
    tmem_st_32x32b_x32(tm_s, p);
    return (sum[0] + sum[1]) + (sum[2] + sum[3]);
}

__global__ void __launch_bounds__(kHabThreads, 1)
hab_tc_kernel(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
              const __grid_constant__ CUtensorMap tmN0, const __grid_constant__ CUtensorMap tmN1, const HabArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sWq = smem + OFF_WQ;
    uint8_t* sWp = smem + OFF_WP;
    uint8_t* sA1 = smem + OFF_A1;
    uint8_t* sQ = smem + OFF_Q;
    uint8_t* sK = smem + OFF_K;
    uint8_t* sV = smem + OFF_V;
    uint8_t* sBias = smem + OFF_BIAS;
    float* sbq = (float*)(smem + OFF_PAR);
    float* sbp = sbq + 288;
    float* sg = sbp + 96;
    float* sbeta = sg + 96;
    float2* sEx = (float2*)(smem + OFF_EX);
    int* sRid = (int*)(smem + OFF_RID);
    uint32_t* sGp = (uint32_t*)(smem + OFF_GP);
    uint64_t* bars = (uint64_t*)(smem + OFF_BAR);
    uint64_t* bar_a1 = bars + 0;    // A1 = xn(i) complete                      count kComputeThreads
    uint64_t* bar_qkv = bars + 1;   // [3] q / k / v part accumulated           tcgen05.commit (issuer g)
    uint64_t* bar_qkvw = bars + 4;  // Q / K / V planes written                 count kComputeThreads
    uint64_t* bar_s = bars + 5;     // [6] S_h + bias accumulated (2 windows)   tcgen05.commit (issuer h % 3)
    uint64_t* bar_p = bars + 11;    // [6] P_h of all 128 rows stored to TMEM   count 128 (group h % 3)
    uint64_t* bar_o = bars + 17;    // [6] O_h accumulated                      tcgen05.commit (issuer h % 3)
    uint64_t* bar_ao = bars + 23;   // normalised O in the Q planes, TMEM read  count kComputeThreads
    uint64_t* bar_t = bars + 24;    // proj accumulated                         tcgen05.commit (issuer 0)
    uint64_t* bar_x = bars + 25;    // x(i) landed in A1, region ids written    count kIoThreads
    uint32_t* tmem_slot = (uint32_t*)(bars + 26);
    int* sMsk = (int*)(bars + 28);  // [2][2]: does window w of the tile in buffer b touch the shift-mask border?

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = a.H, W = a.W, shift = a.shift;
    const int nwx = W / 8, nwy = H / 8;

    if (tid == 0) {
        if (smem_u32(sA1) & 1023u) __trap();  // swizzle pattern = address bits: the dynamic smem base must be 1 KB aligned
        mbar_init(bar_a1, kComputeThreads);
        for (int i = 0; i < 3; ++i) mbar_init(bar_qkv + i, 1);
        mbar_init(bar_qkvw, kComputeThreads);
        for (int i = 0; i < 6; ++i) { mbar_init(bar_s + i, 1); mbar_init(bar_p + i, 128); mbar_init(bar_o + i, 1); }
        mbar_init(bar_ao, kComputeThreads);
        mbar_init(bar_t, 1);
        mbar_init(bar_x, kIoThreads);
        fence_barrier_init();
    }
    if (warp == kIssuerWarp0) { tmem_alloc(tmem_slot, HAB_TMEM_COLS); tmem_relinquish(); }
    if (warp == kIoWarp && lane == 0) {
        tma_prefetch_desc(&tmX0); tma_prefetch_desc(&tmX1); tma_prefetch_desc(&tmN0); tma_prefetch_desc(&tmN1);
    }
    // constants of the launch
    for (int i = tid; i < 12 * 288; i += kHabThreads) {
        const int kg = i / 288, n = i % 288;
        *(uint4*)(sWq + kg * WQ_PITCH + n * 16) = __ldg((const uint4*)a.wqkv + n * 12 + kg);
    }
    for (int i = tid; i < 12 * 96; i += kHabThreads) {
        const int kg = i / 96, n = i % 96;
        *(uint4*)(sWp + kg * WP_PITCH + n * 16) = __ldg((const uint4*)a.wproj + n * 12 + kg);
    }
    for (int i = tid; i < 6 * 64 * 8; i += kHabThreads) ((uint4*)sBias)[i] = __ldg((const uint4*)a.bias + i);
    // Biases: q keeps its bias (pre-scaled like q itself).  The k bias adds q . bk to every logit of a row: softmax is
    // invariant to it, it is dropped.  The v bias passes through the normalised attention unchanged (rows of P / sum
    // add up to 1): it moves behind proj, bproj' = bproj + Wproj bv.
    if (tid < 96) {
        sbq[tid] = a.bqkv[tid] * (0.25f * kLog2e);
        float acc = a.bproj[tid];
        for (int c = 0; c < 96; ++c) acc = fmaf(__bfloat162float(a.wproj[tid * 96 + c]), a.bqkv[192 + c], acc);
        sbp[tid] = acc;
        sg[tid] = a.ln_g[tid];
        sbeta[tid] = a.ln_b[tid];
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) pdl_launch_dependents();
    pdl_wait();  // x comes from the previous kernel in the stream

    const int first = blockIdx.x, step = gridDim.x;
    const int n_my = (a.ntiles > first) ? (a.ntiles - first + step - 1) / step : 0;
    const int nwin = a.B * nwx * nwy;  // an odd window count leaves the second window of the last tile empty
    const uint32_t a1_u = smem_u32(sA1), q_u = smem_u32(sQ), k_u = smem_u32(sK), v_u = smem_u32(sV);
    const uint32_t wq_u = smem_u32(sWq), wp_u = smem_u32(sWp), bias_u = smem_u32(sBias);

    if (warp >= kIssuerWarp0) {
        // ===================== MMA issuer of group g: qkv part g, S + bias and PV of heads g, g+3 (g = 0: proj) =========
        const int g = warp - kIssuerWarp0;
        const uint32_t idesc = umma_idesc_bf16_f32(128, 96);
        const uint32_t idesc_s = umma_idesc_bf16_f32(64, 64);
        const uint32_t idesc_b = umma_idesc_bf16_f32(128, 64);
        const uint32_t idesc_pv = umma_idesc_bf16_f32_major(128, 32, 0, 1);
        auto issue_qkv_part = [&]() {
#pragma unroll
            for (int ks = 0; ks < 6; ++ks)  // k-steps 0-3: the 128B-swizzled chunk (32 bytes per step), 4-5: the 64B one
                umma_bf16_ss(tmem_base + 96 * g,
                             ks < 4 ? umma_smem_desc(a1_u + ks * 32, 16, 1024, UMMA_SW_128B)
                                    : umma_smem_desc(a1_u + A1_C1 + (ks - 4) * 32, 16, 512, UMMA_SW_64B),
                             umma_smem_desc(wq_u + g * 96 * 16 + ks * 2 * WQ_PITCH, WQ_PITCH, 128, UMMA_SW_NONE), idesc,
                             ks > 0);
            umma_commit(bar_qkv + g);
        };
        if (n_my > 0) {
            mbar_wait(bar_a1, 0);
            tc_fence_after();
            if (elect_one()) issue_qkv_part();
            __syncwarp();
        }
        for (int i = 0; i < n_my; ++i) {
            mbar_wait(bar_qkvw, i & 1);  // Q / K / V planes of tile i written, qkv accumulators read out
            tc_fence_after();
            TR(0);
            if (elect_one()) {
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int h = g + 3 * hh;
#pragma unroll
                    for (int w = 0; w < 2; ++w)  // S_h(window w) = Q K^T: M = 64 rows at lane offset 16 w
                        umma_bf16_ss(tmem_base + ((uint32_t)(16 * w) << 16) + 64 * h,
                                     umma_smem_desc(q_u + 2 * h * PP + 64 * w * 16, PP, 128, UMMA_SW_NONE),
                                     umma_smem_desc(k_u + 2 * h * PP + 64 * w * 16, PP, 128, UMMA_SW_NONE), idesc_s, 0);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)  // S_h += E * Bias_h^T for both windows at once: M = 128, lane L reads
                        umma_bf16_ts(tmem_base + 64 * h, tmem_base + COL_E + 8 * ks,  // its own one-hot row of E
                                     umma_smem_desc(bias_u + h * 8192 + ks * 2 * 1024, 1024, 128, UMMA_SW_NONE),
                                     idesc_b, 1);
                    umma_commit(bar_s + h);
                }
            }
            __syncwarp();
            TR(1);
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int h = g + 3 * hh;
                mbar_wait(bar_p + h, i & 1);  // P_h of all 128 rows is in TMEM
                tc_fence_after();
                TR(2 + 2 * hh);
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        umma_bf16_ts(tmem_base + 64 * h + 32, tmem_base + 64 * h + 8 * ks,
                                     umma_smem_desc(v_u + h * 4 * VG + ks * 16 * 16, 128, VG, UMMA_SW_NONE), idesc_pv,
                                     ks > 0);
                    umma_commit(bar_o + h);
                }
                __syncwarp();
                TR(3 + 2 * hh);
            }
            mbar_wait(bar_ao, i & 1);  // normalised O(i) in the Q planes; every TMEM read of tile i's S / P / O is done
            tc_fence_after();
            TR(6);
            if (g == 0) {
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < 6; ++ks)
                        umma_bf16_ss(tmem_base + COL_T, umma_smem_desc(q_u + ks * 2 * PP, PP, 128, UMMA_SW_NONE),
                                     umma_smem_desc(wp_u + ks * 2 * WP_PITCH, WP_PITCH, 128, UMMA_SW_NONE), idesc, ks > 0);
                    umma_commit(bar_t);
                }
                __syncwarp();
            }
            TR(7);
            if (i + 1 < n_my) {
                mbar_wait(bar_a1, (i + 1) & 1);  // LN(i+1) precedes the O read-out of tile i: already complete
                tc_fence_after();
                TR(8);
                if (elect_one()) issue_qkv_part();
                __syncwarp();
                TR(9);
            }
        }
    } else if (warp == kIoWarp) {
        // ===================== I/O warp: x rows HBM -> A1 one tile ahead, xn rows A1 -> HBM, both by TMA =====================
        // Lane p < 16 does the index math of window row p of the tile (window p >> 3, row p & 7); a window row is 8 tokens
        // x 96 channels = 1536 contiguous bytes in HBM and 8 consecutive rows of the swizzled A1 chunks, so half a window
        // (4 window rows) moves as two 3-D boxes {64 ch, 8 tokens, 4 rows} + {32 ch, 8 tokens, 4 rows}.  Only window rows
        // whose 8 tokens wrap around the image edge under the cyclic shift (last window column, shift > 0) go through
        // per-lane 16-byte copies.  cyclic shift = index math: rolled[i] = orig[i + shift] (torch.roll(-shift)).
        int x0_c = 0, gr_c = 0, x0_n = 0, gr_n = 0;  // lanes 0-3: TMA box (window lane >> 1, half lane & 1) of the current / next tile
        bool tma_c = false, tma_n = false;
        uint32_t wrap_c = 0, wrap_n = 0;  // bit p: window row p is copied by the lanes
        int r_wx = 0, r_wy = 0, r_gr = 0;  // lanes 0-15: window row 'lane' of the planned tile
        bool r_have = false;
        // index math of the CTA's next tile, registers only (runs before A1 is free)
        auto plan_regs = [&](int tile) {
            bool wrapx = false;
            r_have = false;
            if (lane < 16) {
                const int wi = lane >> 3, pr = lane & 7;
                const int win = 2 * tile + wi;
                r_have = win < nwin;  // an odd window count leaves the last tile's second window empty
                r_wx = win % nwx;
                const int wq = win / nwx;
                r_wy = wq % nwy;
                int y = r_wy * 8 + pr + shift;  // rolled-image row pr of the window -> image row
                if (y >= H) y -= H;
                r_gr = (wq / nwy) * H + y;
                wrapx = shift > 0 && r_wx == nwx - 1;
            }
            wrap_n = __ballot_sync(0xffffffffu, r_have && wrapx);
            // one TMA box = 4 window rows (half a window: 8-aligned windows shifted by 4 never wrap inside a half); lane
            // L < 4 issues the boxes of window L >> 1, half L & 1 (a TMA instruction costs its issuing warp ~90 cycles)
            const int src = lane < 4 ? 8 * (lane >> 1) + 4 * (lane & 1) : 0;
            x0_n = __shfl_sync(0xffffffffu, r_wx * 8 + shift, src);
            gr_n = __shfl_sync(0xffffffffu, r_gr, src);
            tma_n = __shfl_sync(0xffffffffu, (int)(r_have && !wrapx), src) != 0 && lane < 4;
        };
        // region ids / pixel indices / mask flags of the planned tile -> smem (buffers of tile j: see OFF_RID / OFF_GP)
        auto plan_smem = [&](int j) {
            const int buf = j & 1, gbuf = j % 3;
            if (lane < 16) {
                const int ys = r_wy * 8 + (lane & 7);
                const int rh = ys < H - 8 ? 0 : (ys < H - 4 ? 1 : 2);
                if ((lane & 7) == 0) sMsk[buf * 2 + (lane >> 3)] = (shift > 0 && (r_wx == nwx - 1 || r_wy == nwy - 1)) ? 1 : 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int xs = r_wx * 8 + k;
                    int x = xs + shift;
                    if (x >= W) x -= W;
                    const int rw = xs < W - 8 ? 0 : (xs < W - 4 ? 1 : 2);
                    sRid[buf * 128 + 8 * lane + k] = rh * 3 + rw;
                    sGp[gbuf * 128 + 8 * lane + k] = r_have ? (uint32_t)(r_gr * W + x) : 0xffffffffu;
                }
            }
            __syncwarp();
        };
        // x of the planned tile -> A1 (TMA boxes first, then the bookkeeping, then the wrapped rows); completes bar_x
        auto fetch = [&](int j) {
            const int gbuf = j % 3;
            const uint32_t ntma = __popc(__ballot_sync(0xffffffffu, tma_n));
            if (lane == 0) mbar_expect_tx(bar_x, ntma * 4 * 1536);
            __syncwarp();
            if (tma_n) {
                tma_load_3d(sA1 + lane * 4 * 1024, &tmX0, bar_x, 0, x0_n, gr_n);
                tma_load_3d(sA1 + A1_C1 + lane * 4 * 512, &tmX1, bar_x, 0, x0_n, gr_n);
            }
            plan_smem(j);
            for (uint32_t m = wrap_n; m; m &= m - 1) {
                const int p = __ffs(m) - 1;
#pragma unroll
                for (int c = lane; c < 96; c += 32) {
                    const int tok = c / 12, kg = c % 12;
                    const uint32_t gp = sGp[gbuf * 128 + 8 * p + tok];
                    cp_async_16(sA1 + a1_off(8 * p + tok, kg), (const uint4*)(a.x + (size_t)gp * 96) + kg);
                }
            }
            cp_async_commit_group();
            cp_async_wait_group<0>();
            mbar_arrive(bar_x);
        };
        if (n_my > 0) {
            plan_regs(first);
            fetch(0);
        }
        for (int i = 0; i < n_my; ++i) {
            const int tile = first + i * step;
            x0_c = x0_n; gr_c = gr_n; tma_c = tma_n; wrap_c = wrap_n;
            // xn(i) -> HBM for the CAB branch
            mbar_wait(bar_a1, i & 1);
            TR(0);
            if (tma_c) {
                tma_store_3d(&tmN0, sA1 + lane * 4 * 1024, 0, x0_c, gr_c);
                tma_store_3d(&tmN1, sA1 + A1_C1 + lane * 4 * 512, 0, x0_c, gr_c);
                bulk_commit_group();
            }
            for (uint32_t m = wrap_c; m; m &= m - 1) {
                const int p = __ffs(m) - 1;
#pragma unroll
                for (int c = lane; c < 96; c += 32) {
                    const int tok = c / 12, kg = c % 12;
                    const uint32_t gp = sGp[(i % 3) * 128 + 8 * p + tok];
                    *((uint4*)(a.xn + (size_t)gp * 96) + kg) = *(const uint4*)(sA1 + a1_off(8 * p + tok, kg));
                }
            }
            TR(1);
            if (i + 1 < n_my) {  // x(i+1) -> A1, free once all three qkv parts of tile i have been accumulated
                plan_regs(tile + step);
                bulk_wait_group_read<0>();  // the xn stores have read A1
                for (int part = 0; part < 3; ++part) mbar_wait(bar_qkv + part, i & 1);
                TR(2);
                // (the smem bookkeeping inside fetch() may not run earlier: sRid / sMsk[(i + 1) & 1] are read by the softmax
                // of tile i - 1, which precedes qkv(i))
                fetch(i + 1);
                TR(4);
            }
        }
        bulk_wait_group<0>();
    } else {
        // ===================== 12 compute warps =====================
        const int q = warp & 3, g = warp >> 2;
        const int r = q * 32 + lane;              // token row of the tile == TMEM lane of the M = 128 accumulators
        const int aw = lane >> 4;                 // attention: window of the pair held by this TMEM lane ...
        const int am = 16 * q + (lane & 15);      // ... and query row inside it (M = 64 accumulator layout)
        const int ar = 64 * aw + am;              // the same query as a tile row
        const uint32_t tm_lane = tmem_base + ((uint32_t)(q * 32) << 16);

        if (g == 0) {  // identity rows for the bias MMAs: E[m][k] = (k == m) as bf16 pairs, one row per TMEM lane
            uint32_t e[32];
#pragma unroll
            for (int c = 0; c < 32; ++c) e[c] = (c == (am >> 1)) ? ((am & 1) ? 0x3F800000u : 0x00003F80u) : 0u;
            tmem_st_32x32b_x32(tm_lane + COL_E, e);
            tmem_st_wait();
        }

#ifdef SRB_HAB_PHASES
        long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#define SRB_PHASE(k) do { if (a.dbg && (tid & 31) == 0 && tid < 64) { const long long tn = clock64(); ph[k] += tn - tprev; tprev = tn; } } while (0)
#else
#define SRB_PHASE(k) do { } while (0)
#endif
        // LayerNorm of the x parked in A1, in place (A1 becomes xn = the A operand of qkv)
        auto layer_norm = [&]() {
            uint4 xk[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) xk[u] = *(const uint4*)(sA1 + a1_off(r, 4 * g + u));
            float v[32];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t* w = (const uint32_t*)&xk[u];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = unpackbf(w[k]);
                    v[8 * u + 2 * k] = f.x; v[8 * u + 2 * k + 1] = f.y;
                }
            }
            float s4[4] = {0.f, 0.f, 0.f, 0.f};  // four independent chains instead of one 32-deep one
#pragma unroll
            for (int k = 0; k < 32; ++k) s4[k & 3] += v[k];
            const float mloc = ((s4[0] + s4[1]) + (s4[2] + s4[3])) * (1.f / 32.f);
            float q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 32; ++k) { const float d = v[k] - mloc; q4[k & 3] = fmaf(d, d, q4[k & 3]); }
            const float m2 = (q4[0] + q4[1]) + (q4[2] + q4[3]);
            sEx[g * 128 + r] = make_float2(mloc, m2);
            named_bar_sync(1, kComputeThreads);  // Chan merge of the three channel thirds
            const float2 e0 = sEx[r], e1 = sEx[128 + r], e2 = sEx[256 + r];
            const float mean = (e0.x + e1.x + e2.x) * (1.f / 3.f);
            const float d0 = e0.x - mean, d1 = e1.x - mean, d2 = e2.x - mean;
            const float var = (e0.y + e1.y + e2.y + 32.f * (d0 * d0 + d1 * d1 + d2 * d2)) * (1.f / 96.f);
            const float rs = rsqrtf(var + 1e-5f);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4 g0 = *(const float4*)(sg + 32 * g + 8 * u), g1 = *(const float4*)(sg + 32 * g + 8 * u + 4);
                const float4 b0 = *(const float4*)(sbeta + 32 * g + 8 * u), b1 = *(const float4*)(sbeta + 32 * g + 8 * u + 4);
                const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                uint4 o4;
                uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ow[k] = packbf(fmaf((v[8 * u + 2 * k] - mean) * rs, gg[2 * k], bb[2 * k]),
                                   fmaf((v[8 * u + 2 * k + 1] - mean) * rs, gg[2 * k + 1], bb[2 * k + 1]));
                *(uint4*)(sA1 + a1_off(r, 4 * g + u)) = o4;
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(bar_a1);
        };

        // proj epilogue of tile j: t = x + attn @ Wproj^T + b.  Thread = (row, channels [32 g, +32)) -> four planes of the
        // GROUP'S OWN operand buffer (Q / K / V region: proj, S and PV of tile j are complete; the group itself is the
        // next writer of that buffer), then the group copies its 64 bytes per token out, 4 lanes per token.
        uint8_t* stage = (g == 0) ? sQ : (g == 1) ? sK : sV;
        auto proj_epilogue = [&](int j) {
            // the shortcut x (this thread's row, 32 channels = 64 bytes) comes back from L2 under the proj MMAs: holding
            // it in registers since the LayerNorm cost 16 registers across the softmax and a local-memory spill
            uint4 xk[4];
            {
                const uint32_t gpr = sGp[(j % 3) * 128 + r];
                const uint4* xp = (const uint4*)(a.x + (size_t)(gpr != 0xffffffffu ? gpr : 0u) * 96 + 32 * g);
#pragma unroll
                for (int u = 0; u < 4; ++u) xk[u] = xp[u];  // plain loads: t may alias x (each tile reads its rows first)
            }
            mbar_wait(bar_t, j & 1);
            tc_fence_after();
            { const int i = j; TR(12); }
            uint32_t acc[32];
            tmem_ld_32x32b_x32(tm_lane + COL_T + 32 * g, acc);
            tmem_ld_wait();
            tc_fence_before();
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4 b0 = *(const float4*)(sbp + 32 * g + 8 * u), b1 = *(const float4*)(sbp + 32 * g + 8 * u + 4);
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                const uint32_t* xw = (const uint32_t*)&xk[u];
                uint4 o4;
                uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 xf = unpackbf(xw[k]);
                    ow[k] = packbf(__uint_as_float(acc[8 * u + 2 * k]) + bb[2 * k] + xf.x,
                                   __uint_as_float(acc[8 * u + 2 * k + 1]) + bb[2 * k + 1] + xf.y);
                }
                *(uint4*)(stage + u * PP + r * 16) = o4;
            }
            named_bar_sync(2 + g, 128);  // the group's 128 rows x 64 bytes are staged
            { const int i = j; TR(13); }
            const int gt = tid & 127, cu = gt & 3;
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const int row = 32 * p + (gt >> 2);
                const uint32_t gp = sGp[(j % 3) * 128 + row];
                if (gp != 0xffffffffu)  // (empty second window of an odd last tile)
                    *((uint4*)(a.t + (size_t)gp * 96 + 32 * g) + cu) = *(const uint4*)(stage + cu * PP + row * 16);
            }
            named_bar_sync(2 + g, 128);  // staging read out before the group's next qkv epilogue overwrites it
        };

        if (n_my > 0) {
            mbar_wait(bar_x, 0);
            layer_norm();
        }
        for (int i = 0; i < n_my; ++i) {
            const int tile = first + i * step;
            const bool have_next = (i + 1 < n_my);
            // ---- qkv epilogue: this warp's 32 rows x the 96 columns of part g -> Q / K / V ----
            TR(0);
            mbar_wait(bar_qkv + g, i & 1);
            tc_fence_after();
            TR(1);
            SRB_PHASE(0);
            {
                uint32_t acc[2][32];
                tmem_ld_32x32b_x32(tm_lane + 96 * g, acc[0]);
                if (g == 0) {  // q = (acc + bq) * head_dim^-0.5 * log2(e); the chunk's 32 biases are fetched under the TMEM load
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        float4 bb[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u) bb[u] = *(const float4*)(sbq + 32 * j + 4 * u);
                        tmem_ld_wait();  // chunk j is in registers; fetch chunk j+1 while it is converted
                        if (j < 2) tmem_ld_32x32b_x32(tm_lane + 32 * (j + 1), acc[(j + 1) & 1]);
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const uint32_t* ac = acc[j & 1] + 8 * u;
                            const float4 b0 = bb[2 * u], b1 = bb[2 * u + 1];
                            uint4 o4;
                            o4.x = packbf(fmaf(__uint_as_float(ac[0]), 0.25f * kLog2e, b0.x), fmaf(__uint_as_float(ac[1]), 0.25f * kLog2e, b0.y));
                            o4.y = packbf(fmaf(__uint_as_float(ac[2]), 0.25f * kLog2e, b0.z), fmaf(__uint_as_float(ac[3]), 0.25f * kLog2e, b0.w));
                            o4.z = packbf(fmaf(__uint_as_float(ac[4]), 0.25f * kLog2e, b1.x), fmaf(__uint_as_float(ac[5]), 0.25f * kLog2e, b1.y));
                            o4.w = packbf(fmaf(__uint_as_float(ac[6]), 0.25f * kLog2e, b1.z), fmaf(__uint_as_float(ac[7]), 0.25f * kLog2e, b1.w));
                            *(uint4*)(sQ + (4 * j + u) * PP + r * 16) = o4;
                        }
                    }
                } else {       // k, v: no bias (see the prologue)
                    // V[head][window][half][key]: MN-major operand of the PV MMAs; K: planes like Q
                    uint8_t* dst0 = (g == 2) ? sV + ((r >> 6) * 2) * VG + (r & 63) * 16 : sK + r * 16;
                    const int pitch = (g == 2) ? VG : PP;  // 8-channel group dgi -> V block (dgi >> 1) * 4 + (dgi & 1) / K plane dgi
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        tmem_ld_wait();
                        if (j < 2) tmem_ld_32x32b_x32(tm_lane + 96 * g + 32 * (j + 1), acc[(j + 1) & 1]);
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const uint32_t* ac = acc[j & 1] + 8 * u;
                            const int dgi = 4 * j + u;
                            const int blk = (g == 2) ? (dgi >> 1) * 4 + (dgi & 1) : dgi;
                            uint4 o4;
                            o4.x = packbf(__uint_as_float(ac[0]), __uint_as_float(ac[1]));
                            o4.y = packbf(__uint_as_float(ac[2]), __uint_as_float(ac[3]));
                            o4.z = packbf(__uint_as_float(ac[4]), __uint_as_float(ac[5]));
                            o4.w = packbf(__uint_as_float(ac[6]), __uint_as_float(ac[7]));
                            *(uint4*)(dst0 + blk * pitch) = o4;
                        }
                    }
                }
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(bar_qkvw);
            TR(2);
            SRB_PHASE(1);
            // ---- LN(i+1) in place on A1 while the issuers put S + bias of tile i on the tensor pipe (x(i+1) was fetched
            //      as soon as qkv(i) had been accumulated) ----
            if (have_next) {
                mbar_wait(bar_x, (i + 1) & 1);  // the I/O warp's cp.async of x(i+1) have landed
                TR(3);
                layer_norm();
            }
            TR(4);
            SRB_PHASE(4);
            // ---- softmax numerators of this group's two heads (S + bias arrives from the tensor pipe) ----
            float rowsum[2];
            {
                const bool masked = sMsk[(i & 1) * 2 + aw] != 0;  // (written by the I/O warp with the region ids)
                const int* rid_w = sRid + (i & 1) * 128 + 64 * aw;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int h = g + 3 * hh;
                    if (hh == 0) TR(14);
                    mbar_wait(bar_s + h, i & 1);
                    tc_fence_after();
                    TR(5 + 2 * hh);
                    rowsum[hh] = softmax_row(tm_lane + 64 * h, masked, rid_w, am);
                    tmem_st_wait();
                    tc_fence_before();
                    mbar_arrive(bar_p + h);
                    TR(6 + 2 * hh);
                    if (hh == 0) SRB_PHASE(2);
                }
            }
            SRB_PHASE(3);
            // ---- O / rowsum -> Q planes (rows of head h's two planes): the A operand of proj ----
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int h = g + 3 * hh;
                mbar_wait(bar_o + h, i & 1);
                tc_fence_after();
                TR(9 + hh);
                uint32_t o2[32];  // [window 0 dims | window 1 dims]: the TMEM address must be warp-uniform, so both
                tmem_ld_32x32b_x32(tm_lane + 64 * h + 32, o2);  // halves are loaded and each lane keeps its window's
                tmem_ld_wait();
                float o[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) o[k] = __uint_as_float(aw ? o2[16 + k] : o2[k]);
                const float inv = __fdividef(1.f, rowsum[hh]);
#pragma unroll
                for (int dg = 0; dg < 2; ++dg) {
                    uint4 o4;
                    uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                    for (int k = 0; k < 4; ++k) ow[k] = packbf(o[8 * dg + 2 * k] * inv, o[8 * dg + 2 * k + 1] * inv);
                    *(uint4*)(sQ + (2 * h + dg) * PP + ar * 16) = o4;
                }
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(bar_ao);
            TR(11);
            SRB_PHASE(5);
            // ---- proj epilogue(i) (direct 64-byte row-segment stores) while the tensor pipe runs qkv(i+1) ----
            proj_epilogue(i);
            TR(15);
            SRB_PHASE(6);
        }
#ifdef SRB_HAB_PHASES
        if (a.dbg && (tid & 31) == 0 && tid < 64)  // [0, grid): warp 0;  [grid, 2 grid): warp 1
            for (int k = 0; k < 8; ++k) a.dbg[((tid >> 5) * gridDim.x + blockIdx.x) * 8 + k] = ph[k];
#endif
#undef SRB_PHASE
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kIssuerWarp0) tmem_dealloc(tmem_base, HAB_TMEM_COLS);
}

}  // namespace

extern long long* g_fa_dbg;  // attn.cu (srb_debug_set_phase_buffer)

int hab_attention_tc(const bf16* x, const float* ln_g, const float* ln_b, const bf16* wqkv, const float* bqkv,
                     const bf16* bias_rows_log2e, const bf16* wproj, const float* bproj, bf16* xn, bf16* t, int B, int H,
                     int W, int shift, cudaStream_t s) {
    if (H % 8 || W % 8) { set_error("hab_attention: H=%d and W=%d must be multiples of the window size 8", H, W); return SRB_E_SHAPE; }
    if ((double)B * H * W >= 4294967296.0) { set_error("hab_attention: more than 2^32 tokens"); return SRB_E_SHAPE; }
    static std::atomic<unsigned long long> attr{0};
    SRB_TRY(opt_in_dynamic_smem(hab_tc_kernel, HAB_TC_SMEM, attr));
    HabArgs a;
    a.x = x; a.ln_g = ln_g; a.ln_b = ln_b; a.wqkv = wqkv; a.bqkv = bqkv; a.bias = bias_rows_log2e; a.wproj = wproj;
    a.bproj = bproj; a.xn = xn; a.t = t; a.B = B; a.H = H; a.W = W; a.shift = shift;
    a.ntiles = (B * (H / 8) * (W / 8) + 1) / 2;  // window pairs (an odd count: the last tile's second window is empty)
    a.dbg = g_fa_dbg;
    const int sms = num_sms();
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(a.ntiles < sms ? a.ntiles : sms);
    cfg.blockDim = dim3(kHabThreads);
    cfg.dynamicSmemBytes = HAB_TC_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute attrs[1];
    attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = 1;
    // x / xn as 3-D tensors {channels, W tokens, B * H image rows}, one map per channel chunk: box {64, 8, 4} (half a window) with
    // SWIZZLE_128B over channels [0, 64) and box {32, 8, 4} with SWIZZLE_64B over channels [64, 96)
    CUtensorMap tmX0, tmX1, tmN0, tmN1;
    {
        if ((((uintptr_t)x | (uintptr_t)xn | (uintptr_t)t) & 15)) { set_error("hab_attention: x / xn / t must be 16-byte aligned"); return SRB_E_ARG; }
        cuuint64_t d0[3] = {64, (cuuint64_t)W, (cuuint64_t)B * H}, d1[3] = {32, (cuuint64_t)W, (cuuint64_t)B * H};
        cuuint64_t str[2] = {192, (cuuint64_t)W * 192};
        cuuint32_t b0[3] = {64, 8, 4}, b1[3] = {32, 8, 4};
        SRB_TRY(encode_bf16_map(&tmX0, x, 3, d0, str, b0, CU_TENSOR_MAP_SWIZZLE_128B));
        SRB_TRY(encode_bf16_map(&tmX1, x + 64, 3, d1, str, b1, CU_TENSOR_MAP_SWIZZLE_64B));
        SRB_TRY(encode_bf16_map(&tmN0, xn, 3, d0, str, b0, CU_TENSOR_MAP_SWIZZLE_128B));
        SRB_TRY(encode_bf16_map(&tmN1, xn + 64, 3, d1, str, b1, CU_TENSOR_MAP_SWIZZLE_64B));
    }
    SRB_CUDA(cudaLaunchKernelEx(&cfg, hab_tc_kernel, tmX0, tmX1, tmN0, tmN1, a));
    return SRB_OK;
}

}  // namespace srb
