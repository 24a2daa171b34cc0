// Fused HAB attention branch:  xn = LN1(x),  t = x + proj( (S)W-MSA(xn) ).
// Replaces norm1 / attn.qkv / window attention / attn.proj / residual of the un-vendored HAT block (reached from
// src/architecture.py:78); torch.roll, window_partition / window_reverse are index math.
//
// One persistent CTA per SM walks tiles of 128 tokens = two 8x8 windows.  The two dense GEMMs of the branch run on
// tcgen05 with TMEM accumulators (M = 128 tokens; qkv as three N = 96 parts, proj N = 96, K = 96 = 6 k-steps each);
// the per-head 64x64 attention cores (head_dim 16 = one k-step) stay on warp-level mma.sync, fed by ldmatrix from the
// same shared-memory operand planes the tensor cores use:
//   operand layout: un-swizzled K-major planes [k/8][row][8 bf16] (core matrix = 8 rows x 16 B contiguous, LBO =
//   plane pitch, SBO = 128 B) -> a plane row segment of 8 tokens is exactly one ldmatrix 8x8 tile, and a thread that
//   owns one token row writes 16-byte units without bank conflicts.
// 12 warps (TMEM lane quadrant = warp & 3, channel third = warp >> 2; 3 warps per SM sub-partition so that a thread
// may use up to 168 registers); warp 0 also allocates TMEM and issues the tcgen05 MMAs:
//   LN      thread = (token row, 32 channels): x from HBM -> registers (kept for the shortcut), statistics merged
//           across the three channel thirds through shared memory, xn -> HBM (CAB branch) and -> A1 planes
//   qkv     3 x 6 tcgen05.mma (q | k | v parts, one commit each) -> TMEM columns [0, 288)
//   qkv-epi warp third c reads part c (its 32 rows x 96 columns), + bias, q pre-scaled by head_dim^-0.5 * log2(e),
//           -> Q / K / V planes
//   attn    warp = (window, head): S = bias + Q K^T (bias is the accumulator init), shift mask, exp2 softmax,
//           O = P V, normalised O overwrites the head's two Q planes (= A operand of proj)
//   proj    6 tcgen05.mma -> TMEM columns [288, 384);  epi: + bias + x (registers) -> t -> HBM
#include <cstring>

#include "common.h"
#include "ops.h"
#include "ptx.cuh"

namespace srb {

namespace {

constexpr int kComputeWarps = 12;
constexpr int kComputeThreads = kComputeWarps * 32;
constexpr int kHabThreads = kComputeThreads;  // 3 warps per SM sub-partition -> up to 168 registers per thread
constexpr int PP = 128 * 16 + 16;        // activation plane pitch (bytes): 128 token rows x 16 B + a 16-byte skew that
                                         // spreads the 6-lanes-per-token global I/O pattern over the banks
constexpr int WQ_PITCH = 288 * 16;       // qkv weight plane pitch
constexpr int WP_PITCH = 96 * 16;        // proj weight plane pitch
constexpr int BIAS_ELEMS = 6 * 4 * 8 * 32 * 4;  // bf16, MMA C-fragment order per head (expand_rel_bias_bf16)
constexpr int OFF_WQ = 0;
constexpr int OFF_WP = OFF_WQ + 12 * WQ_PITCH;
constexpr int OFF_A1 = OFF_WP + 12 * WP_PITCH;
constexpr int OFF_Q = OFF_A1 + 12 * PP;
constexpr int OFF_K = OFF_Q + 12 * PP;
constexpr int OFF_V = OFF_K + 12 * PP;
constexpr int OFF_BIAS = OFF_V + 12 * PP;
constexpr int OFF_PAR = OFF_BIAS + BIAS_ELEMS * 2;      // bqkv[288] | bproj[96] | gamma[96] | beta[96]  (fp32)
constexpr int OFF_EX = OFF_PAR + (288 + 96 + 96 + 96) * 4;  // float2 [3][128] LayerNorm partials
constexpr int OFF_RID = OFF_EX + 3 * 128 * 8;           // int [128] shift-mask region ids
constexpr int OFF_BAR = OFF_RID + 128 * 4;
constexpr int HAB_TC_SMEM = OFF_BAR + 64;
constexpr uint32_t COL_QKV = 0, COL_T = 288, HAB_TMEM_COLS = 512;
static_assert(HAB_TC_SMEM <= 227 * 1024, "shared memory budget");

struct HabArgs {
    const bf16* x;        // [B,H,W,96] token stream (shortcut)
    const float *ln_g, *ln_b;
    const bf16* wqkv;     // [288][96]
    const float* bqkv;    // [288]
    const bf16* bias_frag;  // [6][4][8][32][4] bf16, relative-position bias * log2(e) in C-fragment order
    const bf16* wproj;    // [96][96]
    const float* bproj;   // [96]
    bf16* xn;             // [B,H,W,96] LN1(x)
    bf16* t;              // [B,H,W,96] x + proj(attn)
    int B, H, W, shift;
    int ntiles;           // window pairs
    long long* dbg;       // optional [grid][8] per-phase cycle totals of thread 0 (development aid)
};

__device__ __forceinline__ uint32_t packbf(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *(uint32_t*)&h;
}
__device__ __forceinline__ float2 unpackbf(uint32_t v) { return __bfloat1622float2(*(const __nv_bfloat162*)&v); }
// two exp2 per MUFU op on packed bf16 (input and result are bf16x2)
__device__ __forceinline__ uint32_t ex2_bf16x2(uint32_t x) {
    uint32_t y;
    asm("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
    return y;
}
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t saddr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(saddr));
}
__device__ __forceinline__ void ldsm4_trans(uint32_t (&r)[4], uint32_t saddr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(saddr));
}

constexpr float kLog2e = 1.4426950408889634f;

// One (window, head) unit, query rows [32*mh, 32*mh+32) of the window: two m-tiles interleaved.
//   q_u32 / k_u32 / v_u32: shared addresses of plane 2*head of the Q / K / V buffers at the window's first row.
//   bias_h: this head's bias fragments (already * log2 e);  srid: region ids of the window's 64 tokens.
__device__ __forceinline__ void attend_pair(uint32_t q_u32, uint32_t k_u32, uint32_t v_u32, uint8_t* q_ptr,
                                            const bf16* bias_h, const int* srid, bool masked, int lane, int mh) {
    const int g = lane >> 2, t = lane & 3;
    // A fragments: matrices (rows 0-7, kg0) (rows 8-15, kg0) (rows 0-7, kg1) (rows 8-15, kg1)
    const uint32_t a_off = (uint32_t)((lane >> 4) * PP + ((lane & 7) + ((lane >> 3) & 1) * 8) * 16);
    uint32_t qa[2][4];
    ldsm4(qa[0], q_u32 + (32 * mh) * 16 + a_off);
    ldsm4(qa[1], q_u32 + (32 * mh + 16) * 16 + a_off);
    float s[2][8][4];
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        const int mt = 2 * mh + m;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint2 raw = *(const uint2*)(bias_h + ((mt * 8 + j) * 32 + lane) * 4);
            const float2 lo = unpackbf(raw.x), hi = unpackbf(raw.y);
            s[m][j][0] = lo.x; s[m][j][1] = lo.y; s[m][j][2] = hi.x; s[m][j][3] = hi.y;
        }
    }
    // B fragments of two key n-tiles: matrices (keys 0-7, kg0) (keys 0-7, kg1) (keys 8-15, kg0) (keys 8-15, kg1)
    const uint32_t b_off = (uint32_t)(((lane >> 3) & 1) * PP + ((lane & 7) + (lane >> 4) * 8) * 16);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        uint32_t kb[4];
        ldsm4(kb, k_u32 + (8 * j) * 16 + b_off);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            mma16816(s[m][j], qa[m], kb[0], kb[1]);
            mma16816(s[m][j + 1], qa[m], kb[2], kb[3]);
        }
    }
    float mx[2][2];
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        const int mt = 2 * mh + m;
        if (masked) {  // shift mask (-100 across region borders): only windows touching the last window row / column
            const int rid0 = srid[16 * mt + g], rid1 = srid[16 * mt + g + 8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int rc0 = srid[8 * j + 2 * t], rc1 = srid[8 * j + 2 * t + 1];
                if (rc0 != rid0) s[m][j][0] -= 100.f * kLog2e;
                if (rc1 != rid0) s[m][j][1] -= 100.f * kLog2e;
                if (rc0 != rid1) s[m][j][2] -= 100.f * kLog2e;
                if (rc1 != rid1) s[m][j][3] -= 100.f * kLog2e;
            }
        }
        mx[m][0] = mx[m][1] = -1e30f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            mx[m][0] = fmaxf(mx[m][0], fmaxf(s[m][j][0], s[m][j][1]));
            mx[m][1] = fmaxf(mx[m][1], fmaxf(s[m][j][2], s[m][j][3]));
        }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            mx[m][e] = fmaxf(mx[m][e], __shfl_xor_sync(0xffffffffu, mx[m][e], 1));
            mx[m][e] = fmaxf(mx[m][e], __shfl_xor_sync(0xffffffffu, mx[m][e], 2));
        }
    }
    // P = exp2(s - max) as packed bf16 pairs: one MUFU.EX2 per two logits (ex2.approx.ftz.bf16x2) and the result is
    // already the A fragment of the PV MMA.  The row sums come out of the tensor core too: a third n-tile whose B
    // fragment is all ones accumulates sum_k P[row][k] (c0 -> row g, c2 -> row g + 8), from the same bf16 P that
    // multiplies V.
    float o[2][3][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int nt = 0; nt < 3; ++nt) o[m][nt][0] = o[m][nt][1] = o[m][nt][2] = o[m][nt][3] = 0.f;
    // V^T fragments through ldmatrix.trans: matrices (keys 0-7, dims 0-7) (keys 8-15, dims 0-7) (keys 0-7, dims 8-15)
    // (keys 8-15, dims 8-15) of the 16 keys of k-tile kt
    const uint32_t v_off = (uint32_t)((lane >> 4) * PP + ((lane & 7) + ((lane >> 3) & 1) * 8) * 16);
    constexpr uint32_t kOnes = 0x3F803F80u;  // {1.0, 1.0} bf16
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
        uint32_t vb[4];
        ldsm4_trans(vb, v_u32 + (16 * kt) * 16 + v_off);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            uint32_t p[4];
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                const int j = 2 * kt + jj;
                p[2 * jj] = ex2_bf16x2(packbf(s[m][j][0] - mx[m][0], s[m][j][1] - mx[m][0]));
                p[2 * jj + 1] = ex2_bf16x2(packbf(s[m][j][2] - mx[m][1], s[m][j][3] - mx[m][1]));
            }
            mma16816(o[m][0], p, vb[0], vb[1]);
            mma16816(o[m][1], p, vb[2], vb[3]);
            mma16816(o[m][2], p, kOnes, kOnes);
        }
    }
    __syncwarp();  // every lane has read its Q fragments before the rows are overwritten
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        const float i0 = __fdividef(1.f, o[m][2][0]), i1 = __fdividef(1.f, o[m][2][2]);
        const int r0 = 32 * mh + 16 * m + g, r1 = r0 + 8;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {  // head dims 8nt + 2t, +1  -> plane (2h + nt), byte 4t of the row's 16 B
            *(uint32_t*)(q_ptr + nt * PP + r0 * 16 + 4 * t) = packbf(o[m][nt][0] * i0, o[m][nt][1] * i0);
            *(uint32_t*)(q_ptr + nt * PP + r1 * 16 + 4 * t) = packbf(o[m][nt][2] * i1, o[m][nt][3] * i1);
        }
    }
}

__global__ void __launch_bounds__(kHabThreads, 1)
hab_tc_kernel(const HabArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sWq = smem + OFF_WQ;
    uint8_t* sWp = smem + OFF_WP;
    uint8_t* sA1 = smem + OFF_A1;
    uint8_t* sQ = smem + OFF_Q;
    uint8_t* sK = smem + OFF_K;
    uint8_t* sV = smem + OFF_V;
    bf16* sBias = (bf16*)(smem + OFF_BIAS);
    float* sbq = (float*)(smem + OFF_PAR);
    float* sbp = sbq + 288;
    float* sg = sbp + 96;
    float* sbeta = sg + 96;
    float2* sEx = (float2*)(smem + OFF_EX);
    int* sRid = (int*)(smem + OFF_RID);
    uint64_t* bars = (uint64_t*)(smem + OFF_BAR);
    uint64_t* bar_a1 = bars + 0;   // A1 written (count kComputeThreads)
    uint64_t* bar_qkv = bars + 1;  // [3] q / k / v part accumulated (tcgen05.commit)
    uint64_t* bar_o = bars + 4;    // attention output written to the Q planes (count kComputeThreads)
    uint64_t* bar_t = bars + 5;    // proj accumulated (tcgen05.commit)
    uint32_t* tmem_slot = (uint32_t*)(bars + 6);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = a.H, W = a.W, shift = a.shift;
    const int nwx = W / 8, nwy = H / 8;

    if (tid == 0) {
        mbar_init(bar_a1, kComputeThreads);
        for (int i = 0; i < 3; ++i) mbar_init(bar_qkv + i, 1);
        mbar_init(bar_o, kComputeThreads);
        mbar_init(bar_t, 1);
        fence_barrier_init();
    }
    if (warp == 0) { tmem_alloc(tmem_slot, HAB_TMEM_COLS); tmem_relinquish(); }
    // constants of the launch
    for (int i = tid; i < 12 * 288; i += kHabThreads) {
        const int kg = i / 288, n = i % 288;
        *(uint4*)(sWq + kg * WQ_PITCH + n * 16) = __ldg((const uint4*)a.wqkv + n * 12 + kg);
    }
    for (int i = tid; i < 12 * 96; i += kHabThreads) {
        const int kg = i / 96, n = i % 96;
        *(uint4*)(sWp + kg * WP_PITCH + n * 16) = __ldg((const uint4*)a.wproj + n * 12 + kg);
    }
    for (int i = tid; i < BIAS_ELEMS / 8; i += kHabThreads) ((uint4*)sBias)[i] = __ldg((const uint4*)a.bias_frag + i);
    if (tid < 288) sbq[tid] = a.bqkv[tid];
    if (tid < 96) { sbp[tid] = a.bproj[tid]; sg[tid] = a.ln_g[tid]; sbeta[tid] = a.ln_b[tid]; }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) pdl_launch_dependents();
    pdl_wait();  // x comes from the previous kernel in the stream

    const int first = blockIdx.x, step = gridDim.x;
    const int n_my = (a.ntiles > first) ? (a.ntiles - first + step - 1) / step : 0;

    {
        // ===================== all 12 warps (warp 0 also issues the tcgen05 MMAs) =====================
        const int q = warp & 3, c3 = warp >> 2;
        const int r = q * 32 + lane;  // token row of the tile == TMEM lane;  window r >> 6, token r & 63
        const uint32_t tm_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const int un_w = warp / 6, un_h = warp % 6;  // attention unit of this warp: (window of the pair, head)
        // Global I/O mapping ("G"): 6 lanes per token, lane j moves the 16-byte units j and j + 6 -> every request
        // covers 96 contiguous bytes per token (full 32-byte sectors); 2 passes of 384 threads = 128 tokens x 6.
        int g_row[2], g_j[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) { g_row[p] = (p * kComputeThreads + tid) / 6; g_j[p] = (p * kComputeThreads + tid) % 6; }

        // pixel index (cyclic shift = index math) and shift-mask region id of tile row `row`
        auto token_pix = [&](int tile, int row, int& rid) -> size_t {
            const int win = 2 * tile + (row >> 6), n = row & 63;
            const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
            const int ys = wy * 8 + (n >> 3), xs = wx * 8 + (n & 7);  // rolled-image coordinates
            int y = ys + shift, x = xs + shift;                       // torch.roll(-shift): rolled[i] = orig[i + shift]
            if (y >= H) y -= H;
            if (x >= W) x -= W;
            const int rh = ys < H - 8 ? 0 : (ys < H - 4 ? 1 : 2);
            const int rw = xs < W - 8 ? 0 : (xs < W - 4 ? 1 : 2);
            rid = rh * 3 + rw;
            return ((size_t)b * H + y) * W + x;
        };

        uint4 xc[4];        // this tile's 32 channels of x of row r (bf16): LayerNorm input and shortcut
        uint4 xg[2][2];     // next tile's x in the G mapping (in flight from HBM)
        size_t gp_c[2] = {0, 0}, gp_n[2] = {0, 0};  // pixel index of this thread's G rows: current / next tile
        int grid_n[2] = {0, 0};
        const uint32_t idesc = umma_idesc_bf16_f32(128, 96);
        const uint32_t a1_u = smem_u32(sA1), q_u = smem_u32(sQ), wq_u = smem_u32(sWq), wp_u = smem_u32(sWp);
#ifdef SRB_HAB_PHASES
        long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#define SRB_PHASE(k) do { if (a.dbg && tid == 0) { const long long tn = clock64(); ph[k] += tn - tprev; tprev = tn; } } while (0)
#else
#define SRB_PHASE(k) do { } while (0)
#endif
        auto load_x = [&](int tile) {  // HBM -> registers, G mapping
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                gp_n[p] = token_pix(tile, g_row[p], grid_n[p]);
                const uint4* xp = (const uint4*)(a.x + gp_n[p] * 96) + g_j[p];
                xg[p][0] = __ldg(xp);
                xg[p][1] = __ldg(xp + 6);
            }
        };
        auto park_x = [&]() {  // registers -> V planes (free between attention(i) and the qkv epilogue of tile i+1)
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                *(uint4*)(sV + g_j[p] * PP + g_row[p] * 16) = xg[p][0];
                *(uint4*)(sV + (g_j[p] + 6) * PP + g_row[p] * 16) = xg[p][1];
                if (g_j[p] == 0) sRid[g_row[p]] = grid_n[p];
                gp_c[p] = gp_n[p];
            }
        };
        auto copy_out = [&](const uint8_t* planes, bf16* dst, const size_t (&gp)[2]) {  // planes -> HBM, G mapping
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                uint4* op = (uint4*)(dst + gp[p] * 96) + g_j[p];
                op[0] = *(const uint4*)(planes + g_j[p] * PP + g_row[p] * 16);
                op[6] = *(const uint4*)(planes + (g_j[p] + 6) * PP + g_row[p] * 16);
            }
        };

        if (n_my > 0) {
            load_x(first);
            park_x();
            if (n_my > 1) load_x(first + step);
        }
        named_bar_sync(1, kComputeThreads);
        for (int i = 0; i < n_my; ++i) {
            const int tile = first + i * step;
            // ---- LayerNorm: local two-pass statistics over 32 channels, Chan merge of the three thirds ----
#pragma unroll
            for (int u = 0; u < 4; ++u) xc[u] = *(const uint4*)(sV + (4 * c3 + u) * PP + r * 16);
            float v[32];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t* w = (const uint32_t*)&xc[u];
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
            sEx[c3 * 128 + r] = make_float2(mloc, m2);
            named_bar_sync(1, kComputeThreads);
            const float2 e0 = sEx[r], e1 = sEx[128 + r], e2 = sEx[256 + r];
            const float mean = (e0.x + e1.x + e2.x) * (1.f / 3.f);
            const float d0 = e0.x - mean, d1 = e1.x - mean, d2 = e2.x - mean;
            const float var = (e0.y + e1.y + e2.y + 32.f * (d0 * d0 + d1 * d1 + d2 * d2)) * (1.f / 96.f);
            const float rs = rsqrtf(var + 1e-5f);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4 g0 = *(const float4*)(sg + 32 * c3 + 8 * u), g1 = *(const float4*)(sg + 32 * c3 + 8 * u + 4);
                const float4 b0 = *(const float4*)(sbeta + 32 * c3 + 8 * u), b1 = *(const float4*)(sbeta + 32 * c3 + 8 * u + 4);
                const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                uint4 o4;
                uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ow[k] = packbf(fmaf((v[8 * u + 2 * k] - mean) * rs, gg[2 * k], bb[2 * k]),
                                   fmaf((v[8 * u + 2 * k + 1] - mean) * rs, gg[2 * k + 1], bb[2 * k + 1]));
                *(uint4*)(sA1 + (4 * c3 + u) * PP + r * 16) = o4;
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(bar_a1);
            SRB_PHASE(0);
            mbar_wait(bar_a1, i & 1);  // every thread: A1 (= xn) is complete
            if (warp == 0) {           // qkv GEMM: q | k | v parts, one commit each
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int part = 0; part < 3; ++part) {
#pragma unroll
                        for (int ks = 0; ks < 6; ++ks)
                            umma_bf16_ss(tmem_base + COL_QKV + 96 * part,
                                         umma_smem_desc(a1_u + ks * 2 * PP, PP, 128, UMMA_SW_NONE),
                                         umma_smem_desc(wq_u + part * 96 * 16 + ks * 2 * WQ_PITCH, WQ_PITCH, 128, UMMA_SW_NONE),
                                         idesc, ks > 0);
                        umma_commit(bar_qkv + part);
                    }
                }
                __syncwarp();
            }
            // xn -> HBM for the CAB branch while the tensor core works (A1 is only read from here on)
            copy_out(sA1, a.xn, gp_c);
            SRB_PHASE(1);
            // ---- qkv epilogue: this warp's 32 rows x the 96 columns of part c3 -> Q / K / V planes ----
            mbar_wait(bar_qkv + c3, i & 1);
            tc_fence_after();
            {
                uint8_t* dst = (c3 == 0 ? sQ : c3 == 1 ? sK : sV) + r * 16;
                const float scale = (c3 == 0) ? 0.25f * kLog2e : 1.f;
                uint32_t acc[2][32];
                tmem_ld_32x32b_x32(tm_lane + COL_QKV + 96 * c3, acc[0]);
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    tmem_ld_wait();  // chunk j is in registers; fetch chunk j+1 while it is converted
                    if (j < 2) tmem_ld_32x32b_x32(tm_lane + COL_QKV + 96 * c3 + 32 * (j + 1), acc[(j + 1) & 1]);
                    const float* bp = sbq + 96 * c3 + 32 * j;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const float4 b0 = *(const float4*)(bp + 8 * u), b1 = *(const float4*)(bp + 8 * u + 4);
                        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                        uint4 o4;
                        uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            ow[k] = packbf((__uint_as_float(acc[j & 1][8 * u + 2 * k]) + bb[2 * k]) * scale,
                                           (__uint_as_float(acc[j & 1][8 * u + 2 * k + 1]) + bb[2 * k + 1]) * scale);
                        *(uint4*)(dst + (4 * j + u) * PP) = o4;
                    }
                }
            }
            tc_fence_before();
            SRB_PHASE(2);
            named_bar_sync(1, kComputeThreads);  // Q / K / V planes complete
            SRB_PHASE(3);
            // ---- attention: warp = (window un_w, head un_h), 64 query rows in two 32-row passes ----
            {
                const int win = 2 * tile + un_w;
                const int wxx = win % nwx, wyy = (win / nwx) % nwy;
                const bool masked = shift > 0 && (wxx == nwx - 1 || wyy == nwy - 1);
                const uint32_t row0 = (uint32_t)(64 * un_w * 16 + 2 * un_h * PP);
                const bf16* bias_h = sBias + un_h * (4 * 8 * 32 * 4);
#pragma unroll 1
                for (int mh = 0; mh < 2; ++mh)
                    attend_pair(smem_u32(sQ) + row0, smem_u32(sK) + row0, smem_u32(sV) + row0, sQ + row0, bias_h,
                                sRid + 64 * un_w, masked, lane, mh);
            }
            fence_proxy_async_smem();
            mbar_arrive(bar_o);
            SRB_PHASE(4);
            mbar_wait(bar_o, i & 1);  // every warp has left the attention: K / V planes and sRid are free
            if (warp == 0) {          // proj GEMM
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < 6; ++ks)
                        umma_bf16_ss(tmem_base + COL_T, umma_smem_desc(q_u + ks * 2 * PP, PP, 128, UMMA_SW_NONE),
                                     umma_smem_desc(wp_u + ks * 2 * WP_PITCH, WP_PITCH, 128, UMMA_SW_NONE), idesc, ks > 0);
                    umma_commit(bar_t);
                }
                __syncwarp();
            }
            // while the proj GEMM runs: x(i+1) -> V planes (free since every warp has left the attention)
            size_t gp_t[2] = {gp_c[0], gp_c[1]};  // this tile's pixel indices for the t copy-out below
            if (i + 1 < n_my) {
                park_x();
                if (i + 2 < n_my) load_x(tile + 2 * step);
            }
            SRB_PHASE(5);
            // ---- proj epilogue: t = x + attn @ Wproj^T + b : this thread's row x channels [32 c3, +32) -> K planes ----
            mbar_wait(bar_t, i & 1);
            tc_fence_after();
            {
                uint32_t acc[32];
                tmem_ld_32x32b_x32(tm_lane + COL_T + 32 * c3, acc);
                tmem_ld_wait();
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float4 b0 = *(const float4*)(sbp + 32 * c3 + 8 * u), b1 = *(const float4*)(sbp + 32 * c3 + 8 * u + 4);
                    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                    const uint32_t* xw = (const uint32_t*)&xc[u];
                    uint4 o4;
                    uint32_t* ow = (uint32_t*)&o4;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float2 xf = unpackbf(xw[k]);
                        ow[k] = packbf(__uint_as_float(acc[8 * u + 2 * k]) + bb[2 * k] + xf.x,
                                       __uint_as_float(acc[8 * u + 2 * k + 1]) + bb[2 * k + 1] + xf.y);
                    }
                    *(uint4*)(sK + (4 * c3 + u) * PP + r * 16) = o4;
                }
            }
            tc_fence_before();
            named_bar_sync(1, kComputeThreads);  // t staged in the K planes
            copy_out(sK, a.t, gp_t);
            named_bar_sync(1, kComputeThreads);  // x(i+1) parked; K planes read out
            SRB_PHASE(6);
        }
#ifdef SRB_HAB_PHASES
        if (a.dbg && tid == 0)
            for (int k = 0; k < 8; ++k) a.dbg[blockIdx.x * 8 + k] = ph[k];
#endif
#undef SRB_PHASE
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, HAB_TMEM_COLS);
}

}  // namespace

extern long long* g_fa_dbg;  // attn.cu (srb_debug_set_phase_buffer)

int hab_attention_tc(const bf16* x, const float* ln_g, const float* ln_b, const bf16* wqkv, const float* bqkv,
                     const bf16* bias_frag_log2e, const bf16* wproj, const float* bproj, bf16* xn, bf16* t, int B, int H,
                     int W, int shift, cudaStream_t s) {
    if (H % 8 || W % 16) { set_error("hab_attention: H=%d must be a multiple of 8 and W=%d of 16", H, W); return SRB_E_SHAPE; }
    static std::atomic<unsigned long long> attr{0};
    SRB_TRY(opt_in_dynamic_smem(hab_tc_kernel, HAB_TC_SMEM, attr));
    HabArgs a;
    a.x = x; a.ln_g = ln_g; a.ln_b = ln_b; a.wqkv = wqkv; a.bqkv = bqkv; a.bias_frag = bias_frag_log2e; a.wproj = wproj;
    a.bproj = bproj; a.xn = xn; a.t = t; a.B = B; a.H = H; a.W = W; a.shift = shift;
    a.ntiles = B * (H / 8) * (W / 8) / 2;
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
    SRB_CUDA(cudaLaunchKernelEx(&cfg, hab_tc_kernel, a));
    return SRB_OK;
}

}  // namespace srb
