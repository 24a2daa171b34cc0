// (Shifted-)window multi-head self-attention core of HAB from a saved qkv tensor, forward AND backward — the un-fused
// operator pair of the stage-2 training path (upstream hat_arch.py WindowAttention.forward and its
// autograd, reached from src/architecture.py:78).  The inference path keeps the fused tcgen05 kernel (hab_tc.cu).
//
//   window = 8 x 8 tokens, 6 heads x 16 channels, logits = q.k * 0.25 + bias[h][i][j] (+ -100 across shift-mask regions),
//   softmax over the 64 keys, out = P v.  torch.roll / window_partition / window_reverse are index arithmetic.
//
// One WARP per (window, head), one warp per CTA, persistent over items; every contraction is a warp-level
// mma.sync.m16n8k16 (head_dim 16 = one k-step) on ldmatrix fragments of the head's 32-byte q / k / v / dO slices staged by
// cp.async (rows padded to an odd number of 16-byte chunks: conflict-free).  No CTA-wide barrier exists; 8 (backward) to
// 16 (forward) independent warps per SM overlap one another's load / MMA / store phases.
//   forward : S = Q K^T (+bias, mask) -> softmax in registers -> O = P V with P taken straight from the accumulators.
//   backward: per 16-row block: S, P recomputed; dP = dO V^T; dS = P (dP - rowsum(P dP)); dQ = dS K / 4;
//             dK += dS^T Q / 4 and dV += P^T dO with the transposed A fragments produced by movmatrix (register
//             accumulators over the 4 row blocks); the relative-position-bias gradient accumulates dS in a per-warp
//             fp32 [64][64] shared-memory tile across all windows of the warp (each element owned by one lane: plain
//             read-add-write), folded through the index buffer once at the end (one global atomicAdd per table entry).
#include <cuda_bf16.h>

#include "attn_mma.cuh"
#include "common.h"
#include "ops.h"

namespace srb {

namespace {

using namespace amma;

constexpr int RP = 56;  // bf16 pitch of a staged token row: q16 | k16 | v16 | 8 pad = 112 B = 7 chunks (odd)
constexpr int GP = 24;  // bf16 pitch of a staged dO row: 16 | 8 pad = 48 B = 3 chunks (odd)
constexpr int BP = 72;  // bf16 pitch of the CTA's [64][64] bias tile: word address 36 r + 4 j + t -> bank 4 g + t
constexpr int BIAS_TILE = 64 * BP * 2;
constexpr int FWD_WARPS = 4, BWD_WARPS = 8;  // warps per CTA, all on one head (they share its bias tile)

struct WinArgs {
    const bf16* qkv;    // [B,H,W,288]
    const float* bias;  // [6][64][64] relative position bias per head (already gathered from the table)
    const int* index;   // [64][64] relative_position_index (backward only)
    const bf16* dout;   // [B,H,W,96]   (backward)
    bf16* out;          // forward: [B,H,W,96];  backward: dqkv [B,H,W,288]
    float* dtable;      // [225][6] (backward, accumulated with atomicAdd)
    int B, H, W, shift, nwin;
};

// pixel index of window token n (cyclic shift = index math) and its shift-mask region id
__device__ __forceinline__ int win_token(const WinArgs& a, int win, int n, int& rid) {
    const int nwx = a.W / 8, nwy = a.H / 8;
    const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
    const int ys = wy * 8 + (n >> 3), xs = wx * 8 + (n & 7);  // rolled-image coordinates
    int y = ys + a.shift, x = xs + a.shift;                   // torch.roll(-shift): rolled[i] = orig[i + shift]
    if (y >= a.H) y -= a.H;
    if (x >= a.W) x -= a.W;
    const int rh = ys < a.H - 8 ? 0 : (ys < a.H - 4 ? 1 : 2);
    const int rw = xs < a.W - 8 ? 0 : (xs < a.W - 4 ? 1 : 2);
    rid = a.shift > 0 ? rh * 3 + rw : 0;
    return (b * a.H + y) * a.W + x;  // host checks B*H*W < 2^31
}

// stage this (window, head): sT rows = q|k|v slices, optional sG rows = dO slice; returns true when the window crosses
// shift-mask regions
template <bool BWD>
__device__ __forceinline__ bool stage(const WinArgs& a, int win, int h, bf16* sT, bf16* sG, int* sPix, uint8_t* sRid,
                                      int lane) {
    int r0, r1;
    const int p0 = win_token(a, win, lane, r0), p1 = win_token(a, win, lane + 32, r1);
    sPix[lane] = p0; sPix[lane + 32] = p1;
    sRid[lane] = (uint8_t)r0; sRid[lane + 32] = (uint8_t)r1;
    const int first = __shfl_sync(0xffffffffu, r0, 0);
    const bool mixed = __any_sync(0xffffffffu, r0 != first || r1 != first);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        const int i = lane + 32 * k, n = i / 6, c = i % 6;
        cp16(sT + n * RP + c * 8, a.qkv + (size_t)sPix[n] * 288 + (c >> 1) * 96 + h * 16 + (c & 1) * 8);
    }
    if (BWD) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = lane + 32 * k, n = i >> 1, c = i & 1;
            cp16(sG + n * GP + c * 8, a.dout + (size_t)sPix[n] * 96 + h * 16 + c * 8);
        }
    }
    cp_commit_wait();
    __syncwarp();
    return mixed;
}

// softmax probabilities of the 16 query rows of block mt (fp32, accumulator layout: s[j][0..1] row g, s[j][2..3] row g+8)
__device__ __forceinline__ void probs(const bf16* sB, const bf16* sT, const uint8_t* sRid, bool mixed, int mt,
                                      int lane, float (&s)[8][4]) {
    const int g = lane >> 2, t = lane & 3;
    // bias of this CTA's head from its shared-memory tile (bf16, pitch BP: the 32 lanes of a request hit 32 banks).
    // Read from global memory these were 64 float2 loads per item whose 8 row segments are 8 L1 wavefronts each -
    // two thirds of the kernel's LSU time.
    const int r0 = 16 * mt + g, r1 = r0 + 8;
    const bf16* b0 = sB + r0 * BP + 2 * t;
    const bf16* b1 = b0 + 8 * BP;
    float2 bu[8], bv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        bu[j] = __bfloat1622float2(*(const __nv_bfloat162*)(b0 + 8 * j));
        bv[j] = __bfloat1622float2(*(const __nv_bfloat162*)(b1 + 8 * j));
    }
    uint32_t qa[4];
    frag_a(qa, sT, RP, 16 * mt, 0, lane);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        uint32_t kb[4];
        frag_b_nk(kb, sT, RP, 8 * j, 16, lane);
        mma16816(s[j], qa[0], qa[1], qa[2], qa[3], kb[0], kb[1]);
        mma16816(s[j + 1], qa[0], qa[1], qa[2], qa[3], kb[2], kb[3]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        s[j][0] = fmaf(s[j][0], 0.25f, bu[j].x); s[j][1] = fmaf(s[j][1], 0.25f, bu[j].y);
        s[j][2] = fmaf(s[j][2], 0.25f, bv[j].x); s[j][3] = fmaf(s[j][3], 0.25f, bv[j].y);
    }
    if (mixed) {
        const int q0 = sRid[r0], q1 = sRid[r1];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c0 = sRid[8 * j + 2 * t], c1 = sRid[8 * j + 2 * t + 1];
            if (c0 != q0) s[j][0] -= 100.f;
            if (c1 != q0) s[j][1] -= 100.f;
            if (c0 != q1) s[j][2] -= 100.f;
            if (c1 != q1) s[j][3] -= 100.f;
        }
    }
    float m0 = -1e30f, m1 = -1e30f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
        m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
    }
    m0 = quad_max(m0) * LOG2E; m1 = quad_max(m1) * LOG2E;
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        s[j][0] = ex2f(fmaf(s[j][0], LOG2E, -m0)); s[j][1] = ex2f(fmaf(s[j][1], LOG2E, -m0));
        s[j][2] = ex2f(fmaf(s[j][2], LOG2E, -m1)); s[j][3] = ex2f(fmaf(s[j][3], LOG2E, -m1));
        l0 += s[j][0] + s[j][1];
        l1 += s[j][2] + s[j][3];
    }
    const float i0 = 1.f / quad_sum(l0), i1 = 1.f / quad_sum(l1);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j][0] *= i0; s[j][1] *= i0; s[j][2] *= i1; s[j][3] *= i1; }
}

constexpr int FWD_SMEM = 64 * RP * 2 + 64 * 4 + 64;  // per warp

// CTA prologue: head of this CTA, its bf16 bias tile in shared memory (the only CTA-wide barrier of the kernels)
__device__ __forceinline__ void load_bias_tile(const WinArgs& a, int h, bf16* sB) {
    const float4* src = (const float4*)(a.bias + (size_t)h * 4096);
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
        const float4 v = __ldg(src + i);
        const int r = i >> 4, c = (i & 15) * 4;
        *(__nv_bfloat162*)(sB + r * BP + c) = __floats2bfloat162_rn(v.x, v.y);
        *(__nv_bfloat162*)(sB + r * BP + c + 2) = __floats2bfloat162_rn(v.z, v.w);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(32 * FWD_WARPS, 4)
win_attn_fwd_kernel(const WinArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    bf16* sB = (bf16*)smem;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    uint8_t* mine = smem + BIAS_TILE + warp * FWD_SMEM;
    bf16* sT = (bf16*)mine;
    int* sPix = (int*)(mine + 64 * RP * 2);
    uint8_t* sRid = (uint8_t*)(sPix + 64);
    const int h = blockIdx.x % 6, nh = (gridDim.x - h + 5) / 6;  // this CTA's head, CTAs on that head
    load_bias_tile(a, h, sB);
    for (int win = (blockIdx.x / 6) * FWD_WARPS + warp; win < a.nwin; win += nh * FWD_WARPS) {
        const bool mixed = stage<false>(a, win, h, sT, nullptr, sPix, sRid, lane);
#pragma unroll 1
        for (int mt = 0; mt < 4; ++mt) {
            float s[8][4];
            probs(sB, sT, sRid, mixed, mt, lane, s);
            float o[2][4];
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
            for (int kt = 0; kt < 4; ++kt) {
                uint32_t vb[4];
                frag_b_kn(vb, sT, RP, 16 * kt, 32, lane);
                const uint32_t p0 = pack2(s[2 * kt][0], s[2 * kt][1]), p1 = pack2(s[2 * kt][2], s[2 * kt][3]);
                const uint32_t p2 = pack2(s[2 * kt + 1][0], s[2 * kt + 1][1]), p3 = pack2(s[2 * kt + 1][2], s[2 * kt + 1][3]);
                mma16816(o[0], p0, p1, p2, p3, vb[0], vb[1]);
                mma16816(o[1], p0, p1, p2, p3, vb[2], vb[3]);
            }
            __syncwarp();  // every lane holds its Q fragments of this block before the rows are overwritten
            const int r0 = 16 * mt + g, r1 = r0 + 8;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                *(uint32_t*)(sT + r0 * RP + 8 * nt + 2 * t) = pack2(o[nt][0], o[nt][1]);
                *(uint32_t*)(sT + r1 * RP + 8 * nt + 2 * t) = pack2(o[nt][2], o[nt][3]);
            }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = lane + 32 * k, n = i >> 1, c = i & 1;
            *(uint4*)(a.out + (size_t)sPix[n] * 96 + h * 16 + c * 8) = *(const uint4*)(sT + n * RP + c * 8);
        }
        __syncwarp();  // the stage of the next item overwrites sT / sPix
    }
}

constexpr int BWD_SMEM = 64 * RP * 2 + 64 * GP * 2 + 64 * 64 * 4 + 64 * 4 + 64;  // per warp

__global__ void __launch_bounds__(32 * BWD_WARPS, 1)
win_attn_bwd_kernel(const WinArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    bf16* sB = (bf16*)smem;
    const int warp = threadIdx.x >> 5;
    uint8_t* mine = smem + BIAS_TILE + warp * BWD_SMEM;
    bf16* sT = (bf16*)mine;
    bf16* sG = sT + 64 * RP;
    float* sAcc = (float*)(sG + 64 * GP);            // [64][64] sum over this warp's windows of dS (one head)
    int* sPix = (int*)(sAcc + 64 * 64);
    uint8_t* sRid = (uint8_t*)(sPix + 64);
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    // a CTA (and so each of its warps) keeps ONE head for all its items: every sAcc tile is a single-head sum
    const int h = blockIdx.x % 6, nh = (gridDim.x - h + 5) / 6;
    load_bias_tile(a, h, sB);
    for (int i = lane; i < 64 * 64; i += 32) sAcc[i] = 0.f;
    __syncwarp();
    for (int win = (blockIdx.x / 6) * BWD_WARPS + warp; win < a.nwin; win += nh * BWD_WARPS) {
        const bool mixed = stage<true>(a, win, h, sT, sG, sPix, sRid, lane);
        float dk[4][2][4], dv[4][2][4];
#pragma unroll
        for (int kb = 0; kb < 4; ++kb)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int e = 0; e < 4; ++e) dk[kb][nt][e] = dv[kb][nt][e] = 0.f;
#pragma unroll 1
        for (int mt = 0; mt < 4; ++mt) {
            float p[8][4], ds[8][4];
            probs(sB, sT, sRid, mixed, mt, lane, p);
            // dP = dO V^T
            uint32_t ga[4];
            frag_a(ga, sG, GP, 16 * mt, 0, lane);
#pragma unroll
            for (int j = 0; j < 8; ++j) ds[j][0] = ds[j][1] = ds[j][2] = ds[j][3] = 0.f;
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
                uint32_t vb[4];
                frag_b_nk(vb, sT, RP, 8 * j, 32, lane);
                mma16816(ds[j], ga[0], ga[1], ga[2], ga[3], vb[0], vb[1]);
                mma16816(ds[j + 1], ga[0], ga[1], ga[2], ga[3], vb[2], vb[3]);
            }
            float d0 = 0.f, d1 = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                d0 = fmaf(p[j][0], ds[j][0], fmaf(p[j][1], ds[j][1], d0));
                d1 = fmaf(p[j][2], ds[j][2], fmaf(p[j][3], ds[j][3], d1));
            }
            d0 = quad_sum(d0); d1 = quad_sum(d1);
            const int r0 = 16 * mt + g, r1 = r0 + 8;
            uint32_t pp[8][2], sp[8][2];  // bf16 P and dS, accumulator-layout 8x8 blocks: [j][0] rows 0-7, [j][1] rows 8-15
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                ds[j][0] = p[j][0] * (ds[j][0] - d0); ds[j][1] = p[j][1] * (ds[j][1] - d0);
                ds[j][2] = p[j][2] * (ds[j][2] - d1); ds[j][3] = p[j][3] * (ds[j][3] - d1);
                // 8-float column groups XOR-swizzled by the row (r0 & 7 == r1 & 7 == g): the 8 rows of a request hit 8 groups
                float2* q0 = (float2*)(sAcc + r0 * 64 + 8 * (j ^ g) + 2 * t);
                float2* q1 = (float2*)(sAcc + r1 * 64 + 8 * (j ^ g) + 2 * t);
                float2 u = *q0, v = *q1;
                u.x += ds[j][0]; u.y += ds[j][1]; v.x += ds[j][2]; v.y += ds[j][3];
                *q0 = u; *q1 = v;
                pp[j][0] = pack2(p[j][0], p[j][1]); pp[j][1] = pack2(p[j][2], p[j][3]);
                sp[j][0] = pack2(ds[j][0], ds[j][1]); sp[j][1] = pack2(ds[j][2], ds[j][3]);
            }
            // dQ = dS K / 4   (k = keys: 4 k-steps; B[k=key][n=d] = K rows through ldmatrix.trans)
            float dq[2][4];
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) dq[nt][0] = dq[nt][1] = dq[nt][2] = dq[nt][3] = 0.f;
#pragma unroll
            for (int kt = 0; kt < 4; ++kt) {
                uint32_t kb[4];
                frag_b_kn(kb, sT, RP, 16 * kt, 16, lane);
                mma16816(dq[0], sp[2 * kt][0], sp[2 * kt][1], sp[2 * kt + 1][0], sp[2 * kt + 1][1], kb[0], kb[1]);
                mma16816(dq[1], sp[2 * kt][0], sp[2 * kt][1], sp[2 * kt + 1][0], sp[2 * kt + 1][1], kb[2], kb[3]);
            }
            // dK += dS^T Q / 4, dV += P^T dO   (k = the 16 query rows of this block; A = transposed blocks via movmatrix)
            uint32_t qb[4], gb[4];
            frag_b_kn(qb, sT, RP, 16 * mt, 0, lane);
            frag_b_kn(gb, sG, GP, 16 * mt, 0, lane);
#pragma unroll
            for (int kb = 0; kb < 4; ++kb) {
                const uint32_t s0 = movt(sp[2 * kb][0]), s1 = movt(sp[2 * kb + 1][0]);
                const uint32_t s2 = movt(sp[2 * kb][1]), s3 = movt(sp[2 * kb + 1][1]);
                mma16816(dk[kb][0], s0, s1, s2, s3, qb[0], qb[1]);
                mma16816(dk[kb][1], s0, s1, s2, s3, qb[2], qb[3]);
                const uint32_t t0 = movt(pp[2 * kb][0]), t1 = movt(pp[2 * kb + 1][0]);
                const uint32_t t2 = movt(pp[2 * kb][1]), t3 = movt(pp[2 * kb + 1][1]);
                mma16816(dv[kb][0], t0, t1, t2, t3, gb[0], gb[1]);
                mma16816(dv[kb][1], t0, t1, t2, t3, gb[2], gb[3]);
            }
            __syncwarp();  // all lanes hold the Q fragments of this block: its rows now take dQ
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                *(uint32_t*)(sT + r0 * RP + 8 * nt + 2 * t) = pack2(dq[nt][0] * 0.25f, dq[nt][1] * 0.25f);
                *(uint32_t*)(sT + r1 * RP + 8 * nt + 2 * t) = pack2(dq[nt][2] * 0.25f, dq[nt][3] * 0.25f);
            }
        }
        __syncwarp();  // K / V no longer read: their slots take dK / dV
#pragma unroll
        for (int kb = 0; kb < 4; ++kb)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const int k0 = 16 * kb + g, k1 = k0 + 8, c = 8 * nt + 2 * t;
                *(uint32_t*)(sT + k0 * RP + 16 + c) = pack2(dk[kb][nt][0] * 0.25f, dk[kb][nt][1] * 0.25f);
                *(uint32_t*)(sT + k1 * RP + 16 + c) = pack2(dk[kb][nt][2] * 0.25f, dk[kb][nt][3] * 0.25f);
                *(uint32_t*)(sT + k0 * RP + 32 + c) = pack2(dv[kb][nt][0], dv[kb][nt][1]);
                *(uint32_t*)(sT + k1 * RP + 32 + c) = pack2(dv[kb][nt][2], dv[kb][nt][3]);
            }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 12; ++k) {
            const int i = lane + 32 * k, n = i / 6, c = i % 6;
            *(uint4*)(a.out + (size_t)sPix[n] * 288 + (c >> 1) * 96 + h * 16 + (c & 1) * 8) = *(const uint4*)(sT + n * RP + c * 8);
        }
        __syncwarp();
    }
    // fold the dS sums through the index buffer: table gradient [225][6].  sT is free: 225 floats of it hold the fold.
    float* sTab = (float*)sT;
    for (int i = lane; i < 225; i += 32) sTab[i] = 0.f;
    __syncwarp();
    for (int i = lane; i < 64 * 64; i += 32) {  // lanes of one step: one row, 32 consecutive columns -> distinct entries
        const int r = i >> 6, c = i & 63;
        atomicAdd(&sTab[__ldg(a.index + i)], sAcc[r * 64 + 8 * ((c >> 3) ^ (r & 7)) + (c & 7)]);
    }
    __syncwarp();
    for (int i = lane; i < 225; i += 32) atomicAdd(a.dtable + i * 6 + h, sTab[i]);
}

int check_geom(int B, int H, int W, int shift) {
    if (B <= 0 || H % 8 || W % 8 || H < 8 || W < 8 || (shift != 0 && shift != 4)) {
        set_error("win_attn: H=%d W=%d must be multiples of 8, shift=%d must be 0 or 4", H, W, shift);
        return SRB_E_SHAPE;
    }
    if ((size_t)B * H * W >= ((size_t)1 << 31) / 288) { set_error("win_attn: tensor too large for 32-bit indexing"); return SRB_E_SHAPE; }
    return SRB_OK;
}

}  // namespace

int win_attn_fwd(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, int shift, cudaStream_t s) {
    SRB_TRY(check_geom(B, H, W, shift));
    WinArgs a;
    a.qkv = qkv; a.bias = bias; a.index = nullptr; a.dout = nullptr; a.out = out; a.dtable = nullptr;
    a.B = B; a.H = H; a.W = W; a.shift = shift; a.nwin = B * (H / 8) * (W / 8);
    static std::atomic<unsigned long long> smem_set{0};
    constexpr int smem = BIAS_TILE + FWD_WARPS * FWD_SMEM;
    SRB_TRY(opt_in_dynamic_smem(win_attn_fwd_kernel, smem, smem_set));
    int grid = num_sms() * 4;  // 4 CTAs of 4 warps per SM
    const int need = ((a.nwin + FWD_WARPS - 1) / FWD_WARPS) * 6;
    if (grid > need) grid = need;
    win_attn_fwd_kernel<<<grid, 32 * FWD_WARPS, smem, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int win_attn_bwd(const bf16* qkv, const float* bias, const int* index, const bf16* dout, bf16* dqkv, float* dtable, int B,
                 int H, int W, int shift, cudaStream_t s) {
    SRB_TRY(check_geom(B, H, W, shift));
    WinArgs a;
    a.qkv = qkv; a.bias = bias; a.index = index; a.dout = dout; a.out = dqkv; a.dtable = dtable;
    a.B = B; a.H = H; a.W = W; a.shift = shift; a.nwin = B * (H / 8) * (W / 8);
    static std::atomic<unsigned long long> smem_set{0};
    constexpr int smem = BIAS_TILE + BWD_WARPS * BWD_SMEM;
    static_assert(smem <= 227 * 1024, "win_attn_bwd: shared memory");
    SRB_TRY(opt_in_dynamic_smem(win_attn_bwd_kernel, smem, smem_set));
    int grid = num_sms();  // one CTA of 8 warps per SM (255 registers per thread, 26 KB of shared memory per warp)
    const int need = ((a.nwin + BWD_WARPS - 1) / BWD_WARPS) * 6;
    if (grid > need) grid = need;
    win_attn_bwd_kernel<<<grid, 32 * BWD_WARPS, smem, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
