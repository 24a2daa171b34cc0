// Overlapping cross-attention core of OCAB from a saved qkv tensor, forward AND backward — the un-fused operator pair of the
// stage-2 training path (upstream hat_arch.py OCAB.forward and its autograd).  The inference path keeps
// the mma.sync kernel in attn.cu.
//
//   8 x 8 query windows attend to the 12 x 12 key/value window around them (nn.Unfold(kernel 12, stride 8, padding 2) of the
//   PROJECTED k/v maps: keys outside the image are zero vectors that still take part in the softmax with logit = bias).
//   6 heads x 16 channels, logits = q.k * 0.25 + bias[h][i][j], out = P v.
//
// One WARP per (window, head), one warp per CTA, persistent; all contractions on mma.sync.m16n8k16 (attn_mma.cuh), like
// win_attn.cu.  With 144 keys the probabilities of a 16-row block do not fit the register file next to the dK / dV
// accumulators, so the backward is two passes over the staged slices:
//   pass A (per 16-row block): S over all 144 keys -> row max, 1 / row sum, O = P V -> delta = rowsum(dO . O)  (the
//           flash-attention identity for rowsum(P . dP)); the three row statistics go to shared memory.
//   pass B (per chunk of 48 keys, inner loop over the 4 row blocks): P and dP of the chunk recomputed, dS = P (dP - delta),
//           dQ += dS K / 4 (register accumulators of all 4 blocks), dK / dV of the chunk accumulated over the blocks with
//           movmatrix-transposed dS / P.  The bias-table gradient accumulates dS in a per-warp fp32 [64][144] shared-memory
//           tile over all windows of the warp (rotated 8-float groups: conflict-free), folded through the index buffer
//           at the end.
// Keys belong to up to four windows: every (window, head) writes its dK / dV rows to a per-window scratch and a fold kernel
// sums them per pixel in fixed order (deterministic).
#include <cuda_bf16.h>

#include "attn_mma.cuh"
#include "common.h"
#include "ops.h"

namespace srb {

namespace {

using namespace amma;

constexpr int OA_NK = 144, OA_TAB = 361;
constexpr int QP = 24;  // bf16 pitch of a staged q / dO row: 16 | 8 pad = 48 B (3 chunks, odd)
constexpr int KP = 40;  // bf16 pitch of a staged key row: k16 | v16 | 8 pad = 80 B (5 chunks, odd)

struct OcaArgs {
    const bf16* qkv;    // [B,H,W,288]
    const float* bias;  // [6][64][144]
    const float* bias_frag;  // optional: the same values in MMA-fragment order [6][4 mt][18 j][32 lanes][4] (backward)
    const int* index;   // [64][144]  (backward)
    const bf16* dout;   // [B,H,W,96] (backward)
    bf16* out;          // forward: [B,H,W,96];  backward: dqkv [B,H,W,288] (q part written here, k/v part by the fold)
    bf16* dkv_win;      // [nwin][144][192] per-window dK | dV (backward scratch)
    float* dtable;      // [361][6]
    int B, H, W, nwin;
};

__device__ __forceinline__ void oa_unpack8(const uint4& raw, float* f) {
    const __nv_bfloat162* h = (const __nv_bfloat162*)&raw;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float2 v = __bfloat1622float2(h[k]);
        f[2 * k] = v.x;
        f[2 * k + 1] = v.y;
    }
}
__device__ __forceinline__ uint4 oa_pack8(const float* f) {
    uint4 raw;
    __nv_bfloat162* h = (__nv_bfloat162*)&raw;
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(f[2 * k], f[2 * k + 1]);
    return raw;
}

// stage (window, head): sQ rows = q slice, sKV rows = k | v slices of the 12 x 12 neighbourhood (zero outside the image),
// optional sG rows = dO slice; sPix = pixel of every query token
template <bool BWD>
__device__ __forceinline__ void oa_stage(const OcaArgs& a, int win, int h, bf16* sQ, bf16* sKV, bf16* sG, int* sPix, int lane) {
    const int nwx = a.W / 8, nwy = a.H / 8;
    const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
    const int img = b * a.H * a.W;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int n = lane + 32 * k;
        sPix[n] = img + (wy * 8 + (n >> 3)) * a.W + wx * 8 + (n & 7);
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int i = lane + 32 * k, n = i >> 1, c = i & 1;
        cp16(sQ + n * QP + c * 8, a.qkv + (size_t)sPix[n] * 288 + h * 16 + c * 8);
        if (BWD) cp16(sG + n * QP + c * 8, a.dout + (size_t)sPix[n] * 96 + h * 16 + c * 8);
    }
#pragma unroll
    for (int k = 0; k < 18; ++k) {
        const int i = lane + 32 * k, m = i >> 2, c = i & 3;  // c: k chunk 0/1, v chunk 0/1
        const int y = wy * 8 - 2 + m / 12, x = wx * 8 - 2 + m % 12;
        bf16* dst = sKV + m * KP + c * 8;
        if (y >= 0 && y < a.H && x >= 0 && x < a.W)
            cp16(dst, a.qkv + (size_t)(img + y * a.W + x) * 288 + 96 + (c >> 1) * 96 + h * 16 + (c & 1) * 8);
        else
            *(uint4*)dst = make_uint4(0, 0, 0, 0);  // nn.Unfold zero padding of the projected K / V maps
    }
    cp_commit_wait();
    __syncwarp();
}

// logits of NT n-tiles (keys key0 ...) of row block mt: s = q.k / 4 + bias
template <int NT>
__device__ __forceinline__ void oa_logits(const OcaArgs& a, const bf16* sQ, const bf16* sKV, int h, int mt, int key0, int lane,
                                          float (&s)[NT][4]) {
    const int g = lane >> 2, t = lane & 3;
    // bias rows come from L2: issue the loads before the ldmatrix + MMA sequence that covers their latency.  Row-major,
    // the two float2 a lane needs per n-tile are 8 row segments = 8 L1 wavefronts per load; in fragment order one
    // float4 per lane is 512 contiguous bytes per warp = 4 wavefronts for both.
    float2 bu[NT], bv[NT];
    if (a.bias_frag) {
        const float4* bf = (const float4*)a.bias_frag + ((size_t)(h * 4 + mt) * 18 + (key0 >> 3)) * 32 + lane;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            const float4 b4 = __ldg(bf + j * 32);
            bu[j] = make_float2(b4.x, b4.y);
            bv[j] = make_float2(b4.z, b4.w);
        }
    } else {
        const float* b0 = a.bias + ((size_t)h * 64 + 16 * mt + g) * OA_NK + key0 + 2 * t;
        const float* b1 = b0 + 8 * OA_NK;
#pragma unroll
        for (int j = 0; j < NT; ++j) { bu[j] = __ldg((const float2*)(b0 + 8 * j)); bv[j] = __ldg((const float2*)(b1 + 8 * j)); }
    }
    uint32_t qa[4];
    frag_a(qa, sQ, QP, 16 * mt, 0, lane);
#pragma unroll
    for (int j = 0; j < NT; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int j = 0; j < NT; j += 2) {
        uint32_t kb[4];
        frag_b_nk(kb, sKV, KP, key0 + 8 * j, 0, lane);
        mma16816(s[j], qa[0], qa[1], qa[2], qa[3], kb[0], kb[1]);
        mma16816(s[j + 1], qa[0], qa[1], qa[2], qa[3], kb[2], kb[3]);
    }
#pragma unroll
    for (int j = 0; j < NT; ++j) {
        s[j][0] = fmaf(s[j][0], 0.25f, bu[j].x); s[j][1] = fmaf(s[j][1], 0.25f, bu[j].y);
        s[j][2] = fmaf(s[j][2], 0.25f, bv[j].x); s[j][3] = fmaf(s[j][3], 0.25f, bv[j].y);
    }
}

// full-row pass of block mt: un-normalised probabilities in s (18 n-tiles), row maxima (log2 domain) and 1 / row sums,
// o = normalised P V
__device__ __forceinline__ void oa_rows(const OcaArgs& a, const bf16* sQ, const bf16* sKV, int h, int mt, int lane,
                                        float (&o)[2][4], float& m0, float& m1, float& i0, float& i1) {
    float s[18][4];
    oa_logits<18>(a, sQ, sKV, h, mt, 0, lane, s);
    m0 = -1e30f; m1 = -1e30f;
#pragma unroll
    for (int j = 0; j < 18; ++j) {
        m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
        m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
    }
    m0 = quad_max(m0) * LOG2E; m1 = quad_max(m1) * LOG2E;
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
    for (int kt = 0; kt < 9; ++kt) {
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int j = 2 * kt + jj;
            s[j][0] = ex2f(fmaf(s[j][0], LOG2E, -m0)); s[j][1] = ex2f(fmaf(s[j][1], LOG2E, -m0));
            s[j][2] = ex2f(fmaf(s[j][2], LOG2E, -m1)); s[j][3] = ex2f(fmaf(s[j][3], LOG2E, -m1));
            l0 += s[j][0] + s[j][1];
            l1 += s[j][2] + s[j][3];
        }
        uint32_t vb[4];
        frag_b_kn(vb, sKV, KP, 16 * kt, 16, lane);
        const uint32_t p0 = pack2(s[2 * kt][0], s[2 * kt][1]), p1 = pack2(s[2 * kt][2], s[2 * kt][3]);
        const uint32_t p2 = pack2(s[2 * kt + 1][0], s[2 * kt + 1][1]), p3 = pack2(s[2 * kt + 1][2], s[2 * kt + 1][3]);
        mma16816(o[0], p0, p1, p2, p3, vb[0], vb[1]);
        mma16816(o[1], p0, p1, p2, p3, vb[2], vb[3]);
    }
    i0 = 1.f / quad_sum(l0); i1 = 1.f / quad_sum(l1);
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) { o[nt][0] *= i0; o[nt][1] *= i0; o[nt][2] *= i1; o[nt][3] *= i1; }
}

// frag[h][mt][j][lane = 4 g + t] = (bias[h][16 mt + g][8 j + 2 t], [..][8 j + 2 t + 1], bias[h][16 mt + g + 8][8 j + 2 t], [.. + 1])
__global__ void __launch_bounds__(256)
oa_bias_frag_kernel(const float* __restrict__ bias, float4* __restrict__ frag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;  // (h, mt, j, lane)
    if (i >= 6 * 4 * 18 * 32) return;
    const int lane = i & 31, j = (i >> 5) % 18, mt = (i / (32 * 18)) & 3, h = i / (32 * 18 * 4);
    const int g = lane >> 2, t = lane & 3;
    const float* b0 = bias + ((size_t)h * 64 + 16 * mt + g) * OA_NK + 8 * j + 2 * t;
    frag[i] = make_float4(b0[0], b0[1], b0[8 * OA_NK], b0[8 * OA_NK + 1]);
}

constexpr int OA_FWD_SMEM = 64 * QP * 2 + OA_NK * KP * 2 + 64 * 4;

__global__ void __launch_bounds__(32)
ocab_attn_fwd_kernel(const OcaArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    bf16* sQ = (bf16*)smem;
    bf16* sKV = sQ + 64 * QP;
    int* sPix = (int*)(sKV + OA_NK * KP);
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    for (int item = blockIdx.x; item < a.nwin * 6; item += gridDim.x) {
        const int win = item / 6, h = item % 6;
        oa_stage<false>(a, win, h, sQ, sKV, nullptr, sPix, lane);
#pragma unroll 1
        for (int mt = 0; mt < 4; ++mt) {
            float o[2][4], m0, m1, i0, i1;
            oa_rows(a, sQ, sKV, h, mt, lane, o, m0, m1, i0, i1);
            __syncwarp();  // every lane holds its Q fragments of this block before the rows are overwritten
            const int r0 = 16 * mt + g, r1 = r0 + 8;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                *(uint32_t*)(sQ + r0 * QP + 8 * nt + 2 * t) = pack2(o[nt][0], o[nt][1]);
                *(uint32_t*)(sQ + r1 * QP + 8 * nt + 2 * t) = pack2(o[nt][2], o[nt][3]);
            }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = lane + 32 * k, n = i >> 1, c = i & 1;
            *(uint4*)(a.out + (size_t)sPix[n] * 96 + h * 16 + c * 8) = *(const uint4*)(sQ + n * QP + c * 8);
        }
        __syncwarp();
    }
}

constexpr int OA_BWD_SMEM = 64 * QP * 2 * 2 + OA_NK * KP * 2 + 64 * OA_NK * 4 + 64 * 4 * 4 + 64 * 4;

__global__ void __launch_bounds__(32)
ocab_attn_bwd_kernel(const OcaArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    bf16* sQ = (bf16*)smem;
    bf16* sG = sQ + 64 * QP;
    bf16* sKV = sG + 64 * QP;
    float* sAcc = (float*)(sKV + OA_NK * KP);   // [64][144] sum of dS over this warp's windows (one head)
    float4* sStat = (float4*)(sAcc + 64 * OA_NK);  // per query row: max (log2 domain), 1 / sum, delta
    int* sPix = (int*)(sStat + 64);
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    const int h = blockIdx.x % 6;  // a warp keeps one head for all its windows
    for (int i = lane; i < 64 * OA_NK; i += 32) sAcc[i] = 0.f;
    __syncwarp();
    for (int win = blockIdx.x / 6; win < a.nwin; win += gridDim.x / 6) {
        oa_stage<true>(a, win, h, sQ, sKV, sG, sPix, lane);
        // ---- pass A: row statistics
#pragma unroll 1
        for (int mt = 0; mt < 4; ++mt) {
            float o[2][4], m0, m1, i0, i1;
            oa_rows(a, sQ, sKV, h, mt, lane, o, m0, m1, i0, i1);
            const int r0 = 16 * mt + g, r1 = r0 + 8;
            float d0 = 0.f, d1 = 0.f;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const float2 u = __bfloat1622float2(*(const __nv_bfloat162*)(sG + r0 * QP + 8 * nt + 2 * t));
                const float2 v = __bfloat1622float2(*(const __nv_bfloat162*)(sG + r1 * QP + 8 * nt + 2 * t));
                d0 = fmaf(o[nt][0], u.x, fmaf(o[nt][1], u.y, d0));
                d1 = fmaf(o[nt][2], v.x, fmaf(o[nt][3], v.y, d1));
            }
            d0 = quad_sum(d0); d1 = quad_sum(d1);
            if (t == 0) {
                sStat[r0] = make_float4(m0, i0, d0, 0.f);
                sStat[r1] = make_float4(m1, i1, d1, 0.f);
            }
        }
        __syncwarp();
        // ---- pass B: key chunks of 48
        float dq[4][2][4];
#pragma unroll
        for (int mt = 0; mt < 4; ++mt)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) dq[mt][nt][0] = dq[mt][nt][1] = dq[mt][nt][2] = dq[mt][nt][3] = 0.f;
#pragma unroll 1
        for (int c = 0; c < 3; ++c) {
            const int key0 = 48 * c;
            float dk[3][2][4], dv[3][2][4];
#pragma unroll
            for (int kb = 0; kb < 3; ++kb)
#pragma unroll
                for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                    for (int e = 0; e < 4; ++e) dk[kb][nt][e] = dv[kb][nt][e] = 0.f;
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) {
                const int r0 = 16 * mt + g, r1 = r0 + 8;
                float p[6][4], ds[6][4];
                oa_logits<6>(a, sQ, sKV, h, mt, key0, lane, p);
                const float4 st0 = sStat[r0], st1 = sStat[r1];
#pragma unroll
                for (int j = 0; j < 6; ++j) {
                    p[j][0] = ex2f(fmaf(p[j][0], LOG2E, -st0.x)) * st0.y; p[j][1] = ex2f(fmaf(p[j][1], LOG2E, -st0.x)) * st0.y;
                    p[j][2] = ex2f(fmaf(p[j][2], LOG2E, -st1.x)) * st1.y; p[j][3] = ex2f(fmaf(p[j][3], LOG2E, -st1.x)) * st1.y;
                    ds[j][0] = ds[j][1] = ds[j][2] = ds[j][3] = 0.f;
                }
                uint32_t ga[4];
                frag_a(ga, sG, QP, 16 * mt, 0, lane);
#pragma unroll
                for (int j = 0; j < 6; j += 2) {
                    uint32_t vb[4];
                    frag_b_nk(vb, sKV, KP, key0 + 8 * j, 16, lane);
                    mma16816(ds[j], ga[0], ga[1], ga[2], ga[3], vb[0], vb[1]);
                    mma16816(ds[j + 1], ga[0], ga[1], ga[2], ga[3], vb[2], vb[3]);
                }
                uint32_t pp[6][2], sp[6][2];
#pragma unroll
                for (int j = 0; j < 6; ++j) {
                    ds[j][0] = p[j][0] * (ds[j][0] - st0.z); ds[j][1] = p[j][1] * (ds[j][1] - st0.z);
                    ds[j][2] = p[j][2] * (ds[j][2] - st1.z); ds[j][3] = p[j][3] * (ds[j][3] - st1.z);
                    int grp = 6 * c + j + g;  // 8-float groups rotated by the row (r0 & 7 == r1 & 7 == g): conflict-free
                    if (grp >= 18) grp -= 18;
                    float2* q0 = (float2*)(sAcc + r0 * OA_NK + 8 * grp + 2 * t);
                    float2* q1 = (float2*)(sAcc + r1 * OA_NK + 8 * grp + 2 * t);
                    float2 u = *q0, v = *q1;
                    u.x += ds[j][0]; u.y += ds[j][1]; v.x += ds[j][2]; v.y += ds[j][3];
                    *q0 = u; *q1 = v;
                    pp[j][0] = pack2(p[j][0], p[j][1]); pp[j][1] = pack2(p[j][2], p[j][3]);
                    sp[j][0] = pack2(ds[j][0], ds[j][1]); sp[j][1] = pack2(ds[j][2], ds[j][3]);
                }
#pragma unroll
                for (int kt = 0; kt < 3; ++kt) {  // dQ += dS K
                    uint32_t kb[4];
                    frag_b_kn(kb, sKV, KP, key0 + 16 * kt, 0, lane);
                    mma16816(dq[mt][0], sp[2 * kt][0], sp[2 * kt][1], sp[2 * kt + 1][0], sp[2 * kt + 1][1], kb[0], kb[1]);
                    mma16816(dq[mt][1], sp[2 * kt][0], sp[2 * kt][1], sp[2 * kt + 1][0], sp[2 * kt + 1][1], kb[2], kb[3]);
                }
                uint32_t qb[4], gb[4];
                frag_b_kn(qb, sQ, QP, 16 * mt, 0, lane);
                frag_b_kn(gb, sG, QP, 16 * mt, 0, lane);
#pragma unroll
                for (int kb = 0; kb < 3; ++kb) {  // dK += dS^T Q, dV += P^T dO
                    const uint32_t s0 = movt(sp[2 * kb][0]), s1 = movt(sp[2 * kb + 1][0]);
                    const uint32_t s2 = movt(sp[2 * kb][1]), s3 = movt(sp[2 * kb + 1][1]);
                    mma16816(dk[kb][0], s0, s1, s2, s3, qb[0], qb[1]);
                    mma16816(dk[kb][1], s0, s1, s2, s3, qb[2], qb[3]);
                    const uint32_t t0 = movt(pp[2 * kb][0]), t1 = movt(pp[2 * kb + 1][0]);
                    const uint32_t t2 = movt(pp[2 * kb][1]), t3 = movt(pp[2 * kb + 1][1]);
                    mma16816(dv[kb][0], t0, t1, t2, t3, gb[0], gb[1]);
                    mma16816(dv[kb][1], t0, t1, t2, t3, gb[2], gb[3]);
                }
            }
            __syncwarp();  // the K / V rows of this chunk are not read again: they take dK / dV
#pragma unroll
            for (int kb = 0; kb < 3; ++kb)
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int k0 = key0 + 16 * kb + g, k1 = k0 + 8, cc = 8 * nt + 2 * t;
                    *(uint32_t*)(sKV + k0 * KP + cc) = pack2(dk[kb][nt][0] * 0.25f, dk[kb][nt][1] * 0.25f);
                    *(uint32_t*)(sKV + k1 * KP + cc) = pack2(dk[kb][nt][2] * 0.25f, dk[kb][nt][3] * 0.25f);
                    *(uint32_t*)(sKV + k0 * KP + 16 + cc) = pack2(dv[kb][nt][0], dv[kb][nt][1]);
                    *(uint32_t*)(sKV + k1 * KP + 16 + cc) = pack2(dv[kb][nt][2], dv[kb][nt][3]);
                }
            __syncwarp();
        }
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
            const int r0 = 16 * mt + g, r1 = r0 + 8;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                *(uint32_t*)(sQ + r0 * QP + 8 * nt + 2 * t) = pack2(dq[mt][nt][0] * 0.25f, dq[mt][nt][1] * 0.25f);
                *(uint32_t*)(sQ + r1 * QP + 8 * nt + 2 * t) = pack2(dq[mt][nt][2] * 0.25f, dq[mt][nt][3] * 0.25f);
            }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = lane + 32 * k, n = i >> 1, cc = i & 1;
            *(uint4*)(a.out + (size_t)sPix[n] * 288 + h * 16 + cc * 8) = *(const uint4*)(sQ + n * QP + cc * 8);
        }
#pragma unroll
        for (int k = 0; k < 18; ++k) {
            const int i = lane + 32 * k, m = i >> 2, cc = i & 3;
            *(uint4*)(a.dkv_win + ((size_t)win * OA_NK + m) * 192 + (cc >> 1) * 96 + h * 16 + (cc & 1) * 8) =
                *(const uint4*)(sKV + m * KP + cc * 8);
        }
        __syncwarp();
    }
    // fold the dS sums through the index buffer: table gradient [361][6].  sKV is free: 361 floats of it hold the fold.
    float* sTab = (float*)sKV;
    for (int i = lane; i < OA_TAB; i += 32) sTab[i] = 0.f;
    __syncwarp();
    for (int i = lane; i < 64 * OA_NK; i += 32) {
        const int r = i / OA_NK, cidx = i % OA_NK;
        int grp = (cidx >> 3) + (r & 7);
        if (grp >= 18) grp -= 18;
        atomicAdd(&sTab[__ldg(a.index + i)], sAcc[r * OA_NK + 8 * grp + (cidx & 7)]);
    }
    __syncwarp();
    for (int i = lane; i < OA_TAB; i += 32) atomicAdd(a.dtable + i * 6 + h, sTab[i]);
}

// dqkv[pix][96 + c] = sum over the (up to 4) windows whose 12 x 12 key region holds the pixel, in (wy, wx) order
__global__ void __launch_bounds__(256)
ocab_fold_kernel(const bf16* __restrict__ dkv_win, bf16* __restrict__ dqkv, int B, int H, int W) {
    const int nwx = W / 8, nwy = H / 8;
    const size_t n = (size_t)B * H * W * 24;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t pix = i / 24;
        const int u = (int)(i - pix * 24);
        const int x = (int)(pix % W), y = (int)((pix / W) % H), b = (int)(pix / ((size_t)W * H));
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const int wy0 = max(0, (y - 2) >> 3), wy1 = min(nwy - 1, (y + 2) >> 3);
        const int wx0 = max(0, (x - 2) >> 3), wx1 = min(nwx - 1, (x + 2) >> 3);
        for (int wy = wy0; wy <= wy1; ++wy)
            for (int wx = wx0; wx <= wx1; ++wx) {
                const int ky = y - wy * 8 + 2, kx = x - wx * 8 + 2;  // in [0, 12)
                const size_t win = ((size_t)b * nwy + wy) * nwx + wx;
                float f[8];
                oa_unpack8(__ldg((const uint4*)(dkv_win + (win * OA_NK + ky * 12 + kx) * 192) + u), f);
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] += f[k];
            }
        *((uint4*)(dqkv + pix * 288 + 96) + u) = oa_pack8(acc);
    }
}


int oa_check(int B, int H, int W) {
    if (B <= 0 || H % 8 || W % 8 || H < 8 || W < 8) {
        set_error("ocab_attn: H=%d W=%d must be multiples of 8", H, W);
        return SRB_E_SHAPE;
    }
    if ((size_t)B * H * W >= ((size_t)1 << 31) / 288) { set_error("ocab_attn: tensor too large for 32-bit indexing"); return SRB_E_SHAPE; }
    return SRB_OK;
}

}  // namespace

constexpr size_t OA_FRAG_BYTES = (size_t)6 * 64 * OA_NK * sizeof(float);
static size_t oa_dkv_bytes(int B, int H, int W) { return (size_t)B * (H / 8) * (W / 8) * OA_NK * 192 * sizeof(bf16); }
// per-window dK | dV rows, then the bias in fragment order
size_t ocab_attn_bwd_scratch_bytes(int B, int H, int W) { return oa_dkv_bytes(B, H, W) + OA_FRAG_BYTES; }

int ocab_attn_fwd(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, cudaStream_t s) {
    SRB_TRY(oa_check(B, H, W));
    OcaArgs a{};
    a.qkv = qkv; a.bias = bias; a.out = out; a.B = B; a.H = H; a.W = W; a.nwin = B * (H / 8) * (W / 8);
    const int items = a.nwin * 6, cap = num_sms() * 12;
    ocab_attn_fwd_kernel<<<items < cap ? items : cap, 32, OA_FWD_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int ocab_attn_bwd(const bf16* qkv, const float* bias, const int* index, const bf16* dout, bf16* dqkv, float* dtable,
                  void* scratch, int B, int H, int W, cudaStream_t s) {
    SRB_TRY(oa_check(B, H, W));
    OcaArgs a{};
    a.qkv = qkv; a.bias = bias; a.index = index; a.dout = dout; a.out = dqkv; a.dkv_win = (bf16*)scratch; a.dtable = dtable;
    float4* frag = (float4*)((uint8_t*)scratch + oa_dkv_bytes(B, H, W));
    if (((uintptr_t)frag & 15) == 0) {  // (a scratch base off 16 bytes keeps the row-major loads)
        oa_bias_frag_kernel<<<(6 * 4 * 18 * 32 + 255) / 256, 256, 0, s>>>(bias, frag);
        SRB_CUDA(cudaGetLastError());
        a.bias_frag = (const float*)frag;
    }
    a.B = B; a.H = H; a.W = W; a.nwin = B * (H / 8) * (W / 8);
    SRB_CUDA(cudaFuncSetAttribute(ocab_attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OA_BWD_SMEM));
    int warps_per_head = (num_sms() * 4) / 6;  // 4 resident single-warp CTAs per SM (56 KB of shared memory each)
    if (warps_per_head > a.nwin) warps_per_head = a.nwin;
    ocab_attn_bwd_kernel<<<warps_per_head * 6, 32, OA_BWD_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    const size_t n = (size_t)B * H * W * 24;
    size_t g = (n + 255) / 256;
    if (g > (size_t)num_sms() * 16) g = (size_t)num_sms() * 16;
    ocab_fold_kernel<<<(int)g, 256, 0, s>>>((const bf16*)scratch, dqkv, B, H, W);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
