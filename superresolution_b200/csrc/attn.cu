// Window self-attention of HAB ((S)W-MSA, 8x8 windows, 6 heads x 16) and overlapping cross-attention of OCAB
// (8x8 queries x 12x12 zero-padded keys).  SURVEY.md §2.2 rows K12, K15; App. A.2.4-7.
//
// The cyclic shift (torch.roll), window partition/reverse and nn.Unfold are pure index math in the loads/stores:
// nothing is materialised.  One warp per (window, head): S = QK^T and O = PV run on the tensor cores through
// warp-level mma.sync (head_dim 16 == one k-step); softmax (fp32, max-subtracted) lives in registers, P feeds the
// second MMA straight from the accumulator fragments (flash-style).
//  * HAB kernel: persistent CTAs (12 warps = 2 window slots x 6 heads), the relative-position bias of the layer sits in
//    shared memory in MMA-fragment order (bf16, 48 KB, conflict-free 8-byte reads), qkv tiles are double buffered
//    with cp.async, outputs are written back through shared memory as coalesced 16-byte stores.
//  * OCAB kernel: one CTA per window (6 warps), fragment-ordered fp32 bias read straight from L2 (coalesced).
// Attention proper is 7% of the model FLOPs; the tcgen05 budget goes to the convs/linears (conv_tc.cu).
#include "common.h"
#include "ops.h"

namespace srb {

constexpr int AP = 104;  // smem row pitch in bf16 (96 + 8): 52 words -> conflict-free fragment loads

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *(uint32_t*)&h;
}
__device__ __forceinline__ uint32_t pack_raw(bf16 lo, bf16 hi) {
    return (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
// A fragment (16 rows x 16 k) of a row-major [rows][AP] bf16 tile: r = {a0,a1,a2,a3}
__device__ __forceinline__ void load_a_frag(uint32_t (&r)[4], const bf16* tile, int row0, int k0, int lane) {
    ldsm_x4(r, tile + (row0 + (lane & 7) + ((lane >> 3) & 1) * 8) * AP + k0 + (lane >> 4) * 8);
}
// B fragments of TWO adjacent n-tiles (n0..n0+15) x 16 k from a [n][AP] (k contiguous) tile:
// r = {b0,b1 of n-tile 0, b0,b1 of n-tile 1}
__device__ __forceinline__ void load_b_frag2(uint32_t (&r)[4], const bf16* tile, int n0, int k0, int lane) {
    ldsm_x4(r, tile + (n0 + (lane & 7) + (lane >> 4) * 8) * AP + k0 + ((lane >> 3) & 1) * 8);
}
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// One (window, head) attention for the calling warp over query m-tiles [mt_lo, mt_hi).  sQ/sK/sV: [rows][AP] bf16
// tiles of this window.  BIAS_T = bf16 (shared, fragment order, 4 values / lane / (mt, j)) or float (global, fragment
// order).  The normalised output overwrites this head's 16 columns of the processed Q rows (only this warp reads
// them).  All operand fragments come from ldmatrix (V through .trans).
template <int NK, typename BIAS_T>
__device__ __forceinline__ void attend_head(bf16* sQ, const bf16* sK, const bf16* sV, const BIAS_T* bias_h,
                                            const int* srid, bool masked, int h, int lane, int mt_lo = 0,
                                            int mt_hi = 4) {
    const int g = lane >> 2, t = lane & 3;
    constexpr int NT = NK / 8;   // S n-tiles
    constexpr int KT = NK / 16;  // PV k-tiles
#pragma unroll 1
    for (int mt = mt_lo; mt < mt_hi; ++mt) {
        const int r0 = 16 * mt + g, r1 = r0 + 8;
        uint32_t qa[4];
        load_a_frag(qa, sQ, 16 * mt, h * 16, lane);
        float s[NT][4];
#pragma unroll
        for (int j = 0; j < NT; j += 2) {
            uint32_t kb[4];
            load_b_frag2(kb, sK, 8 * j, h * 16, lane);
            s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
            s[j + 1][0] = s[j + 1][1] = s[j + 1][2] = s[j + 1][3] = 0.f;
            mma_bf16_16816(s[j], qa[0], qa[1], qa[2], qa[3], kb[0], kb[1]);
            mma_bf16_16816(s[j + 1], qa[0], qa[1], qa[2], qa[3], kb[2], kb[3]);
        }
        // scale (head_dim^-0.5 = 0.25), relative position bias, shift mask (-100 across region borders)
        int rid0 = 0, rid1 = 0;
        if (masked) { rid0 = srid[r0]; rid1 = srid[r1]; }
        float m0 = -1e30f, m1 = -1e30f;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            float bb[4];
            if (sizeof(BIAS_T) == 2) {
                const uint2 raw = *(const uint2*)((const bf16*)bias_h + ((mt * NT + j) * 32 + lane) * 4);
                const float2 lo = __bfloat1622float2(*(const __nv_bfloat162*)&raw.x);
                const float2 hi = __bfloat1622float2(*(const __nv_bfloat162*)&raw.y);
                bb[0] = lo.x; bb[1] = lo.y; bb[2] = hi.x; bb[3] = hi.y;
            } else {
                const float4 raw = __ldg((const float4*)((const float*)bias_h + ((mt * NT + j) * 32 + lane) * 4));
                bb[0] = raw.x; bb[1] = raw.y; bb[2] = raw.z; bb[3] = raw.w;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) s[j][e] = fmaf(s[j][e], 0.25f, bb[e]);
            if (masked) {
                const int c = 8 * j + 2 * t;
                const int rc0 = srid[c], rc1 = srid[c + 1];
                if (rc0 != rid0) s[j][0] -= 100.f;
                if (rc1 != rid0) s[j][1] -= 100.f;
                if (rc0 != rid1) s[j][2] -= 100.f;
                if (rc1 != rid1) s[j][3] -= 100.f;
            }
            m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
            m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
        }
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            s[j][0] = __expf(s[j][0] - m0);
            s[j][1] = __expf(s[j][1] - m0);
            s[j][2] = __expf(s[j][2] - m1);
            s[j][3] = __expf(s[j][3] - m1);
            l0 += s[j][0] + s[j][1];
            l1 += s[j][2] + s[j][3];
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
        l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        float o[2][4];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
        for (int kt = 0; kt < KT; ++kt) {
            const uint32_t p0 = pack_bf16(s[2 * kt][0], s[2 * kt][1]);
            const uint32_t p1 = pack_bf16(s[2 * kt][2], s[2 * kt][3]);
            const uint32_t p2 = pack_bf16(s[2 * kt + 1][0], s[2 * kt + 1][1]);
            const uint32_t p3 = pack_bf16(s[2 * kt + 1][2], s[2 * kt + 1][3]);
            // V^T fragments: {b0,b1 of dims 0-7, b0,b1 of dims 8-15} for keys [16kt, 16kt+16)
            uint32_t vb[4];
            ldsm_x4_trans(vb, sV + (16 * kt + (lane & 7) + ((lane >> 3) & 1) * 8) * AP + h * 16 + (lane >> 4) * 8);
            mma_bf16_16816(o[0], p0, p1, p2, p3, vb[0], vb[1]);
            mma_bf16_16816(o[1], p0, p1, p2, p3, vb[2], vb[3]);
        }
        const float i0 = 1.f / l0, i1 = 1.f / l1;
        __syncwarp();  // every lane has read its Q fragments of this m-tile before they are overwritten
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            *(uint32_t*)(sQ + r0 * AP + h * 16 + 8 * nt + 2 * t) = pack_bf16(o[nt][0] * i0, o[nt][1] * i0);
            *(uint32_t*)(sQ + r1 * AP + h * 16 + 8 * nt + 2 * t) = pack_bf16(o[nt][2] * i1, o[nt][3] * i1);
        }
    }
}

// Two query m-tiles (rows [32*mh, 32*mh+32)) of one (window, head) at once: the S-MMAs, softmax and PV-MMAs of both
// tiles are interleaved stage by stage so that the tensor pipe, the MUFU (exp) and the FMA pipe overlap inside one
// warp.  bias fragments live in shared memory (bf16, fragment order).  `masked` is warp-uniform.
__device__ __forceinline__ void attend_head_pair(bf16* sQ, const bf16* sK, const bf16* sV, const bf16* bias_h,
                                                 const int* srid, bool masked, int h, int lane, int mh) {
    const int g = lane >> 2, t = lane & 3;
    uint32_t qa[2][4];
    load_a_frag(qa[0], sQ, 32 * mh, h * 16, lane);
    load_a_frag(qa[1], sQ, 32 * mh + 16, h * 16, lane);
    float s[2][8][4];
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        uint32_t kb[4];
        load_b_frag2(kb, sK, 8 * j, h * 16, lane);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            s[m][j][0] = s[m][j][1] = s[m][j][2] = s[m][j][3] = 0.f;
            s[m][j + 1][0] = s[m][j + 1][1] = s[m][j + 1][2] = s[m][j + 1][3] = 0.f;
            mma_bf16_16816(s[m][j], qa[m][0], qa[m][1], qa[m][2], qa[m][3], kb[0], kb[1]);
            mma_bf16_16816(s[m][j + 1], qa[m][0], qa[m][1], qa[m][2], qa[m][3], kb[2], kb[3]);
        }
    }
    float mx[2][2], l[2][2];
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        const int mt = 2 * mh + m;
        mx[m][0] = mx[m][1] = -1e30f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint2 raw = *(const uint2*)(bias_h + ((mt * 8 + j) * 32 + lane) * 4);
            const float2 lo = __bfloat1622float2(*(const __nv_bfloat162*)&raw.x);
            const float2 hi = __bfloat1622float2(*(const __nv_bfloat162*)&raw.y);
            s[m][j][0] = fmaf(s[m][j][0], 0.25f, lo.x);
            s[m][j][1] = fmaf(s[m][j][1], 0.25f, lo.y);
            s[m][j][2] = fmaf(s[m][j][2], 0.25f, hi.x);
            s[m][j][3] = fmaf(s[m][j][3], 0.25f, hi.y);
        }
        if (masked) {  // shift mask (-100 across region borders): only windows touching the last window row / column
            const int rid0 = srid[16 * mt + g], rid1 = srid[16 * mt + g + 8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int rc0 = srid[8 * j + 2 * t], rc1 = srid[8 * j + 2 * t + 1];
                if (rc0 != rid0) s[m][j][0] -= 100.f;
                if (rc1 != rid0) s[m][j][1] -= 100.f;
                if (rc0 != rid1) s[m][j][2] -= 100.f;
                if (rc1 != rid1) s[m][j][3] -= 100.f;
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            mx[m][0] = fmaxf(mx[m][0], fmaxf(s[m][j][0], s[m][j][1]));
            mx[m][1] = fmaxf(mx[m][1], fmaxf(s[m][j][2], s[m][j][3]));
        }
    }
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            mx[m][e] = fmaxf(mx[m][e], __shfl_xor_sync(0xffffffffu, mx[m][e], 1));
            mx[m][e] = fmaxf(mx[m][e], __shfl_xor_sync(0xffffffffu, mx[m][e], 2));
            l[m][e] = 0.f;
        }
    float o[2][2][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) o[m][nt][0] = o[m][nt][1] = o[m][nt][2] = o[m][nt][3] = 0.f;
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
        uint32_t vb[4];
        ldsm_x4_trans(vb, sV + (16 * kt + (lane & 7) + ((lane >> 3) & 1) * 8) * AP + h * 16 + (lane >> 4) * 8);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            float e[2][4];
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                const int j = 2 * kt + jj;
                e[jj][0] = __expf(s[m][j][0] - mx[m][0]);
                e[jj][1] = __expf(s[m][j][1] - mx[m][0]);
                e[jj][2] = __expf(s[m][j][2] - mx[m][1]);
                e[jj][3] = __expf(s[m][j][3] - mx[m][1]);
                l[m][0] += e[jj][0] + e[jj][1];
                l[m][1] += e[jj][2] + e[jj][3];
            }
            const uint32_t p0 = pack_bf16(e[0][0], e[0][1]), p1 = pack_bf16(e[0][2], e[0][3]);
            const uint32_t p2 = pack_bf16(e[1][0], e[1][1]), p3 = pack_bf16(e[1][2], e[1][3]);
            mma_bf16_16816(o[m][0], p0, p1, p2, p3, vb[0], vb[1]);
            mma_bf16_16816(o[m][1], p0, p1, p2, p3, vb[2], vb[3]);
        }
    }
    __syncwarp();  // every lane has read its Q fragments before they are overwritten
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        float l0 = l[m][0], l1 = l[m][1];
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
        l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        const float i0 = 1.f / l0, i1 = 1.f / l1;
        const int r0 = 32 * mh + 16 * m + g, r1 = r0 + 8;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            *(uint32_t*)(sQ + r0 * AP + h * 16 + 8 * nt + 2 * t) = pack_bf16(o[m][nt][0] * i0, o[m][nt][1] * i0);
            *(uint32_t*)(sQ + r1 * AP + h * 16 + 8 * nt + 2 * t) = pack_bf16(o[m][nt][2] * i1, o[m][nt][3] * i1);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// HAB: persistent, 2 window slots x 6 warps
// ---------------------------------------------------------------------------------------------------------------
constexpr int HAB_TILE = 3 * 64 * AP;                   // bf16 elements of one window's q|k|v tile
constexpr int HAB_BIAS = 6 * 4 * 8 * 32 * 4;            // bf16 elements (fragment order)
constexpr int HAB_SMEM = HAB_BIAS * 2 + 4 * HAB_TILE * 2 + 2 * 2 * 128 * 4;

__global__ void __launch_bounds__(384, 1)
hab_attn_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ bias_frag, bf16* __restrict__ out, int B,
                int H, int W, int shift) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    bf16* sbias = (bf16*)smem_raw;
    bf16* tiles = sbias + HAB_BIAS;                 // [slot][buf][q|k|v][64][AP]
    int* meta = (int*)(tiles + 4 * HAB_TILE);       // [slot][buf][128]: pixel index, region id
    const int slot = threadIdx.x / 192, tl = threadIdx.x % 192;
    const int nwx = W / 8, nwy = H / 8, nwin = B * nwx * nwy;

    for (int i = threadIdx.x; i < HAB_BIAS / 8; i += 384) ((uint4*)sbias)[i] = __ldg((const uint4*)bias_frag + i);

    auto issue_load = [&](int win, int buf) {
        bf16* tq = tiles + (slot * 2 + buf) * HAB_TILE;
        int* mt_ = meta + (slot * 2 + buf) * 128;
        const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
        const size_t img = (size_t)b * H * W;
        for (int i = tl; i < 64 * 36; i += 192) {
            const int n = i / 36, ck = i % 36;  // 36 x 16 B per token: q(12) k(12) v(12)
            const int ys = wy * 8 + (n >> 3), xs = wx * 8 + (n & 7);  // coordinates in the rolled image
            const int y = (ys + shift) % H, x = (xs + shift) % W;     // torch.roll(-shift): rolled[i] = orig[i+shift]
            const size_t pix = img + (size_t)y * W + x;
            const int part = ck / 12, c = ck % 12;
            cp_async16(tq + part * 64 * AP + n * AP + c * 8, qkv + pix * 288 + ck * 8);
            if (ck == 0) {
                mt_[n] = (int)(pix - img);
                const int rh = ys < H - 8 ? 0 : (ys < H - 4 ? 1 : 2);
                const int rw = xs < W - 8 ? 0 : (xs < W - 4 ? 1 : 2);
                mt_[64 + n] = rh * 3 + rw;
            }
        }
    };

    const int stride = gridDim.x * 2;
    int win = blockIdx.x * 2 + slot;
    if (win < nwin) issue_load(win, 0);
    cp_async_commit();
    __syncthreads();  // bias resident
    int it = 0;
    for (; win < nwin; win += stride, ++it) {
        const int buf = it & 1;
        if (win + stride < nwin) issue_load(win + stride, buf ^ 1);
        cp_async_commit();
        cp_async_wait<1>();
        bar_sync(1 + slot, 192);
        bf16* tq = tiles + (slot * 2 + buf) * HAB_TILE;
        const int* mt_ = meta + (slot * 2 + buf) * 128;
        const int h = tl >> 5, lane = tl & 31;
        attend_head<64, bf16>(tq, tq + 64 * AP, tq + 128 * AP, sbias + h * (4 * 8 * 32 * 4), mt_ + 64, shift > 0, h, lane);
        bar_sync(1 + slot, 192);
        // coalesced write-back of the 64 x 96 output tile (now sitting in the Q tile)
        const int b = win / (nwx * nwy);
        const size_t img = (size_t)b * H * W;
        for (int i = tl; i < 64 * 12; i += 192) {
            const int n = i / 12, c = i % 12;
            *(uint4*)(out + (img + mt_[n]) * 96 + c * 8) = *(const uint4*)(tq + n * AP + c * 8);
        }
        bar_sync(1 + slot, 192);  // tile + meta free before the next prefetch overwrites them
    }
    cp_async_wait<0>();
}

int window_attention(const bf16* qkv, const bf16* bias_frag, bf16* out, int B, int H, int W, int shift,
                     cudaStream_t s) {
    if (H % 8 || W % 8) { set_error("attention: H=%d W=%d must be multiples of the window size 8", H, W); return SRB_E_SHAPE; }
    static bool attr = false;
    if (!attr) {
        SRB_CUDA(cudaFuncSetAttribute(hab_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HAB_SMEM));
        attr = true;
    }
    const int nwin = B * (H / 8) * (W / 8);
    int grid = (nwin + 1) / 2;
    if (grid > 148) grid = 148;
    hab_attn_kernel<<<grid, 384, HAB_SMEM, s>>>(qkv, bias_frag, out, B, H, W, shift);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// OCAB: one CTA per window, 6 warps; keys/values = 12x12 neighbourhood with ZERO vectors outside the image that
// still take part in the softmax (nn.Unfold pads the projected K/V maps; logit = bias only).  App. A.2.6.
// ---------------------------------------------------------------------------------------------------------------
constexpr int OCA_SMEM = (64 + 2 * 144) * AP * 2 + 64 * 4;

__global__ void __launch_bounds__(192)
ocab_attn_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ bias_frag, bf16* __restrict__ out, int H,
                 int W) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    bf16* sQ = (bf16*)smem_raw;
    bf16* sK = sQ + 64 * AP;
    bf16* sV = sK + 144 * AP;
    int* spix = (int*)(sV + 144 * AP);
    const int nwx = W / 8, nwy = H / 8;
    const int win = blockIdx.x;
    const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
    const size_t img = (size_t)b * H * W;
    for (int i = threadIdx.x; i < 64 * 12; i += 192) {
        const int n = i / 12, ck = i % 12;
        const size_t pix = img + (size_t)(wy * 8 + (n >> 3)) * W + wx * 8 + (n & 7);
        *(uint4*)(sQ + n * AP + ck * 8) = __ldg((const uint4*)(qkv + pix * 288) + ck);
        if (ck == 0) spix[n] = (int)(pix - img);
    }
    for (int i = threadIdx.x; i < 144 * 24; i += 192) {
        const int m = i / 24, ck = i % 24;
        const int y = wy * 8 - 2 + m / 12, x = wx * 8 - 2 + m % 12;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (y >= 0 && y < H && x >= 0 && x < W)
            v = __ldg((const uint4*)(qkv + (img + (size_t)y * W + x) * 288 + 96) + ck);
        bf16* dst = (ck < 12) ? (sK + m * AP + ck * 8) : (sV + m * AP + (ck - 12) * 8);
        *(uint4*)dst = v;
    }
    __syncthreads();
    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    attend_head<144, bf16>(sQ, sK, sV, bias_frag + (size_t)h * (4 * 18 * 32 * 4), nullptr, false, h, lane);
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * 12; i += 192) {
        const int n = i / 12, c = i % 12;
        *(uint4*)(out + (img + spix[n]) * 96 + c * 8) = *(const uint4*)(sQ + n * AP + c * 8);
    }
}

int ocab_attention(const bf16* qkv, const bf16* bias_frag, bf16* out, int B, int H, int W, cudaStream_t s) {
    if (H % 8 || W % 8) { set_error("attention: H=%d W=%d must be multiples of the window size 8", H, W); return SRB_E_SHAPE; }
    static bool attr = false;
    if (!attr) {
        SRB_CUDA(cudaFuncSetAttribute(ocab_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OCA_SMEM));
        attr = true;
    }
    ocab_attn_kernel<<<B * (H / 8) * (W / 8), 192, OCA_SMEM, s>>>(qkv, bias_frag, out, H, W);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Relative-position bias, gathered once at weight-pack time into MMA C-fragment order:
//   out[(((h*4 + mt)*NT + j)*32 + lane)*4 + e] = table[wrap(index[r][c])][h],
//   r = 16*mt + lane/4 + 8*(e>>1), c = 8*j + 2*(lane%4) + (e&1), wrap = python-style negative index into [0,T)
// (App. A.2.4 / A.2.7: the OCA index buffer holds negative values; the buffer comes from the checkpoint.)
// ---------------------------------------------------------------------------------------------------------------
template <typename OUT_T>
__global__ void expand_rel_bias_kernel(const float* __restrict__ table, int T, const int64_t* __restrict__ index,
                                       int NK, OUT_T* __restrict__ out, float scale) {
    const int NT = NK / 8;
    const int n = 6 * 4 * NT * 32 * 4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int e = i & 3, lane = (i >> 2) & 31;
        const int j = (i >> 7) % NT, mt = ((i >> 7) / NT) & 3, h = (i >> 7) / (NT * 4);
        const int r = 16 * mt + (lane >> 2) + 8 * (e >> 1);
        const int c = 8 * j + 2 * (lane & 3) + (e & 1);
        long long idx = index[r * NK + c] % T;
        if (idx < 0) idx += T;
        out[i] = (OUT_T)(table[idx * 6 + h] * scale);
    }
}

int expand_rel_bias_bf16(const float* table, int T, const int64_t* index, int NK, bf16* out, float scale,
                         cudaStream_t s) {
    const int n = 6 * 64 * NK;
    expand_rel_bias_kernel<bf16><<<(n + 255) / 256, 256, 0, s>>>(table, T, index, NK, out, scale);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}
int expand_rel_bias_f32(const float* table, int T, const int64_t* index, int NK, float* out, cudaStream_t s) {
    const int n = 6 * 64 * NK;
    expand_rel_bias_kernel<float><<<(n + 255) / 256, 256, 0, s>>>(table, T, index, NK, out, 1.f);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb

namespace srb {

// ---------------------------------------------------------------------------------------------------------------
// Fused HAB attention branch (App. A.2.4):  t = x + proj( W-MSA( LN1(x) ) ),  and  xn = LN1(x)  for the CAB branch.
// One persistent CTA (6 warps = 6 heads) walks over windows; weights (qkv 288x96, proj 96x96, bf16), the layer's
// relative-position bias (fragment order) and the LayerNorm affine live in shared memory for the whole kernel.
// Per window: cp.async gather of the 64 tokens (cyclic shift = index math, next window prefetched) -> LayerNorm in
// fp32 -> qkv GEMM -> per-head attention -> proj GEMM + bias + shortcut -> coalesced scatter.  Nothing but x is read
// from HBM and only xn and t are written: the qkv / attention-output round trips (1.6 GB per block at B=16) are gone.
// The three small GEMMs (K = 96) run on warp-level mma.sync next to the softmax/LN CUDA-core work of the same warps.
// ---------------------------------------------------------------------------------------------------------------
constexpr int FA_WQKV = 288 * AP, FA_WPROJ = 96 * AP, FA_TILE = 64 * AP;
constexpr int FA_SMEM = (FA_WQKV + FA_WPROJ + HAB_BIAS + 2 * FA_TILE /*x*/ + FA_TILE /*xn*/ + 3 * FA_TILE /*qkv*/) * 2 +
                        2 * 128 * 4 /*meta*/ + (288 + 96 + 96 + 96) * 4;

struct FusedAttnArgs {
    const bf16* x;        // [B,H,W,96] token stream (shortcut)
    const float *ln_g, *ln_b;
    const bf16* wqkv;     // [288][96]
    const float* bqkv;    // [288]
    const bf16* bias_frag;
    const bf16* wproj;    // [96][96]
    const float* bproj;   // [96]
    bf16* xn;             // [B,H,W,96] LN1(x)
    bf16* t;              // [B,H,W,96] x + proj(attn)
    int B, H, W, shift;
    long long* dbg;  // optional [grid][8] per-phase cycle totals (development aid)
};

constexpr int FA_THREADS = 384;  // 12 warps: (m-half, head / column group)

__global__ void __launch_bounds__(FA_THREADS, 1)
hab_fused_attn_kernel(const FusedAttnArgs a) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    bf16* sWq = (bf16*)smem_raw;
    bf16* sWp = sWq + FA_WQKV;
    bf16* sBias = sWp + FA_WPROJ;
    bf16* sX = sBias + HAB_BIAS;        // [2][64][AP]
    bf16* sXN = sX + 2 * FA_TILE;       // [64][AP]
    bf16* sQKV = sXN + FA_TILE;         // [3][64][AP]
    int* meta = (int*)(sQKV + 3 * FA_TILE);  // [2][128]
    float* sbq = (float*)(meta + 256);
    float* sbp = sbq + 288;
    float* sg = sbp + 96;
    float* sb = sg + 96;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tq = lane & 3;
    const int mh = warp / 6, ng = warp % 6;  // m-half (m-tiles 2mh, 2mh+1) and column group / head
    const int H = a.H, W = a.W, shift = a.shift;
    const int nwx = W / 8, nwy = H / 8, nwin = a.B * nwx * nwy;

    // ---- one-time: weights, bias tables, affine parameters ----
    for (int i = tid; i < 288 * 12; i += FA_THREADS) *(uint4*)(sWq + (i / 12) * AP + (i % 12) * 8) = __ldg((const uint4*)a.wqkv + i);
    for (int i = tid; i < 96 * 12; i += FA_THREADS) *(uint4*)(sWp + (i / 12) * AP + (i % 12) * 8) = __ldg((const uint4*)a.wproj + i);
    for (int i = tid; i < HAB_BIAS / 8; i += FA_THREADS) ((uint4*)sBias)[i] = __ldg((const uint4*)a.bias_frag + i);
    if (tid < 288) sbq[tid] = a.bqkv[tid];
    if (tid < 96) { sbp[tid] = a.bproj[tid]; sg[tid] = a.ln_g[tid]; sb[tid] = a.ln_b[tid]; }

    // each thread always moves the same two 16-byte chunks of a window tile: (token n, chunk ck) fixed per thread
    const int ld_n0 = tid / 12, ld_c0 = tid % 12, ld_n1 = (tid + FA_THREADS) / 12, ld_c1 = (tid + FA_THREADS) % 12;
    auto issue_load = [&](int win, int buf) {
        bf16* dst = sX + buf * FA_TILE;
        int* mt_ = meta + buf * 128;
        const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
        const size_t img = (size_t)b * H * W;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int n = k ? ld_n1 : ld_n0, ck = k ? ld_c1 : ld_c0;
            const int ys = wy * 8 + (n >> 3), xs = wx * 8 + (n & 7);  // rolled-image coordinates
            int y = ys + shift, x = xs + shift;                       // torch.roll(-shift): rolled[i] = orig[i+shift]
            if (y >= H) y -= H;
            if (x >= W) x -= W;
            const int pix = y * W + x;
            cp_async16(dst + n * AP + ck * 8, a.x + (img + pix) * 96 + ck * 8);
            if (ck == 0) {
                mt_[n] = pix;
                const int rh = ys < H - 8 ? 0 : (ys < H - 4 ? 1 : 2);
                const int rw = xs < W - 8 ? 0 : (xs < W - 4 ? 1 : 2);
                mt_[64 + n] = rh * 3 + rw;
            }
        }
    };

    int win = blockIdx.x;
    if (win < nwin) issue_load(win, 0);
    cp_async_commit();
    int it = 0;
    long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tprev = clock64();
#define SRB_PHASE(i) do { if (a.dbg && tid == 0) { const long long tn = clock64(); ph[i] += tn - tprev; tprev = tn; } } while (0)
    for (; win < nwin; win += gridDim.x, ++it) {
        const int buf = it & 1;
        if (win + (int)gridDim.x < nwin) issue_load(win + gridDim.x, buf ^ 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        SRB_PHASE(0);
        bf16* x_t = sX + buf * FA_TILE;
        const int* mt_ = meta + buf * 128;
        const size_t img = (size_t)(win / (nwx * nwy)) * H * W;

        // ---- LayerNorm (fp32, eps 1e-5): 4 lanes per token, 24 channels each ----
        if (tid < 256) {
            const int n = tid >> 2, part = tid & 3;
            float v[24];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const uint4 raw = *(const uint4*)(x_t + n * AP + part * 24 + c * 8);
                const __nv_bfloat162* h2 = (const __nv_bfloat162*)&raw;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = __bfloat1622float2(h2[k]);
                    v[c * 8 + 2 * k] = f.x;
                    v[c * 8 + 2 * k + 1] = f.y;
                }
            }
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 24; ++j) s += v[j];
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            const float mean = s * (1.f / 96.f);
            float q = 0.f;
#pragma unroll
            for (int j = 0; j < 24; ++j) q += (v[j] - mean) * (v[j] - mean);
            q += __shfl_xor_sync(0xffffffffu, q, 1);
            q += __shfl_xor_sync(0xffffffffu, q, 2);
            const float rstd = rsqrtf(q * (1.f / 96.f) + 1e-5f);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                uint4 o;
                __nv_bfloat162* h2 = (__nv_bfloat162*)&o;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int ch = part * 24 + c * 8 + 2 * k;
                    h2[k] = __floats2bfloat162_rn((v[c * 8 + 2 * k] - mean) * rstd * sg[ch] + sb[ch],
                                                  (v[c * 8 + 2 * k + 1] - mean) * rstd * sg[ch + 1] + sb[ch + 1]);
                }
                *(uint4*)(sXN + n * AP + part * 24 + c * 8) = o;
            }
        }
        __syncthreads();
        SRB_PHASE(1);
        // xn -> HBM for the CAB branch (coalesced 16-byte chunks)
        for (int i = tid; i < 64 * 12; i += FA_THREADS) {
            const int n = i / 12, c = i % 12;
            *(uint4*)(a.xn + (img + mt_[n]) * 96 + c * 8) = *(const uint4*)(sXN + n * AP + c * 8);
        }

        // ---- qkv = xn @ Wqkv^T + b : warp (mh, ng) owns rows [32mh, 32mh+32) x columns [48ng, 48ng+48) ----
        {
            float acc[2][6][4];
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int n = 0; n < 6; ++n) acc[m][n][0] = acc[m][n][1] = acc[m][n][2] = acc[m][n][3] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 6; ++ks) {
                uint32_t af[2][4], bfr[3][4];
                load_a_frag(af[0], sXN, 32 * mh, 16 * ks, lane);
                load_a_frag(af[1], sXN, 32 * mh + 16, 16 * ks, lane);
#pragma unroll
                for (int n2 = 0; n2 < 3; ++n2) load_b_frag2(bfr[n2], sWq, 48 * ng + 16 * n2, 16 * ks, lane);
#pragma unroll
                for (int m = 0; m < 2; ++m)
#pragma unroll
                    for (int n2 = 0; n2 < 3; ++n2) {
                        mma_bf16_16816(acc[m][2 * n2], af[m][0], af[m][1], af[m][2], af[m][3], bfr[n2][0], bfr[n2][1]);
                        mma_bf16_16816(acc[m][2 * n2 + 1], af[m][0], af[m][1], af[m][2], af[m][3], bfr[n2][2], bfr[n2][3]);
                    }
            }
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                const int r0 = 32 * mh + 16 * m + g, r1 = r0 + 8;
#pragma unroll
                for (int n = 0; n < 6; ++n) {
                    const int c = 48 * ng + 8 * n + 2 * tq;  // output column in [0,288)
                    const int part = c / 96, cc = c % 96;
                    const float b0 = sbq[c], b1 = sbq[c + 1];
                    bf16* dst = sQKV + part * FA_TILE + cc;
                    *(uint32_t*)(dst + r0 * AP) = pack_bf16(acc[m][n][0] + b0, acc[m][n][1] + b1);
                    *(uint32_t*)(dst + r1 * AP) = pack_bf16(acc[m][n][2] + b0, acc[m][n][3] + b1);
                }
            }
        }
        __syncthreads();
        SRB_PHASE(2);

        // ---- per-head attention (head ng, query rows of m-half mh); output overwrites the Q tile ----
        {
            const int wxx = win % nwx, wyy = (win / nwx) % nwy;
            const bool masked = shift > 0 && (wxx == nwx - 1 || wyy == nwy - 1);  // other windows: one region, no mask
            attend_head_pair(sQKV, sQKV + FA_TILE, sQKV + 2 * FA_TILE, sBias + ng * (4 * 8 * 32 * 4), mt_ + 64, masked,
                             ng, lane, mh);
        }
        __syncthreads();
        SRB_PHASE(3);

        // ---- t = x + attn @ Wproj^T + b : warp (mh, ng) owns rows [32mh, +32) x columns [16ng, +16) ----
        {
            float acc[2][2][4];
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int n = 0; n < 2; ++n) acc[m][n][0] = acc[m][n][1] = acc[m][n][2] = acc[m][n][3] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 6; ++ks) {
                uint32_t af[2][4], bfr[4];
                load_a_frag(af[0], sQKV, 32 * mh, 16 * ks, lane);
                load_a_frag(af[1], sQKV, 32 * mh + 16, 16 * ks, lane);
                load_b_frag2(bfr, sWp, 16 * ng, 16 * ks, lane);
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    mma_bf16_16816(acc[m][0], af[m][0], af[m][1], af[m][2], af[m][3], bfr[0], bfr[1]);
                    mma_bf16_16816(acc[m][1], af[m][0], af[m][1], af[m][2], af[m][3], bfr[2], bfr[3]);
                }
            }
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                const int r0 = 32 * mh + 16 * m + g, r1 = r0 + 8;
#pragma unroll
                for (int n = 0; n < 2; ++n) {
                    const int c = 16 * ng + 8 * n + 2 * tq;
                    const float b0 = sbp[c], b1 = sbp[c + 1];
                    bf16* p0 = x_t + r0 * AP + c;
                    bf16* p1 = x_t + r1 * AP + c;
                    const float2 x0 = __bfloat1622float2(*(const __nv_bfloat162*)p0);
                    const float2 x1 = __bfloat1622float2(*(const __nv_bfloat162*)p1);
                    *(uint32_t*)p0 = pack_bf16(acc[m][n][0] + b0 + x0.x, acc[m][n][1] + b1 + x0.y);
                    *(uint32_t*)p1 = pack_bf16(acc[m][n][2] + b0 + x1.x, acc[m][n][3] + b1 + x1.y);
                }
            }
        }
        __syncthreads();
        SRB_PHASE(4);
        for (int i = tid; i < 64 * 12; i += FA_THREADS) {
            const int n = i / 12, c = i % 12;
            *(uint4*)(a.t + (img + mt_[n]) * 96 + c * 8) = *(const uint4*)(x_t + n * AP + c * 8);
        }
        __syncthreads();  // x tile / meta / xn / qkv tiles free for the next window
        SRB_PHASE(5);
    }
    cp_async_wait<0>();
    if (a.dbg && tid == 0)
        for (int i = 0; i < 8; ++i) a.dbg[blockIdx.x * 8 + i] = ph[i];
#undef SRB_PHASE
}

long long* g_fa_dbg = nullptr;  // set through srb_debug_set_phase_buffer (development aid)

int hab_fused_attention(const bf16* x, const float* ln_g, const float* ln_b, const bf16* wqkv, const float* bqkv,
                        const bf16* bias_frag, const bf16* wproj, const float* bproj, bf16* xn, bf16* t, int B, int H,
                        int W, int shift, cudaStream_t s) {
    if (H % 8 || W % 8) { set_error("attention: H=%d W=%d must be multiples of the window size 8", H, W); return SRB_E_SHAPE; }
    static bool attr = false;
    if (!attr) {
        SRB_CUDA(cudaFuncSetAttribute(hab_fused_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM));
        attr = true;
    }
    FusedAttnArgs a;
    a.x = x; a.ln_g = ln_g; a.ln_b = ln_b; a.wqkv = wqkv; a.bqkv = bqkv; a.bias_frag = bias_frag; a.wproj = wproj;
    a.bproj = bproj; a.xn = xn; a.t = t; a.B = B; a.H = H; a.W = W; a.shift = shift;
    a.dbg = g_fa_dbg;
    const int nwin = B * (H / 8) * (W / 8);
    hab_fused_attn_kernel<<<nwin < 148 ? nwin : 148, FA_THREADS, FA_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
