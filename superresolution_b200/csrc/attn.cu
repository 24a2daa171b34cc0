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
    asm volatile(
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
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// One (window, head) attention for the calling warp.  sQ/sK/sV: [rows][AP] bf16 tiles of this window.
// BIAS_T = bf16 (shared, fragment order, 4 values / lane / (mt, j)) or float (global, fragment order).
// The normalised output overwrites this head's 16 columns of sQ (only this warp reads them).
template <int NK, typename BIAS_T>
__device__ __forceinline__ void attend_head(bf16* sQ, const bf16* sK, const bf16* sV, const BIAS_T* bias_h,
                                            const int* srid, bool masked, int h, int lane) {
    const int g = lane >> 2, t = lane & 3;
    constexpr int NT = NK / 8;   // S n-tiles
    constexpr int KT = NK / 16;  // PV k-tiles
#pragma unroll 1
    for (int mt = 0; mt < 4; ++mt) {
        const int r0 = 16 * mt + g, r1 = r0 + 8;
        const uint32_t a0 = *(const uint32_t*)(sQ + r0 * AP + h * 16 + 2 * t);
        const uint32_t a1 = *(const uint32_t*)(sQ + r1 * AP + h * 16 + 2 * t);
        const uint32_t a2 = *(const uint32_t*)(sQ + r0 * AP + h * 16 + 2 * t + 8);
        const uint32_t a3 = *(const uint32_t*)(sQ + r1 * AP + h * 16 + 2 * t + 8);
        float s[NT][4];
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
            const uint32_t b0 = *(const uint32_t*)(sK + (8 * j + g) * AP + h * 16 + 2 * t);
            const uint32_t b1 = *(const uint32_t*)(sK + (8 * j + g) * AP + h * 16 + 2 * t + 8);
            mma_bf16_16816(s[j], a0, a1, a2, a3, b0, b1);
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
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const bf16* vp = sV + (16 * kt + 2 * t) * AP + h * 16 + 8 * nt + g;
                const uint32_t b0 = pack_raw(vp[0], vp[AP]);
                const uint32_t b1 = pack_raw(vp[8 * AP], vp[9 * AP]);
                mma_bf16_16816(o[nt], p0, p1, p2, p3, b0, b1);
            }
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
ocab_attn_kernel(const bf16* __restrict__ qkv, const float* __restrict__ bias_frag, bf16* __restrict__ out, int H,
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
    attend_head<144, float>(sQ, sK, sV, bias_frag + (size_t)h * (4 * 18 * 32 * 4), nullptr, false, h, lane);
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * 12; i += 192) {
        const int n = i / 12, c = i % 12;
        *(uint4*)(out + (img + spix[n]) * 96 + c * 8) = *(const uint4*)(sQ + n * AP + c * 8);
    }
}

int ocab_attention(const bf16* qkv, const float* bias_frag, bf16* out, int B, int H, int W, cudaStream_t s) {
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
                                       int NK, OUT_T* __restrict__ out) {
    const int NT = NK / 8;
    const int n = 6 * 4 * NT * 32 * 4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int e = i & 3, lane = (i >> 2) & 31;
        const int j = (i >> 7) % NT, mt = ((i >> 7) / NT) & 3, h = (i >> 7) / (NT * 4);
        const int r = 16 * mt + (lane >> 2) + 8 * (e >> 1);
        const int c = 8 * j + 2 * (lane & 3) + (e & 1);
        long long idx = index[r * NK + c] % T;
        if (idx < 0) idx += T;
        out[i] = (OUT_T)table[idx * 6 + h];
    }
}

int expand_rel_bias_bf16(const float* table, int T, const int64_t* index, int NK, bf16* out, cudaStream_t s) {
    const int n = 6 * 64 * NK;
    expand_rel_bias_kernel<bf16><<<(n + 255) / 256, 256, 0, s>>>(table, T, index, NK, out);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}
int expand_rel_bias_f32(const float* table, int T, const int64_t* index, int NK, float* out, cudaStream_t s) {
    const int n = 6 * 64 * NK;
    expand_rel_bias_kernel<float><<<(n + 255) / 256, 256, 0, s>>>(table, T, index, NK, out);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
