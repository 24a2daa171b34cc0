// Window self-attention of HAB ((S)W-MSA, 8x8 windows, 6 heads x 16) and overlapping cross-attention of OCAB
// (8x8 queries x 12x12 zero-padded keys).  SURVEY.md §2.2 rows K12, K15; App. A.2.4-7.
//
// One CTA per window, one warp per head.  The cyclic shift (torch.roll), window partition/reverse and nn.Unfold are
// pure index math in the loads/stores: nothing is materialised.  S = QK^T and O = PV run on the tensor cores through
// warp-level mma.sync (head_dim 16 == one k-step); softmax (fp32, max-subtracted) lives in registers, P feeds the
// second MMA straight from the accumulator fragments (flash-style).  Attention proper is 7% of the model FLOPs; the
// tcgen05 budget goes to the convs/linears (conv_tc.cu).  [tcgen05 version of this kernel: planned, DESIGN.md]
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

template <int NK, bool OCA>
__global__ void __launch_bounds__(192)
window_attn_kernel(const bf16* __restrict__ qkv, const float* __restrict__ bias, bf16* __restrict__ out, int H, int W,
                   int shift) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    bf16* sQ = (bf16*)smem_raw;      // [64][AP]
    bf16* sK = sQ + 64 * AP;         // [NK][AP]
    bf16* sV = sK + NK * AP;         // [NK][AP]
    int* spix = (int*)(sV + NK * AP);  // [64] pixel index of each query token
    int* srid = spix + 64;             // [64] shift-mask region id

    const int nwx = W / 8, nwy = H / 8;
    const int win = blockIdx.x;
    const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
    const size_t img = (size_t)b * H * W;

    // ---- load phase: coalesced 16-byte chunks, 12 per (token, part) ----
    for (int i = threadIdx.x; i < 64 * 12; i += blockDim.x) {
        const int n = i / 12, ck = i % 12;
        const int ys = wy * 8 + (n >> 3), xs = wx * 8 + (n & 7);  // coordinates in the rolled image
        const int y = (ys + shift) % H, x = (xs + shift) % W;     // torch.roll(x, -shift): rolled[i] = orig[i+shift]
        const size_t pix = img + (size_t)y * W + x;
        const uint4* src = (const uint4*)(qkv + pix * 288) + ck;
        *(uint4*)(sQ + n * AP + ck * 8) = __ldg(src);
        if (!OCA) {
            *(uint4*)(sK + n * AP + ck * 8) = __ldg(src + 12);
            *(uint4*)(sV + n * AP + ck * 8) = __ldg(src + 24);
        }
        if (ck == 0) {
            spix[n] = (int)(pix - img);
            const int rh = ys < H - 8 ? 0 : (ys < H - 4 ? 1 : 2);
            const int rw = xs < W - 8 ? 0 : (xs < W - 4 ? 1 : 2);
            srid[n] = rh * 3 + rw;
        }
    }
    if (OCA) {
        // nn.Unfold(kernel 12, stride 8, padding 2) on the projected K / V maps: out-of-image slots are ZERO vectors
        // that still take part in the softmax (logit = bias only).  App. A.2.6.
        for (int i = threadIdx.x; i < NK * 12; i += blockDim.x) {
            const int m = i / 12, ck = i % 12;
            const int y = wy * 8 - 2 + m / 12, x = wx * 8 - 2 + m % 12;
            uint4 kk = make_uint4(0, 0, 0, 0), vv = kk;
            if (y >= 0 && y < H && x >= 0 && x < W) {
                const uint4* src = (const uint4*)(qkv + (img + (size_t)y * W + x) * 288) + ck;
                kk = __ldg(src + 12);
                vv = __ldg(src + 24);
            }
            *(uint4*)(sK + m * AP + ck * 8) = kk;
            *(uint4*)(sV + m * AP + ck * 8) = vv;
        }
    }
    __syncthreads();

    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const float* bh = bias + (size_t)h * 64 * NK;
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
        // scale (head_dim^-0.5 = 0.25), relative position bias, shift mask
        const int rid0 = srid[r0], rid1 = srid[r1];
        float m0 = -1e30f, m1 = -1e30f;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            const int c = 8 * j + 2 * t;
            const float2 bb0 = __ldg((const float2*)(bh + r0 * NK + c));
            const float2 bb1 = __ldg((const float2*)(bh + r1 * NK + c));
            s[j][0] = s[j][0] * 0.25f + bb0.x;
            s[j][1] = s[j][1] * 0.25f + bb0.y;
            s[j][2] = s[j][2] * 0.25f + bb1.x;
            s[j][3] = s[j][3] * 0.25f + bb1.y;
            if (!OCA && shift > 0) {
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
        // O = P V
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
        bf16* o0 = out + (img + spix[r0]) * 96 + h * 16 + 2 * t;
        bf16* o1 = out + (img + spix[r1]) * 96 + h * 16 + 2 * t;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            *(uint32_t*)(o0 + 8 * nt) = pack_bf16(o[nt][0] * i0, o[nt][1] * i0);
            *(uint32_t*)(o1 + 8 * nt) = pack_bf16(o[nt][2] * i1, o[nt][3] * i1);
        }
    }
}

template <int NK, bool OCA>
static int launch_attn(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, int shift,
                       cudaStream_t s) {
    if (H % 8 || W % 8) { set_error("attention: H=%d W=%d must be multiples of the window size 8", H, W); return SRB_E_SHAPE; }
    constexpr int smem = (64 + 2 * NK) * AP * 2 + 128 * 4;
    static bool attr = false;
    if (!attr) {
        SRB_CUDA(cudaFuncSetAttribute(window_attn_kernel<NK, OCA>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr = true;
    }
    const int nwin = B * (H / 8) * (W / 8);
    window_attn_kernel<NK, OCA><<<nwin, 192, smem, s>>>(qkv, bias, out, H, W, shift);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int window_attention(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, int shift, cudaStream_t s) {
    return launch_attn<64, false>(qkv, bias, out, B, H, W, shift, s);
}
int ocab_attention(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, cudaStream_t s) {
    return launch_attn<144, true>(qkv, bias, out, B, H, W, 0, s);
}

// bias[h][q][k] = table[index[q][k] wrapped python-style into [0,T)][h]   (App. A.2.4 / A.2.7: the OCA index buffer
// holds negative values that rely on negative-index wrap; the buffer is read from the checkpoint, never hard-coded)
__global__ void expand_rel_bias_kernel(const float* __restrict__ table, int T, const int64_t* __restrict__ index,
                                       int NK, float* __restrict__ bias) {
    const int n = 64 * NK;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n * 6; i += gridDim.x * blockDim.x) {
        const int h = i / n, qk = i % n;
        long long idx = index[qk] % T;
        if (idx < 0) idx += T;
        bias[i] = table[idx * 6 + h];
    }
}

int expand_rel_bias(const float* table, int T, const int64_t* index, int NK, float* bias, cudaStream_t s) {
    const int n = 64 * NK * 6;
    expand_rel_bias_kernel<<<(n + 255) / 256, 256, 0, s>>>(table, T, index, NK, bias);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
