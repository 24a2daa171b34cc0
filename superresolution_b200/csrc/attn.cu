// Overlapping cross-attention core of OCAB (8x8 query windows x 12x12 zero-padded key/value windows, 6 heads x 16)
// and the relative-position-bias expansion shared with the HAB branch (hab_tc.cu).  SURVEY.md §2.2 K15; App. A.2.6-7.
//
// nn.Unfold is pure index math in the loads: the 0.9 GB unfolded key/value buffer of the reference never exists, the
// zero-padded keys stay in the softmax with logit = bias.  One CTA per window, one warp per head: S = QK^T and O = PV
// run on warp-level mma.sync (head_dim 16 == one k-step); softmax (fp32, max-subtracted) lives in registers, P feeds
// the second MMA straight from the accumulator fragments (flash-style); fragment-ordered bf16 bias is read from L2.
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

__device__ __forceinline__ float ex2f(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// One (window, head) attention for the calling warp over the 64 query rows.  sQ/sK/sV: [rows][AP] bf16 tiles of this
// window.  The logits arrive in the log2 domain: q was scaled by head_dim^-0.5 * log2(e) inside the qkv linear (folded
// into its weights at pack time) and bias_h (global, MMA C-fragment order, 4 values / lane / (mt, j)) holds
// bias * log2(e); the bias is the accumulator initialiser of the S MMAs.  The normalised output overwrites this head's
// 16 columns of the Q rows (only this warp reads them).  All operand fragments come from ldmatrix (V through .trans).
template <int NK>
__device__ __forceinline__ void attend_head(bf16* sQ, const bf16* sK, const bf16* sV, const bf16* bias_h, int h,
                                            int lane) {
    const int g = lane >> 2, t = lane & 3;
    constexpr int NT = NK / 8;   // S n-tiles
    constexpr int KT = NK / 16;  // PV k-tiles
#pragma unroll 1
    for (int mt = 0; mt < 4; ++mt) {
        const int r0 = 16 * mt + g, r1 = r0 + 8;
        uint32_t qa[4];
        load_a_frag(qa, sQ, 16 * mt, h * 16, lane);
        float s[NT][4];
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            const uint2 raw = __ldg((const uint2*)(bias_h + ((mt * NT + j) * 32 + lane) * 4));
            const float2 lo = __bfloat1622float2(*(const __nv_bfloat162*)&raw.x);
            const float2 hi = __bfloat1622float2(*(const __nv_bfloat162*)&raw.y);
            s[j][0] = lo.x; s[j][1] = lo.y; s[j][2] = hi.x; s[j][3] = hi.y;
        }
#pragma unroll
        for (int j = 0; j < NT; j += 2) {
            uint32_t kb[4];
            load_b_frag2(kb, sK, 8 * j, h * 16, lane);
            mma_bf16_16816(s[j], qa[0], qa[1], qa[2], qa[3], kb[0], kb[1]);
            mma_bf16_16816(s[j + 1], qa[0], qa[1], qa[2], qa[3], kb[2], kb[3]);
        }
        float m0 = -1e30f, m1 = -1e30f;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
            m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
        }
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
        // row sums on the tensor cores: P x ones (the sum of the bf16-rounded probabilities the PV product really uses;
        // no FADD per element, no shuffles)
        float lsum[4] = {0.f, 0.f, 0.f, 0.f};
        float o[2][4];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
        for (int kt = 0; kt < KT; ++kt) {
            float e[2][4];
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                const int j = 2 * kt + jj;
                e[jj][0] = ex2f(s[j][0] - m0);
                e[jj][1] = ex2f(s[j][1] - m0);
                e[jj][2] = ex2f(s[j][2] - m1);
                e[jj][3] = ex2f(s[j][3] - m1);
            }
            const uint32_t p0 = pack_bf16(e[0][0], e[0][1]), p1 = pack_bf16(e[0][2], e[0][3]);
            const uint32_t p2 = pack_bf16(e[1][0], e[1][1]), p3 = pack_bf16(e[1][2], e[1][3]);
            // V^T fragments: {b0,b1 of dims 0-7, b0,b1 of dims 8-15} for keys [16kt, 16kt+16)
            uint32_t vb[4];
            ldsm_x4_trans(vb, sV + (16 * kt + (lane & 7) + ((lane >> 3) & 1) * 8) * AP + h * 16 + (lane >> 4) * 8);
            mma_bf16_16816(o[0], p0, p1, p2, p3, vb[0], vb[1]);
            mma_bf16_16816(o[1], p0, p1, p2, p3, vb[2], vb[3]);
            mma_bf16_16816(lsum, p0, p1, p2, p3, 0x3F803F80u, 0x3F803F80u);  // B = ones: every column = the row sum
        }
        const float i0 = __fdividef(1.f, lsum[0]), i1 = __fdividef(1.f, lsum[2]);
        __syncwarp();  // every lane has read its Q fragments of this m-tile before they are overwritten
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            *(uint32_t*)(sQ + r0 * AP + h * 16 + 8 * nt + 2 * t) = pack_bf16(o[nt][0] * i0, o[nt][1] * i0);
            *(uint32_t*)(sQ + r1 * AP + h * 16 + 8 * nt + 2 * t) = pack_bf16(o[nt][2] * i1, o[nt][3] * i1);
        }
    }
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
    // every 16-byte chunk of the window is one cp.async: all loads of the CTA are in flight at once.  192 threads = 16
    // rows x 12 chunks (Q) = 8 rows x 24 chunks (K | V): a thread keeps its chunk column and steps through the rows with
    // increments (the div / mod form of this index math was a fifth of the kernel's instructions).
    {
        const int ck = threadIdx.x % 12;
        const bf16* src = qkv + (img + (size_t)(wy * 8) * W + wx * 8) * 288 + ck * 8;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int n = threadIdx.x / 12 + 16 * it;  // window token: row n >> 3, column n & 7
            const int off = (n >> 3) * W + (n & 7);
            cp_async16(sQ + n * AP + ck * 8, src + (size_t)off * 288);
            if (ck == 0) spix[n] = (wy * 8) * W + wx * 8 + off;
        }
    }
    {
        const int ck = threadIdx.x % 24;
        bf16* dst0 = (ck < 12) ? (sK + ck * 8) : (sV + (ck - 12) * 8);
        const bf16* src = qkv + img * 288 + 96 + ck * 8;
        int my = (threadIdx.x / 24) / 12, mx = (threadIdx.x / 24) % 12;  // key (my, mx) of the 12 x 12 neighbourhood
        int m = threadIdx.x / 24;
#pragma unroll 3
        for (int it = 0; it < 18; ++it) {
            const int y = wy * 8 - 2 + my, x = wx * 8 - 2 + mx;
            bf16* dst = dst0 + m * AP;
            if (y >= 0 && y < H && x >= 0 && x < W)
                cp_async16(dst, src + ((size_t)y * W + x) * 288);
            else
                *(uint4*)dst = make_uint4(0, 0, 0, 0);  // nn.Unfold zero padding of the projected K / V maps
            m += 8;
            mx += 8;
            if (mx >= 12) { mx -= 12; ++my; }
        }
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();
    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    attend_head<144>(sQ, sK, sV, bias_frag + (size_t)h * (4 * 18 * 32 * 4), h, lane);
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * 12; i += 192) {
        const int n = i / 12, c = i % 12;
        *(uint4*)(out + (img + spix[n]) * 96 + c * 8) = *(const uint4*)(sQ + n * AP + c * 8);
    }
}

int ocab_attention(const bf16* qkv, const bf16* bias_frag, bf16* out, int B, int H, int W, cudaStream_t s) {
    if (H % 8 || W % 8) { set_error("attention: H=%d W=%d must be multiples of the window size 8", H, W); return SRB_E_SHAPE; }
    static std::atomic<unsigned long long> attr{0};
    SRB_TRY(opt_in_dynamic_smem(ocab_attn_kernel, OCA_SMEM, attr));
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

// Operand-plane variant for the tcgen05 attention core of hab_tc.cu: out[h][r / 8][c][r % 8]
__global__ void expand_rel_bias_rows_kernel(const float* __restrict__ table, int T, const int64_t* __restrict__ index,
                                            int NK, bf16* __restrict__ out, float scale) {
    const int n = 6 * 64 * NK;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int c = i % NK, r = (i / NK) % 64, h = i / (NK * 64);
        long long idx = index[r * NK + c] % T;
        if (idx < 0) idx += T;
        // B operand of the bias MMA (S += I * Bias^T): K-major planes [query / 8][key][query % 8]
        out[((h * 8 + (r >> 3)) * NK + c) * 8 + (r & 7)] = __float2bfloat16_rn(table[idx * 6 + h] * scale);
    }
}

int expand_rel_bias_bf16(const float* table, int T, const int64_t* index, int NK, bf16* out, float scale,
                         cudaStream_t s) {
    const int n = 6 * 64 * NK;
    if (NK == 64) {  // HAB (hab_tc.cu): [6][8][64][8]
        expand_rel_bias_rows_kernel<<<(n + 255) / 256, 256, 0, s>>>(table, T, index, NK, out, scale);
        SRB_CUDA(cudaGetLastError());
        return SRB_OK;
    }
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

long long* g_fa_dbg = nullptr;  // set through srb_debug_set_phase_buffer (development aid: per-phase cycle buffers)

}  // namespace srb
