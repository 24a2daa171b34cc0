// Overlapping cross-attention core of OCAB from a saved qkv tensor, forward AND backward — the un-fused operator pair of the
// stage-2 training path (upstream hat_arch.py OCAB.forward and its autograd).  The inference path keeps
// the mma.sync kernel in attn.cu.
//
//   8 x 8 query windows attend to the 12 x 12 key/value window around them (nn.Unfold(kernel 12, stride 8, padding 2) of the
//   PROJECTED k/v maps: keys outside the image are zero vectors that still take part in the softmax with logit = bias).
//   6 heads x 16 channels, logits = q.k * 0.25 + bias[h][i][j], out = P v.
//
// One CTA per window, one warp per head.  Forward / backward pass 1: a lane owns query rows {lane, lane + 32} with the 144
// probabilities in registers.  Backward pass 2 is key-owned (lane owns keys lane + 32 jj) and RECOMPUTES P from the saved
// per-row log-sum-exp, so no 64 x 144 matrix is ever exchanged (flash-attention style).  Keys belong to up to four windows:
// every window writes its dK/dV rows to a per-window scratch and a fold kernel sums them per pixel in fixed order
// (deterministic); the relative-position-bias table gradient (361 x 6) is accumulated in shared memory per CTA.
#include <cuda_bf16.h>

#include "common.h"
#include "ops.h"

namespace srb {

namespace {

constexpr int OA_THREADS = 192, OA_NK = 144, OA_TAB = 361;
constexpr int OA_Q_BYTES = 64 * 96 * 2, OA_KV_BYTES = OA_NK * 192 * 2, OA_DO_BYTES = 64 * 96 * 2;

struct OcaArgs {
    const bf16* qkv;    // [B,H,W,288]
    const float* bias;  // [6][64][144]
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
__device__ __forceinline__ void oa_load16(const uint8_t* p, float* f) {
    oa_unpack8(*(const uint4*)p, f);
    oa_unpack8(*(const uint4*)(p + 16), f + 8);
}

// stage q rows [64][96], k|v rows [144][192] (zero outside the image) and optionally dO rows [64][96]
__device__ __forceinline__ void oa_load_window(const OcaArgs& a, int win, uint8_t* sQ, uint8_t* sKV, uint8_t* sDO) {
    const int nwx = a.W / 8, nwy = a.H / 8;
    const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
    for (int i = threadIdx.x; i < 64 * 12; i += OA_THREADS) {
        const int n = i / 12, u = i % 12;
        const size_t pix = ((size_t)b * a.H + wy * 8 + (n >> 3)) * a.W + wx * 8 + (n & 7);
        *(uint4*)(sQ + n * 192 + u * 16) = __ldg((const uint4*)(a.qkv + pix * 288) + u);
        if (sDO) *(uint4*)(sDO + n * 192 + u * 16) = __ldg((const uint4*)(a.dout + pix * 96) + u);
    }
    for (int i = threadIdx.x; i < OA_NK * 24; i += OA_THREADS) {
        const int n = i / 24, u = i % 24;
        const int y = wy * 8 - 2 + n / 12, x = wx * 8 - 2 + n % 12;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (y >= 0 && y < a.H && x >= 0 && x < a.W)
            v = __ldg((const uint4*)(a.qkv + (((size_t)b * a.H + y) * a.W + x) * 288 + 96) + u);
        *(uint4*)(sKV + n * 384 + u * 16) = v;
    }
    __syncthreads();
}

// softmax probabilities of query row r of head h: p[144]; returns the row's log-sum-exp
__device__ __forceinline__ float oa_attn_row(const OcaArgs& a, const uint8_t* sQ, const uint8_t* sKV, int h, int r, float* p) {
    float q[16];
    oa_load16(sQ + r * 192 + h * 32, q);
    const float* brow = a.bias + ((size_t)h * 64 + r) * OA_NK;
    float mx = -1e30f;
#pragma unroll
    for (int j = 0; j < OA_NK; ++j) {
        float k[16];
        oa_load16(sKV + j * 384 + h * 32, k);
        float s = 0.f;
#pragma unroll
        for (int d = 0; d < 16; ++d) s = fmaf(q[d], k[d], s);
        s = fmaf(s, 0.25f, __ldg(brow + j));
        p[j] = s;
        mx = fmaxf(mx, s);
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < OA_NK; ++j) {
        p[j] = __expf(p[j] - mx);
        sum += p[j];
    }
    const float inv = 1.f / sum;
#pragma unroll
    for (int j = 0; j < OA_NK; ++j) p[j] *= inv;
    return mx + __logf(sum);
}

__device__ __forceinline__ size_t oa_qpix(const OcaArgs& a, int win, int n) {
    const int nwx = a.W / 8, nwy = a.H / 8;
    const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
    return ((size_t)b * a.H + wy * 8 + (n >> 3)) * a.W + wx * 8 + (n & 7);
}

__global__ void __launch_bounds__(OA_THREADS)
ocab_attn_fwd_kernel(const OcaArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* sQ = smem;
    uint8_t* sKV = sQ + OA_Q_BYTES;
    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int win = blockIdx.x; win < a.nwin; win += gridDim.x) {
        __syncthreads();
        oa_load_window(a, win, sQ, sKV, nullptr);
        for (int rr = 0; rr < 2; ++rr) {
            const int r = lane + 32 * rr;
            float p[OA_NK], o[16];
            oa_attn_row(a, sQ, sKV, h, r, p);
#pragma unroll
            for (int d = 0; d < 16; ++d) o[d] = 0.f;
#pragma unroll
            for (int j = 0; j < OA_NK; ++j) {
                float v[16];
                oa_load16(sKV + j * 384 + 192 + h * 32, v);
#pragma unroll
                for (int d = 0; d < 16; ++d) o[d] = fmaf(p[j], v[d], o[d]);
            }
            uint4* op = (uint4*)(a.out + oa_qpix(a, win, r) * 96 + h * 16);
            op[0] = oa_pack8(o);
            op[1] = oa_pack8(o + 8);
        }
    }
}

__global__ void __launch_bounds__(OA_THREADS)
ocab_attn_bwd_kernel(const OcaArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* sQ = smem;
    uint8_t* sKV = sQ + OA_Q_BYTES;
    uint8_t* sDO = sKV + OA_KV_BYTES;
    float* sLse = (float*)(sDO + OA_DO_BYTES);  // [6][64]
    float* sDelta = sLse + 6 * 64;              // [6][64]
    float* sTab = sDelta + 6 * 64;              // [6][361]
    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 6 * OA_TAB; i += OA_THREADS) sTab[i] = 0.f;
    for (int win = blockIdx.x; win < a.nwin; win += gridDim.x) {
        __syncthreads();
        oa_load_window(a, win, sQ, sKV, sDO);
        // ---- pass 1 (query-owned): P, delta = sum_j P dP, dS = P (dP - delta), dQ = 0.25 dS K, bias-table gradient
        for (int rr = 0; rr < 2; ++rr) {
            const int r = lane + 32 * rr;
            float p[OA_NK], go[16], dq[16];
            const float lse = oa_attn_row(a, sQ, sKV, h, r, p);
            oa_load16(sDO + r * 192 + h * 32, go);
            float delta = 0.f;
#pragma unroll
            for (int j = 0; j < OA_NK; ++j) {
                float v[16];
                oa_load16(sKV + j * 384 + 192 + h * 32, v);
                float dp = 0.f;
#pragma unroll
                for (int d = 0; d < 16; ++d) dp = fmaf(go[d], v[d], dp);
                delta = fmaf(p[j], dp, delta);
            }
#pragma unroll
            for (int d = 0; d < 16; ++d) dq[d] = 0.f;
            const int* irow = a.index + r * OA_NK;
#pragma unroll
            for (int j = 0; j < OA_NK; ++j) {
                float v[16], k[16];
                oa_load16(sKV + j * 384 + 192 + h * 32, v);
                oa_load16(sKV + j * 384 + h * 32, k);
                float dp = 0.f;
#pragma unroll
                for (int d = 0; d < 16; ++d) dp = fmaf(go[d], v[d], dp);
                const float s = p[j] * (dp - delta);
#pragma unroll
                for (int d = 0; d < 16; ++d) dq[d] = fmaf(s, k[d], dq[d]);
                atomicAdd(&sTab[h * OA_TAB + __ldg(irow + j)], s);
            }
#pragma unroll
            for (int d = 0; d < 16; ++d) dq[d] *= 0.25f;
            uint4* op = (uint4*)(a.out + oa_qpix(a, win, r) * 288 + h * 16);
            op[0] = oa_pack8(dq);
            op[1] = oa_pack8(dq + 8);
            sLse[h * 64 + r] = lse;
            sDelta[h * 64 + r] = delta;
        }
        __syncwarp();  // lse / delta of this head are read by the same warp only
        // ---- pass 2 (key-owned): P recomputed from lse; dK_j = 0.25 sum_r dS[r][j] q_r;  dV_j = sum_r P[r][j] dO_r
        for (int jj = 0; jj < 5; ++jj) {
            const int j = lane + 32 * jj;
            if (j >= OA_NK) break;
            float k[16], v[16], dk[16], dv[16];
            oa_load16(sKV + j * 384 + h * 32, k);
            oa_load16(sKV + j * 384 + 192 + h * 32, v);
#pragma unroll
            for (int d = 0; d < 16; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
            const float* bcol = a.bias + (size_t)h * 64 * OA_NK + j;
#pragma unroll 2
            for (int r = 0; r < 64; ++r) {
                float q[16], go[16];
                oa_load16(sQ + r * 192 + h * 32, q);
                oa_load16(sDO + r * 192 + h * 32, go);
                float s = 0.f, dp = 0.f;
#pragma unroll
                for (int d = 0; d < 16; ++d) {
                    s = fmaf(q[d], k[d], s);
                    dp = fmaf(go[d], v[d], dp);
                }
                const float pp = __expf(fmaf(s, 0.25f, __ldg(bcol + r * OA_NK)) - sLse[h * 64 + r]);
                const float ds = pp * (dp - sDelta[h * 64 + r]);
#pragma unroll
                for (int d = 0; d < 16; ++d) {
                    dk[d] = fmaf(ds, q[d], dk[d]);
                    dv[d] = fmaf(pp, go[d], dv[d]);
                }
            }
#pragma unroll
            for (int d = 0; d < 16; ++d) dk[d] *= 0.25f;
            bf16* row = a.dkv_win + ((size_t)win * OA_NK + j) * 192;
            uint4* ok = (uint4*)(row + h * 16);
            ok[0] = oa_pack8(dk);
            ok[1] = oa_pack8(dk + 8);
            uint4* ov = (uint4*)(row + 96 + h * 16);
            ov[0] = oa_pack8(dv);
            ov[1] = oa_pack8(dv + 8);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 6 * OA_TAB; i += OA_THREADS) {
        const int hh = i / OA_TAB, t = i % OA_TAB;
        atomicAdd(a.dtable + t * 6 + hh, sTab[i]);
    }
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

constexpr int OA_FWD_SMEM = OA_Q_BYTES + OA_KV_BYTES;
constexpr int OA_BWD_SMEM = OA_Q_BYTES + OA_KV_BYTES + OA_DO_BYTES + (2 * 6 * 64 + 6 * OA_TAB) * 4;

int oa_check(int B, int H, int W) {
    if (B <= 0 || H % 8 || W % 8 || H < 8 || W < 8) {
        set_error("ocab_attn: H=%d W=%d must be multiples of 8", H, W);
        return SRB_E_SHAPE;
    }
    return SRB_OK;
}

}  // namespace

size_t ocab_attn_bwd_scratch_bytes(int B, int H, int W) { return (size_t)B * (H / 8) * (W / 8) * OA_NK * 192 * sizeof(bf16); }

int ocab_attn_fwd(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, cudaStream_t s) {
    SRB_TRY(oa_check(B, H, W));
    OcaArgs a{};
    a.qkv = qkv; a.bias = bias; a.out = out; a.B = B; a.H = H; a.W = W; a.nwin = B * (H / 8) * (W / 8);
    SRB_CUDA(cudaFuncSetAttribute(ocab_attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OA_FWD_SMEM));
    const int grid = a.nwin < num_sms() * 3 ? a.nwin : num_sms() * 3;
    ocab_attn_fwd_kernel<<<grid, OA_THREADS, OA_FWD_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int ocab_attn_bwd(const bf16* qkv, const float* bias, const int* index, const bf16* dout, bf16* dqkv, float* dtable,
                  void* scratch, int B, int H, int W, cudaStream_t s) {
    SRB_TRY(oa_check(B, H, W));
    OcaArgs a{};
    a.qkv = qkv; a.bias = bias; a.index = index; a.dout = dout; a.out = dqkv; a.dkv_win = (bf16*)scratch; a.dtable = dtable;
    a.B = B; a.H = H; a.W = W; a.nwin = B * (H / 8) * (W / 8);
    SRB_CUDA(cudaFuncSetAttribute(ocab_attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OA_BWD_SMEM));
    const int grid = a.nwin < num_sms() * 2 ? a.nwin : num_sms() * 2;
    ocab_attn_bwd_kernel<<<grid, OA_THREADS, OA_BWD_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    const size_t n = (size_t)B * H * W * 24;
    size_t g = (n + 255) / 256;
    if (g > (size_t)num_sms() * 16) g = (size_t)num_sms() * 16;
    ocab_fold_kernel<<<(int)g, 256, 0, s>>>((const bf16*)scratch, dqkv, B, H, W);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
