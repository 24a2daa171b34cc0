// (Shifted-)window multi-head self-attention core of HAB from a saved qkv tensor, forward AND backward — the un-fused
// operator pair of the stage-2 training path (upstream hat_arch.py WindowAttention.forward and its
// autograd, reached from src/architecture.py:78).  The inference path keeps the fused tcgen05 kernel (hab_tc.cu).
//
//   window = 8 x 8 tokens, 6 heads x 16 channels, logits = q.k * 0.25 + bias[h][i][j] (+ -100 across shift-mask regions),
//   softmax over the 64 keys, out = P v.  torch.roll / window_partition / window_reverse are index arithmetic.
//
// Correctness-first CUDA-core implementation (fp32 math on bf16 operands): one CTA per window, one warp per head, a lane
// owns query rows {lane, lane + 32}; the backward recomputes P, exchanges dS^T / P^T through shared memory (bf16, padded
// rows) for the key-side sums, and accumulates the relative-position-bias TABLE gradient (225 x 6) in shared memory
// across the windows of a persistent CTA (one global atomicAdd per entry and CTA).
#include <cuda_bf16.h>

#include "common.h"
#include "ops.h"

namespace srb {

namespace {

constexpr int WA_THREADS = 192;
constexpr int WA_QKV_BYTES = 64 * 288 * 2;
constexpr int WA_DO_BYTES = 64 * 96 * 2;
constexpr int WA_TP = 66;  // bf16 pitch of the transposed dS / P scratch rows (132 B: conflict-free row-per-lane reads)
constexpr int WA_SCR_BYTES = 2 * 64 * WA_TP * 2;  // per warp: dS^T and P^T
constexpr int WA_TAB = 225;

struct WinArgs {
    const bf16* qkv;    // [B,H,W,288]
    const float* bias;  // [6][64][64] relative position bias per head (already gathered from the table)
    const int* index;   // [64][64] relative_position_index (backward only)
    const bf16* dout;   // [B,H,W,96]   (backward)
    bf16* out;          // forward: [B,H,W,96];  backward: dqkv [B,H,W,288]
    float* dtable;      // [225][6] (backward, accumulated with atomicAdd)
    int B, H, W, shift, nwin;
};

__device__ __forceinline__ void unpack8(const uint4& raw, float* f) {
    const __nv_bfloat162* h = (const __nv_bfloat162*)&raw;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float2 v = __bfloat1622float2(h[k]);
        f[2 * k] = v.x;
        f[2 * k + 1] = v.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
    uint4 raw;
    __nv_bfloat162* h = (__nv_bfloat162*)&raw;
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(f[2 * k], f[2 * k + 1]);
    return raw;
}

// pixel index of window token n (cyclic shift = index math) and its shift-mask region id
__device__ __forceinline__ size_t win_token(const WinArgs& a, int win, int n, int& rid) {
    const int nwx = a.W / 8, nwy = a.H / 8;
    const int wx = win % nwx, wy = (win / nwx) % nwy, b = win / (nwx * nwy);
    const int ys = wy * 8 + (n >> 3), xs = wx * 8 + (n & 7);  // rolled-image coordinates
    int y = ys + a.shift, x = xs + a.shift;                   // torch.roll(-shift): rolled[i] = orig[i + shift]
    if (y >= a.H) y -= a.H;
    if (x >= a.W) x -= a.W;
    const int rh = ys < a.H - 8 ? 0 : (ys < a.H - 4 ? 1 : 2);
    const int rw = xs < a.W - 8 ? 0 : (xs < a.W - 4 ? 1 : 2);
    rid = rh * 3 + rw;
    return ((size_t)b * a.H + y) * a.W + x;
}

// stage the window's qkv rows (and optionally dOut rows) in shared memory; sPix / sRid per token
__device__ __forceinline__ void load_window(const WinArgs& a, int win, uint8_t* sQKV, uint8_t* sDO, size_t* sPix, int* sRid) {
    if (threadIdx.x < 64) {
        int rid;
        sPix[threadIdx.x] = win_token(a, win, threadIdx.x, rid);
        sRid[threadIdx.x] = rid;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * 36; i += WA_THREADS) {
        const int n = i / 36, u = i % 36;
        *(uint4*)(sQKV + n * 576 + u * 16) = __ldg((const uint4*)(a.qkv + sPix[n] * 288) + u);
    }
    if (sDO) {
        for (int i = threadIdx.x; i < 64 * 12; i += WA_THREADS) {
            const int n = i / 12, u = i % 12;
            *(uint4*)(sDO + n * 192 + u * 16) = __ldg((const uint4*)(a.dout + sPix[n] * 96) + u);
        }
    }
    __syncthreads();
}

// softmax probabilities of query row r of head h: p[64] (fp32 registers)
__device__ __forceinline__ void attn_row(const WinArgs& a, const uint8_t* sQKV, const int* sRid, int h, int r, float* q,
                                         float* p) {
    unpack8(*(const uint4*)(sQKV + r * 576 + h * 32), q);
    unpack8(*(const uint4*)(sQKV + r * 576 + h * 32 + 16), q + 8);
    const float* brow = a.bias + ((size_t)h * 64 + r) * 64;
    const int rid = sRid[r];
    float mx = -1e30f;
#pragma unroll
    for (int j = 0; j < 64; ++j) {
        float k[16];
        unpack8(*(const uint4*)(sQKV + j * 576 + 192 + h * 32), k);
        unpack8(*(const uint4*)(sQKV + j * 576 + 192 + h * 32 + 16), k + 8);
        float s = 0.f;
#pragma unroll
        for (int d = 0; d < 16; ++d) s = fmaf(q[d], k[d], s);
        s = fmaf(s, 0.25f, __ldg(brow + j));
        if (a.shift > 0 && sRid[j] != rid) s -= 100.f;
        p[j] = s;
        mx = fmaxf(mx, s);
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 64; ++j) {
        p[j] = __expf(p[j] - mx);
        sum += p[j];
    }
    const float inv = 1.f / sum;
#pragma unroll
    for (int j = 0; j < 64; ++j) p[j] *= inv;
}

__global__ void __launch_bounds__(WA_THREADS)
win_attn_fwd_kernel(const WinArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* sQKV = smem;
    size_t* sPix = (size_t*)(smem + WA_QKV_BYTES);
    int* sRid = (int*)(sPix + 64);
    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int win = blockIdx.x; win < a.nwin; win += gridDim.x) {
        __syncthreads();  // previous window fully consumed
        load_window(a, win, sQKV, nullptr, sPix, sRid);
        for (int rr = 0; rr < 2; ++rr) {
            const int r = lane + 32 * rr;
            float q[16], p[64], o[16];
            attn_row(a, sQKV, sRid, h, r, q, p);
#pragma unroll
            for (int d = 0; d < 16; ++d) o[d] = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
                float v[16];
                unpack8(*(const uint4*)(sQKV + j * 576 + 384 + h * 32), v);
                unpack8(*(const uint4*)(sQKV + j * 576 + 384 + h * 32 + 16), v + 8);
#pragma unroll
                for (int d = 0; d < 16; ++d) o[d] = fmaf(p[j], v[d], o[d]);
            }
            uint4* op = (uint4*)(a.out + sPix[r] * 96 + h * 16);
            op[0] = pack8(o);
            op[1] = pack8(o + 8);
        }
    }
}

__global__ void __launch_bounds__(WA_THREADS)
win_attn_bwd_kernel(const WinArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* sQKV = smem;
    uint8_t* sDO = sQKV + WA_QKV_BYTES;
    uint8_t* sScr = sDO + WA_DO_BYTES;                       // [6 warps][dS^T | P^T]
    float* sTab = (float*)(sScr + 6 * WA_SCR_BYTES);         // [6][225]
    size_t* sPix = (size_t*)(sTab + 6 * WA_TAB + 2);         // (+2 floats keeps 8-byte alignment: 6*225 is even)
    int* sRid = (int*)(sPix + 64);
    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    bf16* dsT = (bf16*)(sScr + h * WA_SCR_BYTES);
    bf16* pT = dsT + 64 * WA_TP;
    for (int i = threadIdx.x; i < 6 * WA_TAB; i += WA_THREADS) sTab[i] = 0.f;
    for (int win = blockIdx.x; win < a.nwin; win += gridDim.x) {
        __syncthreads();  // previous window fully consumed (and sTab zeroed on the first pass)
        load_window(a, win, sQKV, sDO, sPix, sRid);
        // ---- pass 1: query rows. P recomputed; dP = dO V^T; dS = P (dP - sum_j P dP); dQ = 0.25 dS K
        for (int rr = 0; rr < 2; ++rr) {
            const int r = lane + 32 * rr;
            float q[16], p[64], go[16];
            attn_row(a, sQKV, sRid, h, r, q, p);
            unpack8(*(const uint4*)(sDO + r * 192 + h * 32), go);
            unpack8(*(const uint4*)(sDO + r * 192 + h * 32 + 16), go + 8);
            float ds[64];
            float delta = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
                float v[16];
                unpack8(*(const uint4*)(sQKV + j * 576 + 384 + h * 32), v);
                unpack8(*(const uint4*)(sQKV + j * 576 + 384 + h * 32 + 16), v + 8);
                float dp = 0.f;
#pragma unroll
                for (int d = 0; d < 16; ++d) dp = fmaf(go[d], v[d], dp);
                ds[j] = dp;
                delta = fmaf(p[j], dp, delta);
            }
            float dq[16];
#pragma unroll
            for (int d = 0; d < 16; ++d) dq[d] = 0.f;
            const int* irow = a.index + r * 64;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
                const float s = p[j] * (ds[j] - delta);  // dS[r][j] = gradient of the logit (and of the bias entry)
                float k[16];
                unpack8(*(const uint4*)(sQKV + j * 576 + 192 + h * 32), k);
                unpack8(*(const uint4*)(sQKV + j * 576 + 192 + h * 32 + 16), k + 8);
#pragma unroll
                for (int d = 0; d < 16; ++d) dq[d] = fmaf(s, k[d], dq[d]);
                atomicAdd(&sTab[h * WA_TAB + __ldg(irow + j)], s);
                dsT[j * WA_TP + r] = __float2bfloat16_rn(s);
                pT[j * WA_TP + r] = __float2bfloat16_rn(p[j]);
            }
#pragma unroll
            for (int d = 0; d < 16; ++d) dq[d] *= 0.25f;
            uint4* op = (uint4*)(a.out + sPix[r] * 288 + h * 16);
            op[0] = pack8(dq);
            op[1] = pack8(dq + 8);
        }
        __syncwarp();  // dS^T / P^T of this head are complete (written and read by the same warp only)
        // ---- pass 2: key rows. dK_j = 0.25 sum_r dS[r][j] q_r;  dV_j = sum_r P[r][j] dO_r
        for (int jj = 0; jj < 2; ++jj) {
            const int j = lane + 32 * jj;
            float dk[16], dv[16];
#pragma unroll
            for (int d = 0; d < 16; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
#pragma unroll 4
            for (int r = 0; r < 64; ++r) {
                const float s = __bfloat162float(dsT[j * WA_TP + r]), pp = __bfloat162float(pT[j * WA_TP + r]);
                float q[16], go[16];
                unpack8(*(const uint4*)(sQKV + r * 576 + h * 32), q);
                unpack8(*(const uint4*)(sQKV + r * 576 + h * 32 + 16), q + 8);
                unpack8(*(const uint4*)(sDO + r * 192 + h * 32), go);
                unpack8(*(const uint4*)(sDO + r * 192 + h * 32 + 16), go + 8);
#pragma unroll
                for (int d = 0; d < 16; ++d) {
                    dk[d] = fmaf(s, q[d], dk[d]);
                    dv[d] = fmaf(pp, go[d], dv[d]);
                }
            }
#pragma unroll
            for (int d = 0; d < 16; ++d) dk[d] *= 0.25f;
            uint4* ok = (uint4*)(a.out + sPix[j] * 288 + 96 + h * 16);
            ok[0] = pack8(dk);
            ok[1] = pack8(dk + 8);
            uint4* ov = (uint4*)(a.out + sPix[j] * 288 + 192 + h * 16);
            ov[0] = pack8(dv);
            ov[1] = pack8(dv + 8);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 6 * WA_TAB; i += WA_THREADS) {
        const int hh = i / WA_TAB, t = i % WA_TAB;
        atomicAdd(a.dtable + t * 6 + hh, sTab[i]);
    }
}

constexpr int WA_FWD_SMEM = WA_QKV_BYTES + 64 * 8 + 64 * 4;
constexpr int WA_BWD_SMEM = WA_QKV_BYTES + WA_DO_BYTES + 6 * WA_SCR_BYTES + (6 * WA_TAB + 2) * 4 + 64 * 8 + 64 * 4;

int check_geom(int B, int H, int W, int shift) {
    if (B <= 0 || H % 8 || W % 8 || H < 8 || W < 8 || (shift != 0 && shift != 4)) {
        set_error("win_attn: H=%d W=%d must be multiples of 8, shift=%d must be 0 or 4", H, W, shift);
        return SRB_E_SHAPE;
    }
    return SRB_OK;
}

}  // namespace

int win_attn_fwd(const bf16* qkv, const float* bias, bf16* out, int B, int H, int W, int shift, cudaStream_t s) {
    SRB_TRY(check_geom(B, H, W, shift));
    WinArgs a;
    a.qkv = qkv; a.bias = bias; a.index = nullptr; a.dout = nullptr; a.out = out; a.dtable = nullptr;
    a.B = B; a.H = H; a.W = W; a.shift = shift; a.nwin = B * (H / 8) * (W / 8);
    const int grid = a.nwin < num_sms() * 4 ? a.nwin : num_sms() * 4;
    win_attn_fwd_kernel<<<grid, WA_THREADS, WA_FWD_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int win_attn_bwd(const bf16* qkv, const float* bias, const int* index, const bf16* dout, bf16* dqkv, float* dtable, int B,
                 int H, int W, int shift, cudaStream_t s) {
    SRB_TRY(check_geom(B, H, W, shift));
    WinArgs a;
    a.qkv = qkv; a.bias = bias; a.index = index; a.dout = dout; a.out = dqkv; a.dtable = dtable;
    a.B = B; a.H = H; a.W = W; a.shift = shift; a.nwin = B * (H / 8) * (W / 8);
    SRB_CUDA(cudaFuncSetAttribute(win_attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WA_BWD_SMEM));
    const int grid = a.nwin < num_sms() ? a.nwin : num_sms();
    win_attn_bwd_kernel<<<grid, WA_THREADS, WA_BWD_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
