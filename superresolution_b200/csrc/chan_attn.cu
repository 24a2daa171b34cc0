// ChannelAttention of CAB (upstream hat_arch.py: x * sigmoid(conv1x1(relu(conv1x1(avgpool(x)))))), forward and backward,
// as un-fused operators of the stage-2 training path.  The inference path keeps the fused form
// (per-tile sums in the cab.2 conv epilogue + cab_finish + the scale folded into mlp_tc).
//   y: bf16 [B, HW, 96];  pooled[b][c] = mean_px y;  hid = relu(W1 pooled + b1) (3);  s = sigmoid(W2 hid + b2) (96);
//   out = y * s[b][c].
// Backward of out w.r.t. y has a direct term (g * s) and a pooled term ((1/HW) * W1^T relu'(.) W2^T (s(1-s) * sum_px g*y)).
// All reductions are two-level with fixed order (deterministic).
#include <cuda_bf16.h>

#include "common.h"
#include "ops.h"

namespace srb {

namespace {

constexpr int CA_C = 96, CA_HID = 3, CA_ROWS = 256;  // pixels per partial-sum block

// partial[b][blk][c] = sum over the block's pixels of a[px][c] (* b2[px][c] when given)
__global__ void __launch_bounds__(192)
ca_colsum_kernel(const bf16* __restrict__ a, const bf16* __restrict__ m, int HW, float* __restrict__ partial) {
    __shared__ float sm[16][CA_C];
    const int cgp = threadIdx.x % 12, pl = threadIdx.x / 12;  // 12 channel groups x 16 pixel lanes
    const int b = blockIdx.y;
    const int p0 = blockIdx.x * CA_ROWS, p1 = min(HW, p0 + CA_ROWS);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int p = p0 + pl; p < p1; p += 16) {
        const size_t off = ((size_t)b * HW + p) * CA_C + cgp * 8;
        const uint4 ra = __ldg((const uint4*)(a + off));
        const __nv_bfloat162* ha = (const __nv_bfloat162*)&ra;
        if (m) {
            const uint4 rm = __ldg((const uint4*)(m + off));
            const __nv_bfloat162* hm = (const __nv_bfloat162*)&rm;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 fa = __bfloat1622float2(ha[k]), fm = __bfloat1622float2(hm[k]);
                acc[2 * k] = fmaf(fa.x, fm.x, acc[2 * k]);
                acc[2 * k + 1] = fmaf(fa.y, fm.y, acc[2 * k + 1]);
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 fa = __bfloat1622float2(ha[k]);
                acc[2 * k] += fa.x;
                acc[2 * k + 1] += fa.y;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[pl][cgp * 8 + j] = acc[j];
    __syncthreads();
    if (threadIdx.x < CA_C) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 16; ++i) s += sm[i][threadIdx.x];
        partial[((size_t)b * gridDim.x + blockIdx.x) * CA_C + threadIdx.x] = s;
    }
}

// one block per image: pooled = mean, hid, s.  Saves pooled[b][96], hid[b][3], s[b][96] for the backward.
__global__ void __launch_bounds__(CA_C)
ca_fwd_kernel(const float* __restrict__ partial, int nblk, float inv_hw, const float* __restrict__ w1,
              const float* __restrict__ b1, const float* __restrict__ w2, const float* __restrict__ b2,
              float* __restrict__ pooled, float* __restrict__ hid, float* __restrict__ s) {
    __shared__ float sp[CA_C], sh[CA_HID];
    const int b = blockIdx.x, c = threadIdx.x;
    float t = 0.f;
    for (int i = 0; i < nblk; ++i) t += partial[((size_t)b * nblk + i) * CA_C + c];
    sp[c] = t * inv_hw;
    pooled[b * CA_C + c] = sp[c];
    __syncthreads();
    if (c < CA_HID) {
        float a = b1[c];
        for (int k = 0; k < CA_C; ++k) a = fmaf(w1[c * CA_C + k], sp[k], a);
        sh[c] = fmaxf(a, 0.f);
        hid[b * CA_HID + c] = sh[c];
    }
    __syncthreads();
    float z = b2[c];
#pragma unroll
    for (int k = 0; k < CA_HID; ++k) z = fmaf(w2[c * CA_HID + k], sh[k], z);
    s[b * CA_C + c] = 1.f / (1.f + expf(-z));
}

// out[b][px][c] = a * y[b][px][c] * s[b][c] (+ add[b][px][c] / HW-broadcast term dp[b][c] * inv_hw when given)
__global__ void __launch_bounds__(256)
ca_scale_kernel(bf16* __restrict__ out, const bf16* __restrict__ y, const float* __restrict__ s, float a,
                const float* __restrict__ dp, float inv_hw, int HW, size_t ntok) {
    const size_t n = ntok * 12;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t tok = i / 12;
        const int c = (int)(i - tok * 12) * 8;
        const int b = (int)(tok / HW);
        const uint4 raw = __ldg((const uint4*)(y + tok * CA_C + c));
        const __nv_bfloat162* h = (const __nv_bfloat162*)&raw;
        const float4 s0 = *(const float4*)(s + b * CA_C + c), s1 = *(const float4*)(s + b * CA_C + c + 4);
        const float sv[8] = {a * s0.x, a * s0.y, a * s0.z, a * s0.w, a * s1.x, a * s1.y, a * s1.z, a * s1.w};
        float add[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (dp) {
            const float4 d0 = *(const float4*)(dp + b * CA_C + c), d1 = *(const float4*)(dp + b * CA_C + c + 4);
            add[0] = d0.x * inv_hw; add[1] = d0.y * inv_hw; add[2] = d0.z * inv_hw; add[3] = d0.w * inv_hw;
            add[4] = d1.x * inv_hw; add[5] = d1.y * inv_hw; add[6] = d1.z * inv_hw; add[7] = d1.w * inv_hw;
        }
        uint4 o;
        __nv_bfloat162* oh = (__nv_bfloat162*)&o;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 f = __bfloat1622float2(h[k]);
            oh[k] = __floats2bfloat162_rn(fmaf(f.x, sv[2 * k], add[2 * k]), fmaf(f.y, sv[2 * k + 1], add[2 * k + 1]));
        }
        *(uint4*)(out + tok * CA_C + c) = o;
    }
}

// ds[b][c] = sum_blk partial[b][blk][c], written over partial[b][0][c]: eight block slices per (image, channel) folded
// in a fixed order (the single-block kernel below walked all B * nblk partials serially: 250 us at 16 x 256 blocks)
__global__ void __launch_bounds__(CA_C * 8)
ca_fold_kernel(float* __restrict__ partial, int nblk) {
    __shared__ float sm[8][CA_C];
    const int c = threadIdx.x % CA_C, j = threadIdx.x / CA_C, b = blockIdx.x;
    float* pb = partial + (size_t)b * nblk * CA_C;
    float t = 0.f;
    for (int i = j; i < nblk; i += 8) t += pb[(size_t)i * CA_C + c];
    sm[j][c] = t;
    __syncthreads();  // every partial of this image has been read
    if (j == 0) {
        float v = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) v += sm[k][c];
        pb[c] = v;
    }
}

// single block: for every image ds[b][c] = g_scale * sum_px g*y (folded by ca_fold_kernel) -> dz = ds * s(1-s) -> dhid -> dpooled[b][c]; parameter gradients
// summed over the images in order (deterministic).
__global__ void __launch_bounds__(CA_C)
ca_bwd_kernel(const float* __restrict__ partial, int nblk, float g_scale, int B, const float* __restrict__ w1,
              const float* __restrict__ w2, const float* __restrict__ pooled, const float* __restrict__ hid,
              const float* __restrict__ s, float* __restrict__ dpooled, float* __restrict__ dw1, float* __restrict__ db1,
              float* __restrict__ dw2, float* __restrict__ db2, int accumulate) {
    __shared__ float sdz[CA_C], sdh[CA_HID];
    const int c = threadIdx.x;
    float gw2[CA_HID] = {0.f, 0.f, 0.f}, gb2 = 0.f;
    float gw1[CA_HID] = {0.f, 0.f, 0.f};  // dW1[k][c] for k = 0..2 (this thread's column c)
    float gb1 = 0.f;                       // db1[c] for c < 3
    for (int b = 0; b < B; ++b) {
        const float ds = g_scale * partial[(size_t)b * nblk * CA_C + c];
        const float sv = s[b * CA_C + c];
        const float dz = ds * sv * (1.f - sv);
        sdz[c] = dz;
        gb2 += dz;
#pragma unroll
        for (int k = 0; k < CA_HID; ++k) gw2[k] = fmaf(dz, hid[b * CA_HID + k], gw2[k]);
        __syncthreads();
        if (c < CA_HID) {
            float a = 0.f;
            for (int k = 0; k < CA_C; ++k) a = fmaf(w2[k * CA_HID + c], sdz[k], a);
            a = hid[b * CA_HID + c] > 0.f ? a : 0.f;  // ReLU backward
            sdh[c] = a;
            gb1 += a;
        }
        __syncthreads();
        float dpv = 0.f;
#pragma unroll
        for (int k = 0; k < CA_HID; ++k) {
            dpv = fmaf(w1[k * CA_C + c], sdh[k], dpv);
            gw1[k] = fmaf(sdh[k], pooled[b * CA_C + c], gw1[k]);
        }
        dpooled[b * CA_C + c] = dpv;
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < CA_HID; ++k) {
        float* o2 = dw2 + c * CA_HID + k;
        *o2 = accumulate ? *o2 + gw2[k] : gw2[k];
        float* o1 = dw1 + k * CA_C + c;
        *o1 = accumulate ? *o1 + gw1[k] : gw1[k];
    }
    db2[c] = accumulate ? db2[c] + gb2 : gb2;
    if (c < CA_HID) db1[c] = accumulate ? db1[c] + gb1 : gb1;
}

int ca_blocks(int HW) { return (HW + CA_ROWS - 1) / CA_ROWS; }
int ca_grid(size_t n) {
    size_t g = (n + 255) / 256;
    const size_t cap = (size_t)num_sms() * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

// scratch layout (floats): partial[B][nblk][96]
size_t chan_attn_scratch_bytes(int B, int HW) { return (size_t)B * ca_blocks(HW) * CA_C * sizeof(float); }

int chan_attn_fwd(const bf16* y, const float* w1, const float* b1, const float* w2, const float* b2, bf16* out,
                  float* pooled, float* hid, float* s, void* scratch, int B, int HW, cudaStream_t st) {
    const int nblk = ca_blocks(HW);
    ca_colsum_kernel<<<dim3(nblk, B), 192, 0, st>>>(y, nullptr, HW, (float*)scratch);
    SRB_CUDA(cudaGetLastError());
    ca_fwd_kernel<<<B, CA_C, 0, st>>>((const float*)scratch, nblk, 1.f / (float)HW, w1, b1, w2, b2, pooled, hid, s);
    SRB_CUDA(cudaGetLastError());
    if (out) {  // null: the caller folds s[b][c] into a later epilogue (res2_cs of the proj conv) instead of a pass over y
        ca_scale_kernel<<<ca_grid((size_t)B * HW * 12), 256, 0, st>>>(out, y, s, 1.f, nullptr, 0.f, HW, (size_t)B * HW);
        SRB_CUDA(cudaGetLastError());
    }
    return SRB_OK;
}

int chan_attn_bwd(const bf16* y, const bf16* g, const float* w1, const float* w2, const float* pooled, const float* hid,
                  const float* s, bf16* dy, float* dw1, float* db1, float* dw2, float* db2, float* dpooled,
                  void* scratch, int B, int HW, int accumulate, float g_scale, cudaStream_t st) {
    const int nblk = ca_blocks(HW);
    ca_colsum_kernel<<<dim3(nblk, B), 192, 0, st>>>(g, y, HW, (float*)scratch);
    SRB_CUDA(cudaGetLastError());
    ca_fold_kernel<<<B, CA_C * 8, 0, st>>>((float*)scratch, nblk);
    SRB_CUDA(cudaGetLastError());
    ca_bwd_kernel<<<1, CA_C, 0, st>>>((const float*)scratch, nblk, g_scale, B, w1, w2, pooled, hid, s, dpooled, dw1, db1, dw2,
                                      db2, accumulate);
    SRB_CUDA(cudaGetLastError());
    ca_scale_kernel<<<ca_grid((size_t)B * HW * 12), 256, 0, st>>>(dy, g, s, g_scale, dpooled, 1.f / (float)HW, HW, (size_t)B * HW);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
