// Fused token-wise MLP branch of HAB / OCAB (App. A.2.4-5):
//     y0 = t + cab * cs[b,:]          (HAB: attention branch + CAB * sigmoid(...) * conv_scale; OCAB: y0 = t)
//     y  = y0 + fc2( GELU( fc1( LayerNorm(y0) ) ) )
// One warp owns 16 tokens end to end: the combine and LayerNorm (fp32) happen in registers in MMA A-fragment layout,
// fc1 runs 16 hidden columns at a time, bias + exact-erf GELU is applied on the accumulator fragments, which are
// re-packed *in registers* as the A fragments of fc2 (accumulator -> operand reuse, no shared-memory round trip for
// the 192-wide hidden activation), and the fc2 accumulators receive bias + residual before the store.
// Weights (fc1 192x96, fc2 96x192, bf16) stay in shared memory for the whole persistent kernel; there is no block-level
// synchronisation in the main loop.  HBM traffic: read t (+cab), write y — the normalised tensor, the hidden tensor
// (403 MB at B=16) and the combine pass never touch memory.
#include "common.h"
#include "ops.h"

namespace srb {

namespace {

constexpr int P1 = 104;  // fc1 weight row pitch (bf16): 96 + 8  -> conflict-free ldmatrix
constexpr int P2 = 200;  // fc2 weight row pitch (bf16): 192 + 8
constexpr int MLP_SMEM = (192 * P1 + 96 * P2) * 2 + (192 + 96 + 96 + 96) * 4;

__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ uint32_t packbf(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *(uint32_t*)&h;
}
__device__ __forceinline__ float2 unpackbf(uint32_t v) { return __bfloat1622float2(*(const __nv_bfloat162*)&v); }

// erf by Abramowitz-Stegun 7.1.26 (|err| < 1.5e-7): exact-GELU semantics of nn.GELU()
__device__ __forceinline__ float gelu_exact(float x) {
    const float z = fabsf(x) * 0.70710678118654752f;
    const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
    float p = fmaf(1.061405429f, t, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    const float e = 1.0f - p * t * __expf(-z * z);
    return 0.5f * x * (1.0f + copysignf(e, x));
}

struct MlpArgs {
    const bf16* t;
    const bf16* cab;   // nullable
    const float* cs;   // [B][96], nullable with cab
    const float *ln_g, *ln_b;
    const bf16* w1;    // [192][96]
    const float* b1;
    const bf16* w2;    // [96][192]
    const float* b2;
    bf16* y;
    size_t ntok, tok_per_img;
};

__global__ void __launch_bounds__(256, 2)
mlp_fused_kernel(const MlpArgs a) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    bf16* sW1 = (bf16*)smem_raw;       // [192][P1]
    bf16* sW2 = sW1 + 192 * P1;        // [96][P2]
    float* sb1 = (float*)(sW2 + 96 * P2);
    float* sb2 = sb1 + 192;
    float* sg = sb2 + 96;
    float* sb = sg + 96;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tq = lane & 3;

    for (int i = tid; i < 192 * 12; i += 256) *(uint4*)(sW1 + (i / 12) * P1 + (i % 12) * 8) = __ldg((const uint4*)a.w1 + i);
    for (int i = tid; i < 96 * 24; i += 256) *(uint4*)(sW2 + (i / 24) * P2 + (i % 24) * 8) = __ldg((const uint4*)a.w2 + i);
    if (tid < 192) sb1[tid] = a.b1[tid];
    if (tid < 96) { sb2[tid] = a.b2[tid]; sg[tid] = a.ln_g[tid]; sb[tid] = a.ln_b[tid]; }
    __syncthreads();

    const size_t ngroups = a.ntok / 16;
    for (size_t grp = (size_t)blockIdx.x * 8 + warp; grp < ngroups; grp += (size_t)gridDim.x * 8) {
        const size_t ra = grp * 16 + g, rb = ra + 8;
        // ---- y0 = t + cab*cs in A-fragment layout: [ks][0]: cols 16ks+2tq,+1 ; [ks][1]: cols 16ks+8+2tq,+1 ----
        float va[6][4], vb[6][4];  // row a / row b: {c0, c0+1, c1, c1+1}
#pragma unroll
        for (int ks = 0; ks < 6; ++ks) {
            const int c0 = 16 * ks + 2 * tq, c1 = c0 + 8;
            const float2 a0 = unpackbf(*(const uint32_t*)(a.t + ra * 96 + c0));
            const float2 a1 = unpackbf(*(const uint32_t*)(a.t + ra * 96 + c1));
            const float2 b0 = unpackbf(*(const uint32_t*)(a.t + rb * 96 + c0));
            const float2 b1 = unpackbf(*(const uint32_t*)(a.t + rb * 96 + c1));
            va[ks][0] = a0.x; va[ks][1] = a0.y; va[ks][2] = a1.x; va[ks][3] = a1.y;
            vb[ks][0] = b0.x; vb[ks][1] = b0.y; vb[ks][2] = b1.x; vb[ks][3] = b1.y;
        }
        if (a.cab) {
            const float* cs = a.cs + (grp * 16 / a.tok_per_img) * 96;
#pragma unroll
            for (int ks = 0; ks < 6; ++ks) {
                const int c0 = 16 * ks + 2 * tq, c1 = c0 + 8;
                const float2 s0 = __ldg((const float2*)(cs + c0)), s1 = __ldg((const float2*)(cs + c1));
                const float2 a0 = unpackbf(*(const uint32_t*)(a.cab + ra * 96 + c0));
                const float2 a1 = unpackbf(*(const uint32_t*)(a.cab + ra * 96 + c1));
                const float2 b0 = unpackbf(*(const uint32_t*)(a.cab + rb * 96 + c0));
                const float2 b1 = unpackbf(*(const uint32_t*)(a.cab + rb * 96 + c1));
                va[ks][0] = fmaf(a0.x, s0.x, va[ks][0]); va[ks][1] = fmaf(a0.y, s0.y, va[ks][1]);
                va[ks][2] = fmaf(a1.x, s1.x, va[ks][2]); va[ks][3] = fmaf(a1.y, s1.y, va[ks][3]);
                vb[ks][0] = fmaf(b0.x, s0.x, vb[ks][0]); vb[ks][1] = fmaf(b0.y, s0.y, vb[ks][1]);
                vb[ks][2] = fmaf(b1.x, s1.x, vb[ks][2]); vb[ks][3] = fmaf(b1.y, s1.y, vb[ks][3]);
            }
        }
        // ---- LayerNorm statistics (fp32): each row is spread over the 4 lanes of a quad ----
        float sa = 0.f, sbm = 0.f;
#pragma unroll
        for (int ks = 0; ks < 6; ++ks)
#pragma unroll
            for (int e = 0; e < 4; ++e) { sa += va[ks][e]; sbm += vb[ks][e]; }
        sa += __shfl_xor_sync(0xffffffffu, sa, 1); sa += __shfl_xor_sync(0xffffffffu, sa, 2);
        sbm += __shfl_xor_sync(0xffffffffu, sbm, 1); sbm += __shfl_xor_sync(0xffffffffu, sbm, 2);
        const float ma = sa * (1.f / 96.f), mb = sbm * (1.f / 96.f);
        float qa = 0.f, qb = 0.f;
#pragma unroll
        for (int ks = 0; ks < 6; ++ks)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                qa += (va[ks][e] - ma) * (va[ks][e] - ma);
                qb += (vb[ks][e] - mb) * (vb[ks][e] - mb);
            }
        qa += __shfl_xor_sync(0xffffffffu, qa, 1); qa += __shfl_xor_sync(0xffffffffu, qa, 2);
        qb += __shfl_xor_sync(0xffffffffu, qb, 1); qb += __shfl_xor_sync(0xffffffffu, qb, 2);
        const float rsa = rsqrtf(qa * (1.f / 96.f) + 1e-5f), rsb = rsqrtf(qb * (1.f / 96.f) + 1e-5f);
        // ---- normalised A fragments (bf16) and the residual y0 (packed bf16, the precision the stream is stored in) ----
        uint32_t xn[6][4], y0a[6][2], y0b[6][2];
#pragma unroll
        for (int ks = 0; ks < 6; ++ks) {
            const int c0 = 16 * ks + 2 * tq, c1 = c0 + 8;
            const float2 g0 = *(const float2*)(sg + c0), g1 = *(const float2*)(sg + c1);
            const float2 e0 = *(const float2*)(sb + c0), e1 = *(const float2*)(sb + c1);
            xn[ks][0] = packbf((va[ks][0] - ma) * rsa * g0.x + e0.x, (va[ks][1] - ma) * rsa * g0.y + e0.y);
            xn[ks][1] = packbf((vb[ks][0] - mb) * rsb * g0.x + e0.x, (vb[ks][1] - mb) * rsb * g0.y + e0.y);
            xn[ks][2] = packbf((va[ks][2] - ma) * rsa * g1.x + e1.x, (va[ks][3] - ma) * rsa * g1.y + e1.y);
            xn[ks][3] = packbf((vb[ks][2] - mb) * rsb * g1.x + e1.x, (vb[ks][3] - mb) * rsb * g1.y + e1.y);
            y0a[ks][0] = packbf(va[ks][0], va[ks][1]); y0a[ks][1] = packbf(va[ks][2], va[ks][3]);
            y0b[ks][0] = packbf(vb[ks][0], vb[ks][1]); y0b[ks][1] = packbf(vb[ks][2], vb[ks][3]);
        }
        // ---- fc1 (16 hidden columns at a time) -> GELU -> fc2 accumulation ----
        float acc2[12][4];
#pragma unroll
        for (int n = 0; n < 12; ++n) acc2[n][0] = acc2[n][1] = acc2[n][2] = acc2[n][3] = 0.f;
#pragma unroll 1
        for (int kk = 0; kk < 12; ++kk) {
            float h0[4] = {0.f, 0.f, 0.f, 0.f}, h1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 6; ++ks) {
                uint32_t w[4];
                ldsm4(w, sW1 + (16 * kk + (lane & 7) + (lane >> 4) * 8) * P1 + 16 * ks + ((lane >> 3) & 1) * 8);
                mma16816(h0, xn[ks], w[0], w[1]);
                mma16816(h1, xn[ks], w[2], w[3]);
            }
            const int hc = 16 * kk + 2 * tq;
            const float2 ba = *(const float2*)(sb1 + hc), bb = *(const float2*)(sb1 + hc + 8);
            uint32_t hf[4];
            hf[0] = packbf(gelu_exact(h0[0] + ba.x), gelu_exact(h0[1] + ba.y));
            hf[1] = packbf(gelu_exact(h0[2] + ba.x), gelu_exact(h0[3] + ba.y));
            hf[2] = packbf(gelu_exact(h1[0] + bb.x), gelu_exact(h1[1] + bb.y));
            hf[3] = packbf(gelu_exact(h1[2] + bb.x), gelu_exact(h1[3] + bb.y));
#pragma unroll
            for (int n2 = 0; n2 < 6; ++n2) {
                uint32_t w[4];
                ldsm4(w, sW2 + (16 * n2 + (lane & 7) + (lane >> 4) * 8) * P2 + 16 * kk + ((lane >> 3) & 1) * 8);
                mma16816(acc2[2 * n2], hf, w[0], w[1]);
                mma16816(acc2[2 * n2 + 1], hf, w[2], w[3]);
            }
        }
        // ---- y = y0 + fc2 + b2 ----
#pragma unroll
        for (int ks = 0; ks < 6; ++ks) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int n = 2 * ks + half, c = 8 * n + 2 * tq;
                const float2 bb = *(const float2*)(sb2 + c);
                const float2 ya = unpackbf(y0a[ks][half]), yb = unpackbf(y0b[ks][half]);
                *(uint32_t*)(a.y + ra * 96 + c) = packbf(acc2[n][0] + bb.x + ya.x, acc2[n][1] + bb.y + ya.y);
                *(uint32_t*)(a.y + rb * 96 + c) = packbf(acc2[n][2] + bb.x + yb.x, acc2[n][3] + bb.y + yb.y);
            }
        }
    }
}

}  // namespace

int mlp_fused(const bf16* t, const bf16* cab, const float* cs, const float* ln_g, const float* ln_b, const bf16* w1,
              const float* b1, const bf16* w2, const float* b2, bf16* y, size_t ntok, size_t tok_per_img,
              cudaStream_t s) {
    if (ntok % 16 || tok_per_img % 16) { set_error("mlp_fused: token counts must be multiples of 16"); return SRB_E_SHAPE; }
    static bool attr = false;
    if (!attr) {
        SRB_CUDA(cudaFuncSetAttribute(mlp_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MLP_SMEM));
        attr = true;
    }
    MlpArgs a;
    a.t = t; a.cab = cab; a.cs = cs; a.ln_g = ln_g; a.ln_b = ln_b; a.w1 = w1; a.b1 = b1; a.w2 = w2; a.b2 = b2; a.y = y;
    a.ntok = ntok; a.tok_per_img = tok_per_img;
    size_t need = (ntok / 16 + 7) / 8;
    int grid = need < 296 ? (int)need : 296;
    mlp_fused_kernel<<<grid, 256, MLP_SMEM, s>>>(a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
