// Hardware probe (development tool, not product code): establishes the tcgen05 facts the tensor-core attention
// core of hab_tc / ocab_tc relies on.  One process per test id (a faulting form cannot poison later tests).
//   0  SS  M=64  N=64 K=16, D at lane offset 0      -> which TMEM lanes hold which rows
//   1  SS  M=64  N=64 K=16, D at lane offset 16
//   2  TS  M=128 N=64 K=16, A (bf16 pairs) written to TMEM with tcgen05.st 32x32b
//   3  TS  M=64  N=64 K=16, A lanes 0.., D lane offset 0
//   4  TS  M=64  N=64 K=16, A at lane offset 16, D at lane offset 16
//   5  SS  M=128 N=16 K=16 with an MN-major B operand (V planes [dim/8][key][8 dims]); b_major bit = 1
//   6  SS  M=128 N=16 K=64 MN-major B, 4 k-steps (LBO = 128: next 8 keys)
//   7  rates: cycles per MMA for the shapes of the attention core
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I superresolution_b200/csrc -o tools/bin/probe_attn tools/probe_attn.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ptx.cuh"
using namespace srb;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); exit(2); } } while (0)

__device__ __forceinline__ void tmem_st_32x32b_x8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

constexpr int PP = 128 * 16 + 16;  // plane pitch of the product kernels
__device__ __forceinline__ uint32_t idesc_mn(uint32_t M, uint32_t N, int b_mn) {
    return umma_idesc_bf16_f32(M, N) | ((uint32_t)b_mn << 16);
}

// A[r][k] (K-major planes), B[n][k] K-major planes or V[k][n] MN-major planes; D dumped for all 128 lanes x 64 columns
__global__ void __launch_bounds__(128, 1) probe(int test, int pass, float* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA = smem;                 // 8 planes x PP
    uint8_t* sB = smem + 8 * PP;        // 8 planes x PP
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    for (int i = tid; i < 16 * PP / 4; i += 128) ((uint32_t*)smem)[i] = 0;
    __syncthreads();
    // A rows: pass 0 value = row id (all k); pass 1 value = k
    for (int i = tid; i < 128 * 64; i += 128) {
        const int r = i / 64, k = i % 64;
        const float v = pass == 0 ? (float)r : (float)k;
        *(__nv_bfloat16*)(sA + (k / 8) * PP + r * 16 + (k % 8) * 2) = __float2bfloat16(v);
    }
    if (test <= 4) {  // K-major B [n][k] = (n % 16 == k): D[r][n] = A[r][n % 16]
        for (int i = tid; i < 128 * 16; i += 128) {
            const int n = i / 16, k = i % 16;
            *(__nv_bfloat16*)(sB + (k / 8) * PP + n * 16 + (k % 8) * 2) = __float2bfloat16((n % 16) == k ? 1.f : 0.f);
        }
    } else {  // V[key][dim] planes [dim/8][key][8 dims]; V[k][n] = 16 * (k % 16) + n  (exact in bf16: <= 255)
        for (int i = tid; i < 128 * 16; i += 128) {
            const int k = i / 16, n = i % 16;
            *(__nv_bfloat16*)(sB + (n / 8) * PP + k * 16 + (n % 8) * 2) = __float2bfloat16((float)(16 * (k % 16) + n));
        }
        // A = one-hot: A[r][k] = (k == r % K)
        const int K = test == 5 ? 16 : 64;
        for (int i = tid; i < 128 * 64; i += 128) {
            const int r = i / 64, k = i % 64;
            *(__nv_bfloat16*)(sA + (k / 8) * PP + r * 16 + (k % 8) * 2) = __float2bfloat16(k == (r % K) ? 1.f : 0.f);
        }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = slot;
    // clear D region (columns 0..63 of every lane) and stage A in TMEM columns [256, 264) for the TS tests
    {
        uint32_t z[8] = {0x7fc00000u, 0x7fc00000u, 0x7fc00000u, 0x7fc00000u, 0x7fc00000u, 0x7fc00000u, 0x7fc00000u, 0x7fc00000u};  // NaN marks untouched cells
        for (int c = 0; c < 64; c += 8) tmem_st_32x32b_x8(tm + ((uint32_t)(warp * 32) << 16) + c, z);
        uint32_t a[8];
        const int r = warp * 32 + lane;
        for (int j = 0; j < 8; ++j) {
            const float lo = pass == 0 ? (float)r : (float)(2 * j), hi = pass == 0 ? (float)r : (float)(2 * j + 1);
            __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
            a[j] = *(uint32_t*)&h;
        }
        tmem_st_32x32b_x8(tm + ((uint32_t)(warp * 32) << 16) + 256, a);
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0) {
        if (elect_one()) {
            const uint32_t a_u = smem_u32(sA), b_u = smem_u32(sB);
            const uint64_t ad = umma_smem_desc(a_u, PP, 128, UMMA_SW_NONE), bd = umma_smem_desc(b_u, PP, 128, UMMA_SW_NONE);
            if (test == 0) umma_bf16_ss(tm, ad, bd, umma_idesc_bf16_f32(64, 64), 0);
            if (test == 1) umma_bf16_ss(tm + (16u << 16), ad, bd, umma_idesc_bf16_f32(64, 64), 0);
            if (test == 2) umma_bf16_ts(tm, tm + 256, bd, umma_idesc_bf16_f32(128, 64), 0);
            if (test == 3) umma_bf16_ts(tm, tm + 256, bd, umma_idesc_bf16_f32(64, 64), 0);
            if (test == 4) umma_bf16_ts(tm + (16u << 16), tm + (16u << 16) + 256, bd, umma_idesc_bf16_f32(64, 64), 0);
            if (test == 5)  // MN-major B: 8 dims contiguous, next dim group = next plane (SBO), next 8 keys = +128 B (LBO)
                umma_bf16_ss(tm, ad, umma_smem_desc(b_u, 128, PP, UMMA_SW_NONE), idesc_mn(128, 16, 1), 0);
            if (test == 6)
                for (int ks = 0; ks < 4; ++ks)
                    umma_bf16_ss(tm, umma_smem_desc(a_u + ks * 2 * PP, PP, 128, UMMA_SW_NONE),
                                 umma_smem_desc(b_u + ks * 16 * 16, 128, PP, UMMA_SW_NONE), idesc_mn(128, 16, 1), ks > 0);
            umma_commit(&bar);
        }
        __syncwarp();
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tm + ((uint32_t)(warp * 32) << 16) + c0, r);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 64 + c0 + j] = __uint_as_float(r[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}

// rates: one CTA per SM, lane 0 of warp 0 issues `iters` MMAs (descriptors precomputed, 8x unrolled)
template <int SHAPE>
__global__ void __launch_bounds__(128, 1) rates(int iters, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    fence_proxy_async_smem();
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tm = slot;
    if (threadIdx.x == 0) {
        const uint32_t a_u = smem_u32(smem), b_u = smem_u32(smem + 32 * 1024);
        const uint64_t ad = umma_smem_desc(a_u, PP, 128, UMMA_SW_NONE), bd = umma_smem_desc(b_u, PP, 128, UMMA_SW_NONE);
        const uint64_t bmn = umma_smem_desc(b_u, 128, PP, UMMA_SW_NONE);
        constexpr int M = (SHAPE == 1 || SHAPE == 5 || SHAPE == 12 || SHAPE == 13) ? 64 : 128;
        constexpr int N = SHAPE == 0 ? 64 : SHAPE == 1 ? 64 : SHAPE == 2 ? 128 : SHAPE == 3 ? 16 : SHAPE == 4 ? 32 : SHAPE == 5 ? 16
                        : SHAPE == 6 ? 16 : SHAPE == 7 ? 16 : SHAPE == 8 ? 96 : SHAPE == 9 ? 64 : SHAPE == 10 ? 256
                        : SHAPE == 11 ? 128 : SHAPE == 12 ? 16 : 24;
        constexpr bool TS = (SHAPE >= 3 && SHAPE <= 6) || (SHAPE >= 8 && SHAPE <= 10) || SHAPE == 12 || SHAPE == 13;
        constexpr bool MN = SHAPE == 6 || SHAPE == 12 || SHAPE == 13;
        const uint32_t idesc = idesc_mn(M, N, MN ? 1 : 0);
        const uint64_t b = MN ? bmn : bd;
        const long long t0 = clock64();
        for (int i = 0; i < iters; i += 8) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (TS) umma_bf16_ts(tm, tm + 256 + 8 * (u & 3), b, idesc, 1);
                else umma_bf16_ss(tm, ad, b, idesc, 1);
            }
        }
        const long long t1 = clock64();
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        const long long t2 = clock64();
        out[blockIdx.x * 2] = t1 - t0;
        out[blockIdx.x * 2 + 1] = t2 - t0;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}

template <int SHAPE>
static void run_rate(const char* what) {
    long long* d;
    CK(cudaMalloc(&d, 16 * 148));
    CK(cudaFuncSetAttribute(rates<SHAPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    const int iters = 4000;
    for (int rep = 0; rep < 2; ++rep) rates<SHAPE><<<148, 128, 64 * 1024>>>(iters, d);
    CK(cudaDeviceSynchronize());
    long long h[296];
    CK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
    double ci = 0, ce = 0;
    for (int i = 0; i < 148; ++i) { ci += h[2 * i]; ce += h[2 * i + 1]; }
    printf("rate %-34s: issue %.1f execute %.1f cycles per MMA\n", what, ci / 148 / iters, ce / 148 / iters);
    CK(cudaFree(d));
}

int main(int argc, char** argv) {
    if (argc < 2) { printf("usage: probe_attn <test>\n"); return 1; }
    const int test = atoi(argv[1]);
    if (test == 7) {
        run_rate<0>("SS M128 N64");
        run_rate<1>("SS M64 N64 (S)");
        run_rate<2>("SS M128 N128");
        run_rate<11>("SS M128 N128 again");
        run_rate<7>("SS M128 N16");
        run_rate<3>("TS M128 N16");
        run_rate<4>("TS M128 N32");
        run_rate<5>("TS M64 N16");
        run_rate<6>("TS M128 N16 MN-major B");
        run_rate<12>("TS M64 N16 MN-major B (PV)");
        run_rate<13>("TS M64 N24 MN-major B");
        run_rate<8>("TS M128 N96 (proj)");
        run_rate<9>("TS M128 N64");
        run_rate<10>("TS M128 N256");
        return 0;
    }
    std::vector<float> res[2];
    float* dO;
    CK(cudaMalloc(&dO, 128 * 64 * 4));
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16 * PP + 1024));
    for (int pass = 0; pass < 2; ++pass) {
        CK(cudaMemset(dO, 0, 128 * 64 * 4));
        probe<<<1, 128, 16 * PP + 1024>>>(test, pass, dO);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        res[pass].resize(128 * 64);
        CK(cudaMemcpy(res[pass].data(), dO, 128 * 64 * 4, cudaMemcpyDeviceToHost));
    }
    printf("== test %d ==\n", test);
    if (test <= 4) {
        // pass 0: D value = source row id; pass 1: value = k = n % 16.  Print per lane: row id seen at col 0 / col 17 / col 63, k at col 5, 21
        for (int l = 0; l < 128; ++l) {
            printf("lane %3d: row@c0 %6.1f row@c17 %6.1f row@c63 %6.1f | k@c5 %5.1f k@c21 %5.1f k@c63 %5.1f\n", l, res[0][l * 64], res[0][l * 64 + 17],
                   res[0][l * 64 + 63], res[1][l * 64 + 5], res[1][l * 64 + 21], res[1][l * 64 + 63]);
        }
    } else {
        const int K = test == 5 ? 16 : 64;
        int bad = 0;
        for (int r = 0; r < 128; ++r)
            for (int n = 0; n < 16; ++n) {
                const float e = (float)(16 * ((r % K) % 16) + n);
                if (res[0][r * 64 + n] != e) ++bad;
            }
        printf("MN-major B K=%d: %s (bad=%d)\n", K, bad ? "MISMATCH" : "MATCH", bad);
        if (bad)
            for (int r : {0, 1, 2, 9, 17, 70}) {
                printf(" r=%3d:", r);
                for (int n = 0; n < 16; ++n) printf(" %5.0f", res[0][r * 64 + n]);
                printf("\n");
            }
    }
    return 0;
}
