// Microbenchmark: tcgen05.mma issue/execute rate per SM for the shapes the conv / HAB / MLP kernels use
// (kind::f16, bf16 x bf16 -> fp32, M = 128, K = 16 per instruction, SS operands), one CTA per SM, one issuing thread.
//   mode 0: canonical SWIZZLE_128B K-major tiles (SBO 1024)            - the GEMM layout
//   mode 1: conv-style A: SWIZZLE_128B rows with SBO = 34*128 (patch row pitch), shifted start per "tap"
//   mode 2: un-swizzled K-major planes [k/8][row][8] (LBO = plane pitch, SBO = 128)   - hab_tc / mlp_tc layout
// Optional: `ld` epilogue warps hammer tcgen05.ld on a different TMEM column range while the MMAs run.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I superresolution_b200/csrc -o tools/bin/bench_umma tools/bench_umma.cu
#include <cstdio>
#include "ptx.cuh"
using namespace srb;

__global__ void __launch_bounds__(288, 1) k(int N, int mode, int iters, int ldwarps, long long* out, uint32_t* sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x3c003c00u + i;  // finite bf16
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tm = slot;
    const uint32_t a_u32 = smem_u32(smem), b_u32 = smem_u32(smem + 128 * 1024);
    __shared__ volatile int stop;
    if (threadIdx.x == 0) stop = 0;
    __syncthreads();
    if (warp == 0) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16_f32(128, N);
            uint64_t ad[4], bd[4];
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                if (mode == 0) {
                    ad[ks] = umma_smem_desc(a_u32 + ks * 32, 16, 1024, UMMA_SW_128B);
                    bd[ks] = umma_smem_desc(b_u32 + ks * 32, 16, 1024, UMMA_SW_128B);
                } else if (mode == 1) {
                    ad[ks] = umma_smem_desc(a_u32 + (34 + 1) * 128 + ks * 32, 16, 34 * 128, UMMA_SW_128B);
                    bd[ks] = umma_smem_desc(b_u32 + ks * 32, 16, 1024, UMMA_SW_128B);
                } else {
                    ad[ks] = umma_smem_desc(a_u32 + ks * 2 * 2064, 2064, 128, UMMA_SW_NONE);
                    bd[ks] = umma_smem_desc(b_u32 + ks * 2 * (N * 16), N * 16, 128, UMMA_SW_NONE);
                }
            }
            const long long t0 = clock64();
            for (int i = 0; i < iters; i += 4) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) umma_bf16_ss(tm, ad[ks], bd[ks], idesc, 1u);
            }
            const long long t1 = clock64();
            umma_commit(&bar);
            mbar_wait(&bar, 0);
            const long long t2 = clock64();
            out[blockIdx.x * 2] = t1 - t0;
            out[blockIdx.x * 2 + 1] = t2 - t0;
            stop = 1;
        }
        __syncwarp();
    } else if (warp - 1 < ldwarps) {
        const uint32_t base = tm + ((uint32_t)(((warp - 1) & 3) * 32) << 16) + 256;
        uint32_t acc = 0;
        while (!stop) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(base + (((warp - 1) >> 2) * 32), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) acc ^= r[j];
        }
        if (acc == 0x12345678) sink[0] = acc;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(slot, 512);
}

int main() {
    long long* d; uint32_t* s;
    cudaMalloc(&d, 16 * 148); cudaMalloc(&s, 4);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 4000;
    const char* names[3] = {"sw128 canonical", "sw128 conv patch (SBO 4352, tap shifts)", "no-swizzle planes"};
    for (int ld : {0, 8})
        for (int mode = 0; mode < 3; ++mode)
            for (int N : {32, 64, 96, 128, 192, 256}) {
                for (int rep = 0; rep < 2; ++rep) k<<<148, 288, 200 * 1024>>>(N, mode, iters, ld, d, s);
                cudaDeviceSynchronize();
                long long h[296]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
                double ci = 0, ce = 0; for (int i = 0; i < 148; ++i) { ci += h[2 * i]; ce += h[2 * i + 1]; }
                ci /= 148; ce /= 148;
                printf("%-40s ld-warps %d N %3d: issue %.1f, execute %.1f cycles per MMA (math floor %.0f, smem model %.0f) %s\n",
                       names[mode], ld, N, ci / iters, ce / iters, N / 2.0, (128 * 32 + N * 32) / 128.0,
                       cudaGetErrorString(cudaGetLastError()));
            }
    return 0;
}
