// Hardware probe (development tool): is the ~45-cycle floor per small tcgen05.mma a per-issuing-thread limit?
// NW warps each issue `iters` MMAs into their own TMEM column range; reports cycles per MMA (aggregate).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I superresolution_b200/csrc -o tools/bin/probe_issue tools/probe_issue.cu
#include <cstdio>
#include "ptx.cuh"
using namespace srb;
constexpr int PP = 128 * 16 + 16;
template <int N, bool TS, int SWZ, int M = 128>
__global__ void __launch_bounds__(128, 1) k(int nw, int iters, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    fence_proxy_async_smem();
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tm = slot;
    long long t0 = 0, t2 = 0;
    if (warp < nw && lane == 0) {
        const uint32_t a_u = smem_u32(smem) + warp * 8 * 1024, b_u = smem_u32(smem + 48 * 1024) + warp * 8 * 1024;
        const uint64_t ad = SWZ ? umma_smem_desc(a_u, 16, 1024, UMMA_SW_128B) : umma_smem_desc(a_u, PP, 128, UMMA_SW_NONE);
        const uint64_t bd = SWZ ? umma_smem_desc(b_u, 16, 1024, UMMA_SW_128B) : umma_smem_desc(b_u, PP, 128, UMMA_SW_NONE);
        const uint32_t idesc = umma_idesc_bf16_f32(M, N);
        const uint32_t d = tm + warp * 64;
        t0 = clock64();
        for (int i = 0; i < iters; i += 8) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (TS) umma_bf16_ts(d, tm + 256 + 8 * (u & 3), bd, idesc, 1);
                else umma_bf16_ss(d, ad, bd, idesc, 1);
            }
        }
        umma_commit(&bar[warp]);
        mbar_wait(&bar[warp], 0);
        t2 = clock64();
        out[blockIdx.x * 4 + warp] = t2 - t0;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}
template <int N, bool TS, int SWZ, int M = 128>
void run(const char* what) {
    long long* d; cudaMalloc(&d, 32 * 148);
    cudaFuncSetAttribute(k<N, TS, SWZ, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    const int iters = 4000;
    for (int nw = 1; nw <= 4; ++nw) {
        cudaMemset(d, 0, 32 * 148);
        for (int rep = 0; rep < 2; ++rep) k<N, TS, SWZ, M><<<148, 128, 96 * 1024>>>(nw, iters, d);
        cudaDeviceSynchronize();
        long long h[148 * 4]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        double c = 0; for (int i = 0; i < 148; ++i) { long long m = 0; for (int w = 0; w < nw; ++w) m = h[4 * i + w] > m ? h[4 * i + w] : m; c += m; }
        c /= 148;
        printf("%-28s issuers %d: %.1f cycles per MMA (aggregate), %s\n", what, nw, c / (iters * nw), cudaGetErrorString(cudaGetLastError()));
    }
    cudaFree(d);
}
int main() {
    run<16, true, 0>("TS M128 N16");
    run<32, false, 0>("SS M128 N32 planes");
    run<32, false, 1>("SS M128 N32 sw128");
    run<64, false, 1>("SS M128 N64 sw128");
    run<96, false, 0>("SS M128 N96 planes");
    run<64, true, 0>("TS M128 N64 (bias)");
    run<32, true, 0>("TS M128 N32 (PV)");
    run<64, false, 0, 64>("SS M64 N64 planes (QK)");
    run<96, true, 0>("TS M128 N96");
    return 0;
}
