// Microbenchmark: tcgen05.ld throughput (TMEM -> registers) per SM for 4 / 8 / 16 warps and x8 / x16 / x32 shapes.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I superresolution_b200/csrc -o tools/bin/bench_tmem tools/bench_tmem.cu
#include <cstdio>
#include "ptx.cuh"
using namespace srb;

template <int X>
__global__ void k(int iters, long long* out, uint32_t* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        const uint32_t col = ((i * 32) + (warp >> 2) * 64) & 255;
        if (X == 32) { uint32_t r[32]; tmem_ld_32x32b_x32(base + col, r); tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) acc ^= r[j]; }
        if (X == 16) { uint32_t r[16]; tmem_ld_32x32b_x16(base + col, r); tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) acc ^= r[j]; }
        if (X == 8) { uint32_t r[8]; tmem_ld_32x32b_x8(base + col, r); tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 8; ++j) acc ^= r[j]; }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (acc == 0x12345678) sink[0] = acc;
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(slot, 512);
}

template <int X>
void run(int warps) {
    long long* d; uint32_t* s; cudaMalloc(&d, 8 * 148); cudaMalloc(&s, 4);
    const int iters = 2000;
    k<X><<<148, warps * 32>>>(iters, d, s);
    k<X><<<148, warps * 32>>>(iters, d, s);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = 0; for (int i = 0; i < 148; ++i) cyc += h[i]; cyc /= 148;
    const double bytes = (double)iters * warps * 32 * X * 4;
    printf("x%-2d warps %2d: %.0f cycles, %.1f B/clk/SM, %.1f cycles per warp-load (%s)\n", X, warps, cyc, bytes / cyc,
           cyc / iters, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d); cudaFree(s);
}
int main() {
    for (int w : {4, 8, 16}) { run<32>(w); run<16>(w); run<8>(w); }
    return 0;
}
