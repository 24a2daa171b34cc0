// MUFU throughput probe (B200): cycles per warp-instruction per SM sub-partition for ex2 / tanh in f32, f16x2, bf16x2.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/probe_mufu tools/probe_mufu.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__device__ __forceinline__ uint32_t op(uint32_t x) {
    uint32_t y;
    if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=r"(y) : "r"(x));
    else if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x));
    else if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
    else if (MODE == 3) asm volatile("tanh.approx.f32 %0, %1;" : "=r"(y) : "r"(x));
    else if (MODE == 4) asm volatile("tanh.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x));
    else asm volatile("tanh.approx.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
    return y;
}

template <int MODE>
__global__ void k(uint32_t* out, long long* cyc, int iters) {
    uint32_t v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = threadIdx.x * 8 + j + 0x3c003c00u;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = op<MODE>(v[j]);  // 8 independent chains per thread
    }
    const long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s ^= v[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, uint32_t* out, long long* cyc) {
    const int iters = 2000;
    for (int warps : {4, 8, 16}) {
        k<MODE><<<1, warps * 32>>>(out, cyc, iters);
        cudaDeviceSynchronize();
        long long c;
        cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        // warp-instructions per SMSP = warps/4 * iters * 8
        printf("%-22s %2d warps: %.2f cycles per warp-instruction per SMSP\n", name, warps, (double)c / (warps / 4.0 * iters * 8));
    }
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
    run<0>("ex2.approx.ftz.f32", out, cyc);
    run<1>("ex2.approx.f16x2", out, cyc);
    run<2>("ex2.approx.ftz.bf16x2", out, cyc);
    run<3>("tanh.approx.f32", out, cyc);
    run<4>("tanh.approx.f16x2", out, cyc);
    run<5>("tanh.approx.bf16x2", out, cyc);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
