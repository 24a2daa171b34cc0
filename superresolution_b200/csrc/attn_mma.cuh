// Warp-level building blocks shared by the training-path attention kernels (win_attn.cu, ocab_attn_train.cu):
// mma.sync.m16n8k16 bf16, ldmatrix fragment loaders for padded row-major shared-memory tiles, movmatrix transpose,
// cp.async staging, quad reductions.
#pragma once
#include <cuda_bf16.h>

#include "common.h"

namespace srb {
namespace amma {

constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *(uint32_t*)&h;
}
__device__ __forceinline__ void cp16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_commit_wait() {
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
// 8x8 b16 transpose across the warp: accumulator-layout block -> A-fragment block of the transposed matrix
__device__ __forceinline__ uint32_t movt(uint32_t a) {
    uint32_t d;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
    return d;
}
__device__ __forceinline__ float ex2f(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float quad_max(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// A fragment: rows row0..row0+15, 16 k-columns starting at element col0 of a [rows][pitch] bf16 tile
__device__ __forceinline__ void frag_a(uint32_t (&r)[4], const bf16* tile, int pitch, int row0, int col0, int lane) {
    ldsm4(r, tile + (row0 + (lane & 7) + ((lane >> 3) & 1) * 8) * pitch + col0 + (lane >> 4) * 8);
}
// B fragments of two adjacent n-tiles (rows n0..n0+15 of the tile are the n index, k contiguous): {b0,b1 | b0,b1}
__device__ __forceinline__ void frag_b_nk(uint32_t (&r)[4], const bf16* tile, int pitch, int n0, int col0, int lane) {
    ldsm4(r, tile + (n0 + (lane & 7) + (lane >> 4) * 8) * pitch + col0 + ((lane >> 3) & 1) * 8);
}
// B fragments for k = rows k0..k0+15 of the tile, n = 16 columns from col0 (two n-tiles): {b0,b1 of n 0-7 | b0,b1 of n 8-15}
__device__ __forceinline__ void frag_b_kn(uint32_t (&r)[4], const bf16* tile, int pitch, int k0, int col0, int lane) {
    ldsm4t(r, tile + (k0 + (lane & 7) + ((lane >> 3) & 1) * 8) * pitch + col0 + (lane >> 4) * 8);
}

}  // namespace amma
}  // namespace srb
