// Hardware probe (development tool, not product code): decodes which shared-memory bytes a tcgen05.mma
// reads for a given shared-memory matrix descriptor. B is an identity so D[r][n] == A_row_r[n]; running with
// "pixel id" data and "channel id" data reveals the exact (row, k) -> smem address map.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_umma probe_umma.cu
// Usage: probe_umma <test-id>   (one process per test so a faulting descriptor cannot poison later tests)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda.h>
#include "../superresolution_b200/csrc/ptx.cuh"

using namespace srb;

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e_ = (x);                                                          \
        if (e_ != cudaSuccess) {                                                       \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                   \
        }                                                                              \
    } while (0)

struct Cfg {
    const char* name;
    int fill;        // 0 TMA 2D sw128, 1 manual sw128, 2 manual no-swizzle planes, 3 TMA 4D patch, 4 TMA 2D chan-OOB
    int layout;      // UMMA_SW_*
    int start_rows;  // start shift in rows (128 B for sw128, 16 B for no-swizzle)
    int sbo;         // bytes
    int lbo;         // bytes (no-swizzle: plane pitch; sw128: 16)
    int base_mode;   // 0: base_offset=0, 1: (addr>>7)&7
    int N;
};

static const int R = 256;           // rows of X resident in smem
static const int PLANE = R * 16 + 16;  // no-swizzle plane pitch (padded)

static Cfg cfgs[] = {
    {"tma sw128 s=0 (baseline)", 0, UMMA_SW_128B, 0, 1024, 16, 0, 64},
    {"manual sw128 s=0", 1, UMMA_SW_128B, 0, 1024, 16, 0, 64},
    {"tma sw128 s=1 base0", 0, UMMA_SW_128B, 1, 1024, 16, 0, 64},
    {"tma sw128 s=2 base0", 0, UMMA_SW_128B, 2, 1024, 16, 0, 64},
    {"tma sw128 s=5 base0", 0, UMMA_SW_128B, 5, 1024, 16, 0, 64},
    {"tma sw128 s=8 base0", 0, UMMA_SW_128B, 8, 1024, 16, 0, 64},
    {"tma sw128 s=9 base0", 0, UMMA_SW_128B, 9, 1024, 16, 0, 64},
    {"tma sw128 s=1 baseF", 0, UMMA_SW_128B, 1, 1024, 16, 1, 64},
    {"tma sw128 s=2 baseF", 0, UMMA_SW_128B, 2, 1024, 16, 1, 64},
    {"tma sw128 s=5 baseF", 0, UMMA_SW_128B, 5, 1024, 16, 1, 64},
    {"tma sw128 s=9 baseF", 0, UMMA_SW_128B, 9, 1024, 16, 1, 64},
    {"tma sw128 s=0 sbo=1280", 0, UMMA_SW_128B, 0, 1280, 16, 0, 64},
    {"tma sw128 s=3 sbo=1280 base0", 0, UMMA_SW_128B, 3, 1280, 16, 0, 64},
    {"tma sw128 s=3 sbo=1280 baseF", 0, UMMA_SW_128B, 3, 1280, 16, 1, 64},
    {"nosw planes s=0", 2, UMMA_SW_NONE, 0, 128, PLANE, 0, 64},
    {"nosw planes s=1", 2, UMMA_SW_NONE, 1, 128, PLANE, 0, 64},
    {"nosw planes s=13", 2, UMMA_SW_NONE, 13, 128, PLANE, 0, 64},
    {"nosw planes s=3 sbo=160", 2, UMMA_SW_NONE, 3, 160, PLANE, 0, 64},
    {"tma4d patch (halo, OOB zero)", 3, UMMA_SW_128B, 0, 1024, 16, 0, 64},
    {"tma2d channel OOB (C=32 of 64)", 4, UMMA_SW_128B, 0, 1024, 16, 0, 64},
    {"tma sw128 N=32", 0, UMMA_SW_128B, 0, 1024, 16, 0, 32},
    {"tma sw128 N=96", 0, UMMA_SW_128B, 0, 1024, 16, 0, 96},
    {"tma sw128 N=256", 0, UMMA_SW_128B, 0, 1024, 16, 0, 256},
    {"tma sw128 s=0 sbo=2048", 0, UMMA_SW_128B, 0, 2048, 16, 0, 64},
};
static const int NCFG = sizeof(cfgs) / sizeof(cfgs[0]);

struct KParams {
    int fill, layout, start_rows, sbo, lbo, base_mode, N;
};

// smem: A region 40 KB (R rows * 128 B + slack), B region 32 KB (256 rows * 128 B)
static const int A_BYTES = 40 * 1024;
static const int B_BYTES = 33 * 1024;

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __nv_bfloat16* __restrict__ X, const __nv_bfloat16* __restrict__ Wt, float* __restrict__ out,
             KParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB = smem + A_BYTES;
    __shared__ uint64_t bar_full, bar_mma;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid / 32;
    const int N = p.N;

    if (tid == 0) {
        mbar_init(&bar_full, 1);
        mbar_init(&bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(&tmem_base_s, 256);
        tmem_relinquish();
    }
    // zero the A region so unfilled rows read as 0
    for (int i = tid; i < A_BYTES / 16; i += 128) ((uint4*)sA)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;

    // ---- manual fills first (generic proxy), then TMA fills (async proxy) ----
    uint32_t tx = 0;
    if (p.fill == 1) {
        for (int i = tid; i < R * 8; i += 128) {
            int row = i / 8, ch = i % 8;
            uint4 v = ((const uint4*)X)[row * 8 + ch];
            *(uint4*)(sA + row * 128 + ((ch ^ (row & 7)) * 16)) = v;
        }
    } else if (p.fill == 2) {
        for (int i = tid; i < R * 8; i += 128) {
            int row = i / 8, kg = i % 8;
            uint4 v = ((const uint4*)X)[row * 8 + kg];
            *(uint4*)(sA + kg * PLANE + row * 16) = v;
        }
    }
    if (p.layout == UMMA_SW_NONE) {
        // B planes: kg * (N*16) + n*16
        for (int i = tid; i < N * 8; i += 128) {
            int n = i / 8, kg = i % 8;
            uint4 v = ((const uint4*)Wt)[n * 8 + kg];
            *(uint4*)(sB + kg * (N * 16) + n * 16) = v;
        }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
        if (p.fill == 0 || p.fill == 4) tx += R * 128;
        if (p.fill == 3) tx += 100 * 128;
        if (p.layout != UMMA_SW_NONE) tx += N * 128;
        mbar_arrive_expect_tx(&bar_full, tx);
        if (p.fill == 0 || p.fill == 4) {
            tma_load_2d(sA, &tmA, &bar_full, 0, 0);                // rows 0..127
            tma_load_2d(sA + 128 * 128, &tmA, &bar_full, 0, 128);  // rows 128..255
        } else if (p.fill == 3) {
            tma_load_4d(sA, &tmA, &bar_full, 0, -1, -1, 1);  // 10x10 patch of image b=1 with halo
        }
        if (p.layout != UMMA_SW_NONE) tma_load_2d(sB, &tmB, &bar_full, 0, 0);
    }

    // ---- MMA ----
    if (warp == 0) {
        mbar_wait(&bar_full, 0);
        tc_fence_after();
        if (elect_one()) {
            const uint32_t idesc = umma_idesc_bf16_f32(128, N);
            const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
            for (int ks = 0; ks < 4; ++ks) {
                uint32_t aaddr, baddr;
                uint64_t ad, bd;
                if (p.layout == UMMA_SW_NONE) {
                    aaddr = a0 + p.start_rows * 16 + ks * 2 * p.lbo;
                    baddr = b0 + ks * 2 * (N * 16);
                    ad = umma_smem_desc(aaddr, p.lbo, p.sbo, UMMA_SW_NONE, 0);
                    bd = umma_smem_desc(baddr, N * 16, 128, UMMA_SW_NONE, 0);
                } else {
                    aaddr = a0 + p.start_rows * 128 + ks * 32;
                    baddr = b0 + ks * 32;
                    uint32_t bo = p.base_mode ? ((aaddr >> 7) & 7) : 0;
                    ad = umma_smem_desc(aaddr, p.lbo, p.sbo, p.layout, bo);
                    bd = umma_smem_desc(baddr, 16, 1024, p.layout, 0);
                }
                umma_bf16_ss(tmem_base, ad, bd, idesc, ks > 0);
            }
            umma_commit(&bar_mma);
        }
        __syncwarp();
    }
    // ---- epilogue: all 4 warps ----
    mbar_wait(&bar_mma, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, r);
        tmem_ld_wait();
        int row = warp * 32 + (tid & 31);
        for (int j = 0; j < 32; ++j) out[row * N + c0 + j] = __uint_as_float(r[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 256);
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn get_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn || qres != cudaDriverEntryPointSuccess) {
        printf("cuTensorMapEncodeTiled not found\n");
        exit(2);
    }
    return (EncodeFn)fn;
}

static CUtensorMap make_map(EncodeFn enc, void* ptr, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                            const cuuint32_t* box) {
    CUtensorMap m;
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, ptr, dims, strides_b, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
        exit(2);
    }
    return m;
}

static float bf(float x) { return x; }

int main(int argc, char** argv) {
    if (argc < 2) {
        printf("%d\n", NCFG);
        return 0;
    }
    int id = atoi(argv[1]);
    if (id < 0 || id >= NCFG) return 1;
    Cfg c = cfgs[id];
    EncodeFn enc = get_encode();
    const int N = c.N;

    // X variants: pass 0 = pixel id (+1), pass 1 = channel id (+1)
    std::vector<float> res[2];
    for (int pass = 0; pass < 2; ++pass) {
        std::vector<__nv_bfloat16> hX(R * 64), hW(256 * 64);
        for (int r = 0; r < R; ++r)
            for (int k = 0; k < 64; ++k) hX[r * 64 + k] = __float2bfloat16(pass == 0 ? (float)(r % 250 + 1) : (float)(k + 1));
        for (int n = 0; n < 256; ++n)
            for (int k = 0; k < 64; ++k) hW[n * 64 + k] = __float2bfloat16((n % 64) == k ? 1.f : 0.f);
        __nv_bfloat16 *dX, *dW;
        float* dO;
        CK(cudaMalloc(&dX, hX.size() * 2));
        CK(cudaMalloc(&dW, hW.size() * 2));
        CK(cudaMalloc(&dO, 128 * 256 * 4));
        CK(cudaMemcpy(dX, hX.data(), hX.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dW, hW.data(), hW.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemset(dO, 0xFF, 128 * 256 * 4));

        CUtensorMap tmA, tmB;
        if (c.fill == 3) {
            // X viewed as NHWC [B=4][H=8][W=8][C=64] (256 pixels)
            cuuint64_t dims[4] = {64, 8, 8, 4};
            cuuint64_t str[3] = {128, 8 * 128, 64 * 128};
            cuuint32_t box[4] = {64, 10, 10, 1};
            tmA = make_map(enc, dX, 4, dims, str, box);
        } else if (c.fill == 4) {
            cuuint64_t dims[2] = {32, (cuuint64_t)R};
            cuuint64_t str[1] = {128};
            cuuint32_t box[2] = {64, 128};
            tmA = make_map(enc, dX, 2, dims, str, box);
        } else {
            cuuint64_t dims[2] = {64, (cuuint64_t)R};
            cuuint64_t str[1] = {128};
            cuuint32_t box[2] = {64, 128};
            tmA = make_map(enc, dX, 2, dims, str, box);
        }
        {
            cuuint64_t dims[2] = {64, 256};
            cuuint64_t str[1] = {128};
            cuuint32_t box[2] = {64, (cuuint32_t)N};
            tmB = make_map(enc, dW, 2, dims, str, box);
        }
        KParams kp = {c.fill, c.layout, c.start_rows, c.sbo, c.lbo, c.base_mode, N};
        size_t smem = A_BYTES + B_BYTES + 1024;
        CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        probe_kernel<<<1, 128, smem>>>(tmA, tmB, dX, dW, dO, kp);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        res[pass].resize(128 * N);
        CK(cudaMemcpy(res[pass].data(), dO, 128 * N * 4, cudaMemcpyDeviceToHost));
    }
    // decode + check expectation
    int G = (c.layout == UMMA_SW_NONE) ? c.sbo / 16 : c.sbo / 128;  // rows per 8-row group pitch
    int bad = 0, first_bad = -1;
    for (int r = 0; r < 128; ++r) {
        for (int n = 0; n < N; ++n) {
            int src = c.start_rows + (r / 8) * G + (r % 8);
            float epix, ech = (float)((n % 64) + 1);
            if (c.fill == 3) {
                int py = src / 10, px = src % 10;
                int y = py - 1, x = px - 1;
                bool in = (src < 100 && y >= 0 && y < 8 && x >= 0 && x < 8);
                epix = in ? (float)((64 + y * 8 + x) % 250 + 1) : 0.f;
                if (!in) ech = 0.f;
            } else {
                epix = (src < R) ? (float)(src % 250 + 1) : 0.f;
                if (src >= R) ech = 0.f;
                if (c.fill == 4 && (n % 64) >= 32) { epix = 0.f; ech = 0.f; }
            }
            float gp = res[0][r * N + n], gc = res[1][r * N + n];
            if (gp != bf(epix) || gc != bf(ech)) {
                if (first_bad < 0) first_bad = r * N + n;
                ++bad;
            }
        }
    }
    printf("[%2d] %-34s : %s (bad=%d)\n", id, c.name, bad ? "MISMATCH" : "MATCH", bad);
    if (bad) {
        printf("     first bad at r=%d n=%d\n", first_bad / N, first_bad % N);
        int rows[] = {0, 1, 2, 7, 8, 9, 15, 16, 127};
        for (int r : rows) {
            printf("     r=%3d pix:", r);
            for (int n = 0; n < 64 && n < N; n += 8) printf(" %5.0f", res[0][r * N + n]);
            printf(" | ch:");
            for (int n = 0; n < 64 && n < N; n += 8) printf(" %3.0f", res[1][r * N + n]);
            printf("\n");
        }
    }
    return 0;
}
