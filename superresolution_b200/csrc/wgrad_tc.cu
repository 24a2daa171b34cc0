// Weight gradients of the conv / linear layers on tcgen05 (reference: autograd of nn.Conv2d /
// nn.Linear inside BasicSR RRDBNet and HAT, reached from scripts/Modello_supporto.py:133).
//
//   dW[co][ci][tap] = sum_px dY[px][co] * X[px + tap][ci],   db[co] = sum_px dY[px][co]
//
// is a GEMM whose contraction runs over PIXELS, with both operands channel-contiguous in memory ([px][c], NHWC).  That
// is exactly an "MN-major" UMMA operand: a TMA box of 64 channels x R pixels with SWIZZLE_128B lands as R rows of 128
// bytes, 8 rows = one 1 KB swizzle atom, and the MMA walks it with LBO = distance between 64-channel atoms and
// SBO = 1 KB between 8-pixel groups.  No transposition pass, no ldmatrix.trans: the tensors are consumed as stored.
// The nine taps of a 3x3 layer are nine start-address shifts into one haloed X patch (the swizzle is a function of
// the shared-memory address, so a shifted view reads what TMA wrote).
//
// One CTA owns a split of the pixels for the whole launch and accumulates in TMEM (fp32); three operand roles:
//   mode L   (linear, 1 tap)   M = dY (Cout in tiles of 128 rows), N = X (Cin <= 256);  D tile t at columns t * Cin
//   mode C1  (3x3, Cout <= 64) M = X chunk of 128 channels, N = 32 channels of dY, dY shifted:
//                              dW[co][ci][tap] = sum_q X[q][ci] dY[q - tap][co]
//   mode C2  (3x3, Cout > 64)  M = dY chunk of 128 channels, N = 32 channels of X, X shifted
// In the 3x3 modes the shifted (narrow) operand is a haloed patch of 32-channel rows (64 bytes, SWIZZLE_64B), and ONE MMA
// covers the three horizontal taps of a kernel row: its N = 96 is three 32-channel atoms whose leading-dimension
// offset is 64 bytes = one pixel, i.e. three overlapping views of the patch one pixel apart.  Three N = 96 MMAs per
// k-step instead of nine N = 32 ones (which are issue-bound at ~45 cycles each).  D block b = 3 * row + atom at
// columns 32 b holds tap b (C2) or tap 8 - b (C1: the shift is on dY, so the view order is mirrored).
// The bias gradient is one more MMA per k-step against an all-ones operand (any layout: every element is 1.0) into 16
// or 32 spare TMEM columns.  Split-K partials go to the same [split][tap][co][ci] / [split][co] scratch layout as the
// mma.sync kernels of conv_bwd.cu, so wgrad_reduce / bias_grad_final (fixed order, deterministic) finish both.
//
// Warps: 0 = TMA producer, 1 = MMA issuer (+ TMEM allocation), 2-5 = epilogue (TMEM lane quadrant = warp & 3), 6 = second
// MMA issuer of the 3x3 modes (their N = 32 MMAs are issue-bound at ~45 cycles per instruction and thread: the two
// issuers take alternate taps).
#include <cuda.h>

#include <cstring>

#include "common.h"
#include "ops.h"
#include "ptx.cuh"
#include "tmap.h"

namespace srb {

namespace {

constexpr int WT_THREADS = 224;  // 7 warps
constexpr int WT_TW = 32, WT_TH = 4;                 // 3x3 chunk: 4 image rows x 32 pixels (8 k-steps of 16 pixels)
constexpr int WT_PW = WT_TW + 2, WT_PH = WT_TH + 2;  // haloed X patch of the chunk
constexpr int WT_KP = 64;                            // linear chunk: 64 consecutive pixels (4 k-steps)
constexpr int WT_FLAT_ATOM = WT_KP * 128;            // bytes of one 64-channel atom (128-byte rows): flat chunk,
constexpr int WT_TILE_ATOM = WT_TW * WT_TH * 128;    //   3x3 chunk of the un-shifted operand
constexpr int WT_NPATCH_BYTES = WT_PW * WT_PH * 64;  // haloed patch of the shifted operand: 32-channel rows of 64 bytes
constexpr int WT_NPATCH_ATOM = (WT_NPATCH_BYTES + 1023) / 1024 * 1024;
constexpr int WT_ONES_BYTES = 4096;
constexpr int WT_MAX_STAGES = 6;
constexpr int WT_SMEM_LIMIT = 227 * 1024;
enum { WT_L = 0, WT_C1 = 1, WT_C2 = 2 };

struct WtArgs {
    float* partial;   // [ksplit][taps][Cout][Cin]
    float* bpartial;  // [ksplit][Cout] or nullptr
    int Cin, Cout, mode;
    int nchunks, chunks_per_split;
    int tiles_x, tiles_y;          // 3x3 modes: chunks per image row / column
    int n_n_chunks;                // blockIdx.y = m_chunk * n_n_chunks + n_chunk
    int m_tiles;                   // mode L: 128-row tiles of dY
    int m_atoms, m_atom_bytes;     // atoms carved per stage on the M side (2 per tile) and their pitch
    int n_atoms, n_atom_bytes;
    int stage_bytes, nstages;
};

template <int MODE>
__global__ void __launch_bounds__(WT_THREADS, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmM, const __grid_constant__ CUtensorMap tmN, const WtArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* ones = smem + a.nstages * a.stage_bytes;
    uint64_t* full = (uint64_t*)(ones + WT_ONES_BYTES);
    uint64_t* empty = full + WT_MAX_STAGES;
    uint64_t* acc_full = empty + WT_MAX_STAGES;
    uint32_t* tmem_slot = (uint32_t*)(acc_full + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    const int m_chunk = blockIdx.y / a.n_n_chunks, n_chunk = blockIdx.y % a.n_n_chunks;
    const int Cm = MODE == WT_C1 ? a.Cin : a.Cout;           // channels of the M-side tensor
    const int m_c0 = MODE == WT_L ? 0 : 128 * m_chunk;       // first channel of this CTA's M rows
    const int m_loaded = min(a.m_atoms, (Cm - m_c0 + 63) / 64);  // atoms that hold channels (the rest stays zero)
    const int c_begin = blockIdx.x * a.chunks_per_split;
    const int c_end = min(a.nchunks, c_begin + a.chunks_per_split);

    if (threadIdx.x == 0) {
        const uint32_t n_iss = MODE == WT_L ? 1 : 2;  // MMA issuer warps that release a stage / publish the accumulators
        for (int i = 0; i < WT_MAX_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], n_iss); }
        mbar_init(acc_full, n_iss);
        fence_barrier_init();
        tma_prefetch_desc(&tmM);
        tma_prefetch_desc(&tmN);
    }
    // atoms beyond the tensor's channels are never written by TMA and must multiply as zeros; the ones tile is 1.0
    for (int i = threadIdx.x; i < a.nstages * a.stage_bytes / 16; i += WT_THREADS) ((uint4*)smem)[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < WT_ONES_BYTES / 4; i += WT_THREADS) ((uint32_t*)ones)[i] = 0x3F803F80u;
    fence_proxy_async_smem();
    if (warp == 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        const uint32_t tx_bytes = MODE == WT_L ? (uint32_t)(m_loaded + a.n_atoms) * WT_FLAT_ATOM
                                               : (uint32_t)m_loaded * WT_TILE_ATOM + WT_NPATCH_BYTES;
        uint32_t it = 0;
        for (int c = c_begin; c < c_end; ++c, ++it) {
            const uint32_t s = it % a.nstages;
            mbar_wait(&empty[s], ((it / a.nstages) & 1) ^ 1);
            if (elect_one()) {
                uint8_t* sm_m = smem + s * a.stage_bytes;
                uint8_t* sm_n = sm_m + a.m_atoms * a.m_atom_bytes;
                mbar_arrive_expect_tx(&full[s], tx_bytes);
                if (MODE == WT_L) {
                    for (int k = 0; k < m_loaded; ++k) tma_load_2d(sm_m + k * a.m_atom_bytes, &tmM, &full[s], 64 * k, c * WT_KP);
                    for (int k = 0; k < a.n_atoms; ++k) tma_load_2d(sm_n + k * a.n_atom_bytes, &tmN, &full[s], 64 * k, c * WT_KP);
                } else {
                    const int per_img = a.tiles_x * a.tiles_y;
                    const int b = c / per_img, r = c % per_img;
                    const int y0 = (r / a.tiles_x) * WT_TH, x0 = (r % a.tiles_x) * WT_TW;
                    for (int k = 0; k < m_loaded; ++k)
                        tma_load_4d(sm_m + k * a.m_atom_bytes, &tmM, &full[s], m_c0 + 64 * k, x0, y0, b);
                    tma_load_4d(sm_n, &tmN, &full[s], 32 * n_chunk, x0 - 1, y0 - 1, b);  // haloed: origin (-1, -1)
                }
            }
            __syncwarp();
        }
    } else if (warp == 1 || (warp == 6 && MODE != WT_L)) {
        // ===================== MMA issuer(s) =====================
        const int iss = warp == 1 ? 0 : 1;
        const uint32_t ones_u = smem_u32(ones);
        const uint64_t ones_desc = umma_smem_desc(ones_u, 2048, 128, UMMA_SW_NONE);  // K-major planes, all 1.0
        uint32_t it = 0;
        for (int c = c_begin; c < c_end; ++c, ++it) {
            const uint32_t s = it % a.nstages;
            mbar_wait(&full[s], (it / a.nstages) & 1);
            tc_fence_after();
            const uint32_t m_u = smem_u32(smem + s * a.stage_bytes);
            const uint32_t n_u = m_u + a.m_atoms * a.m_atom_bytes;
            const bool leader = elect_one();
            if (MODE == WT_L) {
                const uint32_t idesc = umma_idesc_bf16_f32_major(128, a.Cin, 1, 1);
                const uint32_t idesc_b = umma_idesc_bf16_f32_major(128, 16, 1, 0);
#pragma unroll
                for (int ks = 0; ks < WT_KP / 16; ++ks) {
                    const uint32_t acc = (c > c_begin || ks > 0);
                    const uint64_t bd = umma_smem_desc(n_u + ks * 2048, a.n_atom_bytes, 1024, UMMA_SW_128B);
                    for (int t = 0; t < a.m_tiles; ++t) {
                        const uint64_t ad = umma_smem_desc(m_u + 2 * t * a.m_atom_bytes + ks * 2048, a.m_atom_bytes, 1024, UMMA_SW_128B);
                        if (leader) {
                            umma_bf16_ss(tmem_base + t * a.Cin, ad, bd, idesc, acc);
                            if (a.bpartial) umma_bf16_ss(tmem_base + a.m_tiles * a.Cin + 16 * t, ad, ones_desc, idesc_b, acc);
                        }
                    }
                }
            } else {
                const uint32_t idesc = umma_idesc_bf16_f32_major(128, 96, 1, 1);
                // bias: C1 = ones (A, K-major) x the un-shifted view of dY (B);  C2 = dY (A) x ones (B, K-major); one CTA
                // column of the grid adds it
                const bool bias = a.bpartial && (MODE == WT_C1 ? m_chunk == 0 : n_chunk == 0);
                const uint32_t idesc_b = MODE == WT_C1 ? umma_idesc_bf16_f32_major(128, 32, 0, 1)
                                                       : umma_idesc_bf16_f32_major(128, 16, 1, 0);
#pragma unroll 1
                for (int ks = 0; ks < WT_TW * WT_TH / 16; ++ks) {
                    const int yy = ks >> 1, xs = (ks & 1) * 16;
                    const uint32_t acc = (c > c_begin || ks > 0);
                    const uint64_t md = umma_smem_desc(m_u + (yy * WT_TW + xs) * 128, a.m_atom_bytes, 1024, UMMA_SW_128B);
                    // patch rows yy + j (j = 0..2), pixels xs + atom (atom = 0..2 through LBO = 64 bytes)
                    const uint64_t nd0 = umma_smem_desc(n_u + (yy * WT_PW + xs) * 64, 64, 512, UMMA_SW_64B);
#pragma unroll
                    for (int j = 0; j < 3; ++j)
                        if (leader && (j >> 1) == iss)  // issuer 0: kernel rows 0, 1; issuer 1: row 2 + the bias
                            umma_bf16_ss(tmem_base + 96 * j, md, nd0 + (uint64_t)((j * WT_PW * 64) >> 4), idesc, acc);
                    if (bias && leader && iss == 1) {
                        if (MODE == WT_C1) umma_bf16_ss(tmem_base + 288, ones_desc, nd0 + (uint64_t)(((WT_PW + 1) * 64) >> 4), idesc_b, acc);
                        else umma_bf16_ss(tmem_base + 288, md, ones_desc, idesc_b, acc);
                    }
                }
            }
            if (leader) umma_commit(&empty[s]);
            __syncwarp();
        }
        if (elect_one()) umma_commit(acc_full);
        __syncwarp();
    } else if (warp >= 2 && warp < 6) {
        // ===================== epilogue: TMEM -> split-K partials =====================
        mbar_wait(acc_full, 0);
        tc_fence_after();
        const int row = (warp & 3) * 32 + lane;
        const uint32_t tm_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const bool any = c_end > c_begin;  // a split without chunks writes zeros (its TMEM was never written)
        uint32_t v[32];
        if (MODE == WT_L) {
            float* dst = a.partial + (size_t)blockIdx.x * a.Cout * a.Cin;
            for (int t = 0; t < a.m_tiles; ++t) {
                const int co = 128 * t + row;
                for (int c0 = 0; c0 < a.Cin; c0 += 32) {
                    tmem_ld_32x32b_x32(tm_lane + t * a.Cin + c0, v);
                    tmem_ld_wait();
                    if (co < a.Cout) {
                        float4* o = (float4*)(dst + (size_t)co * a.Cin + c0);
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            o[j] = any ? make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                                     __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]))
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
                if (a.bpartial) {
                    uint32_t b8[8];
                    tmem_ld_32x32b_x8(tm_lane + a.m_tiles * a.Cin + 16 * t, b8);
                    tmem_ld_wait();
                    if (co < a.Cout) a.bpartial[(size_t)blockIdx.x * a.Cout + co] = any ? __uint_as_float(b8[0]) : 0.f;
                }
            }
        } else {
            float* dst = a.partial + (size_t)blockIdx.x * 9 * a.Cout * a.Cin;
            const int mc = m_c0 + row;  // C1: input channel, C2: output channel
            for (int blk = 0; blk < 9; ++blk) {
                const int tap = MODE == WT_C1 ? 8 - blk : blk;
                tmem_ld_32x32b_x32(tm_lane + 32 * blk, v);
                tmem_ld_wait();
                if (MODE == WT_C1) {
                    if (mc < a.Cin) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            dst[((size_t)tap * a.Cout + 32 * n_chunk + j) * a.Cin + mc] = any ? __uint_as_float(v[j]) : 0.f;
                    }
                } else if (mc < a.Cout) {
                    float4* o = (float4*)(dst + ((size_t)tap * a.Cout + mc) * a.Cin + 32 * n_chunk);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        o[j] = any ? make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                                 __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]))
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            if (a.bpartial) {
                if (MODE == WT_C1 && m_chunk == 0) {  // every row of the 32 bias columns holds the sums: row 0's warp writes
                    tmem_ld_32x32b_x32(tm_lane + 288, v);
                    tmem_ld_wait();
                    if ((warp & 3) == 0) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (lane == j) a.bpartial[(size_t)blockIdx.x * a.Cout + 32 * n_chunk + j] = any ? __uint_as_float(v[j]) : 0.f;
                    }
                } else if (MODE == WT_C2 && n_chunk == 0) {
                    uint32_t b8[8];
                    tmem_ld_32x32b_x8(tm_lane + 288, b8);
                    tmem_ld_wait();
                    if (mc < a.Cout) a.bpartial[(size_t)blockIdx.x * a.Cout + mc] = any ? __uint_as_float(b8[0]) : 0.f;
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
// Number of pixel splits the tcgen05 kernel uses for this layer, 0 = shape not covered (conv_bwd.cu falls back to its
// mma.sync kernels: Cout 8 of the padded conv_last, Cin 16 of conv_first, ...).  SRB_WGRAD_TC=0 disables the path.
int wgrad_tc_splits(int B, int H, int W, int Cin, int Cout, int taps) {
    static const bool on = env_flag("SRB_WGRAD_TC", true);
    if (!on) return 0;
    long long nchunks;
    int tiles;
    if (taps == 1) {
        const int mt = (Cout + 127) / 128;
        if (Cin % 16 || Cin < 64 || Cin > 256 || Cout % 8 || Cout < 64 || mt > 3 || mt * Cin + 16 * mt > 512) return 0;
        tiles = 1;
        nchunks = ((long long)B * H * W + WT_KP - 1) / WT_KP;
    } else if (taps == 9) {
        if (Cout % 32) return 0;
        if (Cout <= 64) {
            if (Cin % 8 || Cin < 64) return 0;
            tiles = ((Cin + 127) / 128) * (Cout / 32);
        } else {
            if (Cin % 32) return 0;
            tiles = ((Cout + 127) / 128) * (Cin / 32);
        }
        nchunks = (long long)B * ((H + WT_TH - 1) / WT_TH) * ((W + WT_TW - 1) / WT_TW);
    } else {
        return 0;
    }
    if (nchunks > 0x7fffffff / WT_KP) return 0;
    long long ks = num_sms() / tiles;
    if (ks < 1) ks = 1;
    if (ks > nchunks) ks = nchunks;
    return (int)ks;
}

template <int MODE>
static int launch_wt(const CUtensorMap& tmM, const CUtensorMap& tmN, const WtArgs& a, dim3 grid, cudaStream_t s) {
    static std::atomic<unsigned long long> smem_set{0};
    SRB_TRY(opt_in_dynamic_smem(wgrad_tc_kernel<MODE>, WT_SMEM_LIMIT, smem_set));
    const int smem = a.nstages * a.stage_bytes + WT_ONES_BYTES + 256 + 1024;
    wgrad_tc_kernel<MODE><<<grid, WT_THREADS, smem, s>>>(tmM, tmN, a);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// partial: [ksplit][taps][Cout][Cin] fp32, bpartial: [ksplit][Cout] fp32 or null; ksplit = wgrad_tc_splits(...) > 0
int wgrad_tc_launch(const bf16* x, int x_pitch, const bf16* dy, int dy_pitch, float* partial, float* bpartial, int ksplit,
                    int B, int H, int W, int Cin, int Cout, int taps, cudaStream_t s) {
    WtArgs a;
    memset(&a, 0, sizeof(a));
    a.partial = partial; a.bpartial = bpartial; a.Cin = Cin; a.Cout = Cout;
    a.mode = taps == 1 ? WT_L : (Cout <= 64 ? WT_C1 : WT_C2);
    const bf16* m_ptr = a.mode == WT_C1 ? x : dy;
    const bf16* n_ptr = a.mode == WT_C1 ? dy : x;
    const int m_pitch = a.mode == WT_C1 ? x_pitch : dy_pitch, n_pitch = a.mode == WT_C1 ? dy_pitch : x_pitch;
    const int Cm = a.mode == WT_C1 ? Cin : Cout, Cn = a.mode == WT_C1 ? Cout : Cin;
    CUtensorMap tmM, tmN;
    int tiles = 1;
    if (a.mode == WT_L) {
        const size_t npx = (size_t)B * H * W;
        a.nchunks = (int)((npx + WT_KP - 1) / WT_KP);
        a.m_tiles = (Cout + 127) / 128;
        a.m_atoms = 2 * a.m_tiles; a.m_atom_bytes = WT_FLAT_ATOM;
        a.n_atoms = (Cin + 63) / 64; a.n_atom_bytes = WT_FLAT_ATOM;
        a.n_n_chunks = 1;
        cuuint32_t box[2] = {64, (cuuint32_t)WT_KP};
        cuuint64_t dm[2] = {(cuuint64_t)Cm, (cuuint64_t)npx}, sm[1] = {(cuuint64_t)m_pitch * 2};
        cuuint64_t dn[2] = {(cuuint64_t)Cn, (cuuint64_t)npx}, sn[1] = {(cuuint64_t)n_pitch * 2};
        SRB_TRY(encode_bf16_map(&tmM, m_ptr, 2, dm, sm, box, CU_TENSOR_MAP_SWIZZLE_128B));
        SRB_TRY(encode_bf16_map(&tmN, n_ptr, 2, dn, sn, box, CU_TENSOR_MAP_SWIZZLE_128B));
    } else {
        a.tiles_x = (W + WT_TW - 1) / WT_TW;
        a.tiles_y = (H + WT_TH - 1) / WT_TH;
        a.nchunks = B * a.tiles_x * a.tiles_y;
        a.m_tiles = 1;
        a.m_atoms = 2; a.n_atoms = 1;
        a.m_atom_bytes = WT_TILE_ATOM;
        a.n_atom_bytes = WT_NPATCH_ATOM;
        a.n_n_chunks = Cn / 32;
        tiles = ((Cm + 127) / 128) * a.n_n_chunks;
        cuuint32_t box_m[4] = {64, (cuuint32_t)WT_TW, (cuuint32_t)WT_TH, 1}, box_n[4] = {32, (cuuint32_t)WT_PW, (cuuint32_t)WT_PH, 1};
        cuuint64_t dm[4] = {(cuuint64_t)Cm, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
        cuuint64_t sm[3] = {(cuuint64_t)m_pitch * 2, (cuuint64_t)W * m_pitch * 2, (cuuint64_t)H * W * m_pitch * 2};
        cuuint64_t dn[4] = {(cuuint64_t)Cn, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
        cuuint64_t sn[3] = {(cuuint64_t)n_pitch * 2, (cuuint64_t)W * n_pitch * 2, (cuuint64_t)H * W * n_pitch * 2};
        SRB_TRY(encode_bf16_map(&tmM, m_ptr, 4, dm, sm, box_m, CU_TENSOR_MAP_SWIZZLE_128B));
        SRB_TRY(encode_bf16_map(&tmN, n_ptr, 4, dn, sn, box_n, CU_TENSOR_MAP_SWIZZLE_64B));
    }
    a.stage_bytes = a.m_atoms * a.m_atom_bytes + a.n_atoms * a.n_atom_bytes;
    a.nstages = (WT_SMEM_LIMIT - 1024 - WT_ONES_BYTES - 256) / a.stage_bytes;
    if (a.nstages > WT_MAX_STAGES) a.nstages = WT_MAX_STAGES;
    if (a.nstages < 2) { set_error("wgrad_tc: stage of %d bytes does not fit twice", a.stage_bytes); return SRB_E_SHAPE; }
    a.chunks_per_split = (a.nchunks + ksplit - 1) / ksplit;
    const dim3 grid(ksplit, tiles);
    if (a.mode == WT_L) return launch_wt<WT_L>(tmM, tmN, a, grid, s);
    if (a.mode == WT_C1) return launch_wt<WT_C1>(tmM, tmN, a, grid, s);
    return launch_wt<WT_C2>(tmM, tmN, a, grid, s);
}

}  // namespace srb
