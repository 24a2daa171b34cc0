// Halo-blend stitch for tiled whole-field inference (BASELINE config 3).  The reference only iterates
// pre-cut patches (scripts/Modello_5.py:72-85), so the tile/blend semantics are defined here and restated in
// oracle/tiling.py: every tile output is weighted by a separable ramp w(i) = min(i+1, n-i, ov)/ov, accumulated into
// fp32 (sum w*y, sum w) planes in a fixed order (deterministic), then normalised, clipped and cast like
// scripts/Modello_5.py:29-33.
// Tiles are independent: a rank blends only the rows of its band -> no collective on the data path.
#include "common.h"
#include "ops.h"

namespace srb {

__device__ __forceinline__ float ramp(int i, int n, int ov) {
    const int m = min(min(i + 1, n - i), ov);
    return (float)m / (float)ov;
}

// ONE tile per launch: its pixels are distinct output pixels, so the accumulation is a plain read-add-write, and the
// order in which overlapping tiles reach a pixel is the launch order (the stream's), i.e. the plan's raster order:
// the stitched frame is bit-reproducible run to run.  (One launch over a whole batch of tiles needed float atomics,
// whose order varies.)
__global__ void __launch_bounds__(256)
blend_tile_kernel(const float* __restrict__ tile, const int* __restrict__ oy, const int* __restrict__ ox, int t,
                  int th, int tw, int ov, float* __restrict__ acc, float* __restrict__ wsum, int FH, int FW,
                  int row0, int row1) {
    const int n = th * tw;
    const int y0 = oy[t], x0 = ox[t];
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
        const int ty = r / tw, tx = r % tw;
        const int y = y0 + ty, x = x0 + tx;
        if (y < row0 || y >= row1 || y >= FH || x >= FW || x < 0) continue;
        const float w = ramp(ty, th, ov) * ramp(tx, tw, ov);
        const size_t o = (size_t)(y - row0) * FW + x;
        acc[o] = fmaf(w, tile[r], acc[o]);
        wsum[o] += w;
    }
}

__global__ void __launch_bounds__(256)
blend_finalize_kernel(const float* __restrict__ acc, const float* __restrict__ wsum, void* __restrict__ out,
                      int out_dtype, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float w = wsum[i];
        const float v = w > 0.f ? acc[i] / w : 0.f;
        if (out_dtype == SRB_U16) ((uint16_t*)out)[i] = (uint16_t)(fminf(fmaxf(v, 0.f), 1.f) * 65535.f);
        else ((float*)out)[i] = v;
    }
}

int blend_tiles(const float* tiles, const int* oy, const int* ox, int T, int th, int tw, int ov, float* acc,
                float* wsum, int FH, int FW, int row0, int row1, cudaStream_t s) {
    if (ov <= 0 || ov > th || ov > tw) { set_error("blend_tiles: bad overlap %d", ov); return SRB_E_ARG; }
    const size_t per = (size_t)th * tw;
    if (per > 0x7fffffff) { set_error("blend_tiles: tile too large"); return SRB_E_SHAPE; }
    int grid = (int)((per + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    for (int t = 0; t < T; ++t) {
        blend_tile_kernel<<<grid, 256, 0, s>>>(tiles + (size_t)t * per, oy, ox, t, th, tw, ov, acc, wsum, FH, FW, row0, row1);
        SRB_CUDA(cudaGetLastError());
    }
    return SRB_OK;
}

// Tile gather: tiles[t] = frame[iy[t] : iy[t] + tile, ix[t] : ix[t] + tile] for T tiles whose origins live on the device
// (one launch per batch instead of one slice copy per tile; the batch then feeds srb_forward directly).
template <typename T>
__global__ void __launch_bounds__(256)
gather_tiles_kernel(const T* __restrict__ frame, int FW, const int* __restrict__ iy, const int* __restrict__ ix, int nt,
                    int tile, T* __restrict__ out) {
    const size_t per = (size_t)tile * tile;
    const size_t n = per * nt;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int t = (int)(i / per), r = (int)(i % per);
        out[i] = frame[(size_t)(iy[t] + r / tile) * FW + ix[t] + r % tile];
    }
}

int gather_tiles(const void* frame, int dtype, int FH, int FW, const int* iy, const int* ix, int T, int tile, void* out,
                 cudaStream_t s) {
    if (tile <= 0 || tile > FH || tile > FW) { set_error("gather_tiles: tile %d does not fit the %dx%d frame", tile, FH, FW); return SRB_E_SHAPE; }
    const size_t n = (size_t)T * tile * tile;
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    if (dtype == SRB_U16) gather_tiles_kernel<uint16_t><<<grid, 256, 0, s>>>((const uint16_t*)frame, FW, iy, ix, T, tile, (uint16_t*)out);
    else if (dtype == SRB_F32) gather_tiles_kernel<float><<<grid, 256, 0, s>>>((const float*)frame, FW, iy, ix, T, tile, (float*)out);
    else { set_error("gather_tiles: frame dtype must be U16 or F32"); return SRB_E_DTYPE; }
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int blend_finalize(const float* acc, const float* wsum, void* out, int out_dtype, size_t n, cudaStream_t s) {
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    blend_finalize_kernel<<<grid, 256, 0, s>>>(acc, wsum, out, out_dtype, n);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // namespace srb
