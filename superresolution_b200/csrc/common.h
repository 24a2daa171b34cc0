// Shared host-side declarations for the srb CUDA library (error convention, launch parameter structs).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stddef.h>

#include <atomic>

#include "../../include/srb.h"

namespace srb {

typedef __nv_bfloat16 bf16;

// thread-local error string behind srb_last_error()
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);  // records the message, returns SRB_E_CUDA

#define SRB_CUDA(x)                                          \
    do {                                                     \
        cudaError_t e__ = (x);                               \
        if (e__ != cudaSuccess) return cuda_fail(e__, #x);   \
    } while (0)
#define SRB_TRY(x)                \
    do {                          \
        int r__ = (x);            \
        if (r__ != SRB_OK) return r__; \
    } while (0)

enum Act : int { ACT_NONE = 0, ACT_LRELU = 1, ACT_GELU = 2 };

// One tensor-core conv / linear launch.  All activation tensors are NHWC ("pixel-major") with an explicit channel
// pitch so that a layer can read / write a channel slice of a wider buffer (dense-block concatenation is free).
//   v = acc + bias[c]; v = act(v); v = v*alpha1 + res1; v = v*alpha2 + res2 * (res2_cs ? res2_cs[b,c] : 1)
struct ConvArgs {
    int B, H, W;      // output (= input) spatial size; stride 1, zero "same" padding
    int Cin, Cout;    // reduction channels / output channels
    int taps;         // 9 (3x3) or 1 (1x1 == token-wise linear)
    const bf16* in;   // input base already offset to the first input channel
    int in_pitch;     // channels per pixel in the input buffer
    const bf16* w;    // packed weights [taps][Cout][Cin] bf16
    const float* bias;  // [Cout] or null
    int act;
    float slope;      // LeakyReLU slope
    float alpha1;
    const void* res1; int res1_f32; int res1_pitch;   // per pixel, channel index = output channel
    float alpha2;
    const void* res2; int res2_f32; int res2_pitch;
    const float* res2_cs;                             // optional [B][Cout] scale on res2
    bf16* out; int out_pitch;                         // bf16 output (base offset to first out channel) or null
    float* out32; int out32_pitch;                    // fp32 output or null
    int pixel_shuffle;                                // 1: out is [B,2H,2W,Cout/4] (PixelShuffle(2) fused in the store)
    float* colsum;                                    // optional [B * tiles_per_image][Cout] fp32: per-tile channel sums of the
                                                      // epilogue result (CAB average pool fused into cab.2; needs Cout == N tile)
    // training path (both optional): pre_out keeps the pre-activation acc + bias (bf16) next to the activated output;
    // gelu_grad_of multiplies the result by GELU'(x) of that saved pre-activation (the dgrad conv feeding a GELU)
    bf16* pre_out; int pre_pitch;
    const bf16* gelu_grad_of; int gelu_grad_pitch;
    const float* alpha1_samples;                      // optional [B]: per-image alpha1 (DropPath factor of the branch)
    // 1: weights / bias were produced by a kernel enqueued just before this launch (training path: pack kernels run
    // per step) -> every warp waits for the programmatic dependency before touching them.  0 (inference): weights are
    // constants packed long before, the weight TMA and the bias prologue overlap the previous layer's tail.
    int dyn_weights;
};

int conv_tc_launch(const ConvArgs& a, cudaStream_t stream);
// chain of dependent 3x3 convs (Cout 32 / 64, common B/H/W) in one persistent kernel with grid barriers (conv_chain.cu)
int conv_chain_launch(const ConvArgs* layers, int n, unsigned int* counters, cudaStream_t stream);
int conv_chain_max_layers();
int num_sms();  // SM count of the current device (cached per device)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device property of a kernel: opt in once per (kernel, device).
// `done` is the kernel's own bit mask of devices already configured (thread-safe; nn.DataParallel drives several
// devices from one process, reference scripts/Modello_supporto.py:103-105).
template <typename Kernel>
inline int opt_in_dynamic_smem(Kernel* kernel, int bytes, std::atomic<unsigned long long>& done) {
    int dev = 0;
    SRB_CUDA(cudaGetDevice(&dev));
    const unsigned long long bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_acquire) & bit)) {
        SRB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        done.fetch_or(bit, std::memory_order_release);
    }
    return SRB_OK;
}

}  // namespace srb
