// Host-side TMA tensor-map encoding shared by the tensor-core conv kernels.
#pragma once
#include <cuda.h>

namespace srb {

// bf16 tiled tensor map (driver entry point resolved at run time: libsrb.so does not link libcuda)
int encode_bf16_map(CUtensorMap* m, const void* ptr, int rank, const cuuint64_t* dims, const cuuint64_t* strides_bytes,
                    const cuuint32_t* box, CUtensorMapSwizzle swz);
bool env_flag(const char* name, bool dflt);

}  // namespace srb
