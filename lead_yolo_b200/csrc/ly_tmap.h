// Host-side TMA descriptor encoding (implemented in capi.cu).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace ly {
// bf16 tensor, 128B swizzle, zero fill out of bounds. dims/box: `rank` entries (innermost first);
// strides_bytes: rank-1 entries (stride of dim 1..rank-1).
int tmap_encode_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box);
}  // namespace ly
