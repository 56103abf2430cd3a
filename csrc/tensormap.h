// Host-side TMA descriptor construction (cuTensorMapEncodeTiled resolved at run time, so the
// library has no link-time dependency on libcuda and builds on a box without a driver).
#pragma once
#include <cuda.h>
#include <cstdint>

namespace b200 {

enum class TmDtype { BF16, F32 };

// Encode a rank-`rank` tiled tensor map with 128-byte swizzle.  dims/box are innermost-first,
// strides_bytes has rank-1 entries (stride of dims 1..rank-1).  Returns 0 on success.
int encode_tensor_map(CUtensorMap* out, TmDtype dt, int rank, const void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box, bool swizzle128 = true);

}  // namespace b200
