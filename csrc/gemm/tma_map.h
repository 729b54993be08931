// 5-D TMA tensor-map construction shared by the kernels' host sides (defined in gemm_api.cu).
#pragma once
#include <cuda.h>
#include <cstdint>

namespace ddp {
// dims / box innermost first; strides in bytes for dims 1..4 (multiples of 16).  128-byte swizzle
// (the 32-byte-atom flavour for MN-major tf32 operands).  Returns 0 or a negative error code.
int make_map(CUtensorMap* m, int dt, void* ptr, const uint64_t dims[5], const uint64_t strides[4],
             const uint32_t box[5], bool mn_major);
}  // namespace ddp
