// Host-side helpers shared by the translation units of libddp_b200.so.
#pragma once
#include <cuda_runtime.h>

namespace ddp {
// Records a thread-local error string and returns `code` (negative).
int set_error(int code, const char* msg);
int cuda_check(cudaError_t e, const char* what);
}  // namespace ddp
