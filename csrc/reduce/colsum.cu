// Column sum out[c] = sum_r x[r][c]: the bias gradient of the projection layers (autograd of
// reference module.py:43-45,75 with add_bias=True).  x is [rows][cols] (tall and thin: rows = B*T/N).
//
// Two stages so that the whole GPU reads x: stage 1 splits the rows over `splits` CTAs per 32-column block and
// writes fp32 partial sums to a scratch buffer [splits][cols]; stage 2 adds the partials in fixed order
// (deterministic) and converts to the output dtype.  Without scratch a single-stage kernel is used.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/ddp_b200.h"
#include "../common/host_util.h"

namespace ddp {

template <typename T> __device__ __forceinline__ float cs_to_f(T v);
template <> __device__ __forceinline__ float cs_to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float cs_to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float cs_to_f<__half>(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T cs_from_f(float v);
template <> __device__ __forceinline__ float cs_from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 cs_from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half cs_from_f<__half>(float v) { return __float2half_rn(v); }

// block (32, 8); grid (ceil(cols/32), splits).  partial == nullptr -> gridDim.y must be 1 and `out` is written.
template <typename T>
__global__ void colsum_stage1(const T* __restrict__ x, T* __restrict__ out, float* __restrict__ partial, long long rows,
                              int cols, long long ld) {
  __shared__ float part[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const long long per = (rows + gridDim.y - 1) / gridDim.y;
  const long long r0 = per * blockIdx.y;
  const long long r1 = (r0 + per < rows) ? r0 + per : rows;
  float s = 0.f;
  if (c < cols)
    for (long long r = r0 + threadIdx.y; r < r1; r += 8) s += cs_to_f(x[r * ld + c]);
  part[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += part[i][threadIdx.x];
    if (partial) {
      partial[static_cast<size_t>(blockIdx.y) * cols + c] = t;
    } else {
      out[c] = cs_from_f<T>(t);
    }
  }
}

template <typename T>
__global__ void colsum_stage2(const float* __restrict__ partial, T* __restrict__ out, int splits, int cols) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float t = 0.f;
  for (int s = 0; s < splits; ++s) t += partial[static_cast<size_t>(s) * cols + c];
  out[c] = cs_from_f<T>(t);
}

template <typename T>
static int run_colsum(const ddp_tensor* x, void* out, float* scratch, size_t scratch_bytes, cudaStream_t st) {
  const int cols = static_cast<int>(x->cols);
  const long long rows = x->rows;
  const dim3 block(32, 8);
  const unsigned col_blocks = static_cast<unsigned>((cols + 31) / 32);
  long long splits = 1;
  if (scratch) {
    const long long by_rows = (rows + 255) / 256;                        // >= 256 rows per CTA
    const long long by_grid = (4LL * ddp_num_sms() + col_blocks - 1) / col_blocks;  // ~4 CTAs per SM
    const long long by_mem = static_cast<long long>(scratch_bytes / (sizeof(float) * static_cast<size_t>(cols)));
    splits = by_rows < by_grid ? by_rows : by_grid;
    if (splits > by_mem) splits = by_mem;
    if (splits < 1) splits = 1;
  }
  if (splits == 1) {
    colsum_stage1<T><<<dim3(col_blocks, 1), block, 0, st>>>(static_cast<const T*>(x->ptr), static_cast<T*>(out), nullptr,
                                                           rows, cols, x->ld);
  } else {
    colsum_stage1<T><<<dim3(col_blocks, static_cast<unsigned>(splits)), block, 0, st>>>(
        static_cast<const T*>(x->ptr), static_cast<T*>(out), scratch, rows, cols, x->ld);
    colsum_stage2<T><<<(cols + 127) / 128, 128, 0, st>>>(scratch, static_cast<T*>(out), static_cast<int>(splits), cols);
  }
  return cuda_check(cudaGetLastError(), "colsum launch");
}

static int dispatch_colsum(const ddp_tensor* x, void* out, float* scratch, size_t scratch_bytes, cudaStream_t st) {
  if (x->rows < 1 || x->cols < 1) return set_error(-4, "colsum: empty input");
  switch (x->dtype) {
    case DDP_F32: return run_colsum<float>(x, out, scratch, scratch_bytes, st);
    case DDP_BF16: return run_colsum<__nv_bfloat16>(x, out, scratch, scratch_bytes, st);
    case DDP_F16: return run_colsum<__half>(x, out, scratch, scratch_bytes, st);
    default: return set_error(-4, "colsum: unsupported dtype");
  }
}

}  // namespace ddp

using namespace ddp;

extern "C" size_t ddp_colsum_scratch_bytes(int64_t rows, int64_t cols) {
  const long long col_blocks = (cols + 31) / 32;
  long long splits = (rows + 255) / 256;
  const long long by_grid = (4LL * ddp_num_sms() + col_blocks - 1) / col_blocks;
  if (splits > by_grid) splits = by_grid;
  if (splits < 2) return 0;
  return static_cast<size_t>(splits) * static_cast<size_t>(cols) * sizeof(float);
}

extern "C" int ddp_colsum(const ddp_tensor* x, void* out, void* scratch, size_t scratch_bytes, void* stream) {
  return dispatch_colsum(x, out, static_cast<float*>(scratch), scratch_bytes, static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------------------
// Sum of S partial products (split-K weight gradients, ops/linear.py): out[i] = sum_s part[s][i], fp32 accumulate,
// 16-byte vectors.  n must be a multiple of 16 / sizeof(T) and the pointers 16-byte aligned.
namespace ddp {
template <typename T>
__global__ void sum_partials_kernel(const T* __restrict__ part, T* __restrict__ out, int splits, long long n_vec,
                                    long long stride_vec) {
  constexpr int VE = 16 / sizeof(T);
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_vec) return;
  float acc[VE];
#pragma unroll
  for (int e = 0; e < VE; ++e) acc[e] = 0.f;
  const uint4* src = reinterpret_cast<const uint4*>(part) + i;
  for (int s = 0; s < splits; ++s) {
    const uint4 v = src[s * stride_vec];
    const T* t = reinterpret_cast<const T*>(&v);
#pragma unroll
    for (int e = 0; e < VE; ++e) acc[e] += static_cast<float>(t[e]);
  }
  uint4 o;
  T* t = reinterpret_cast<T*>(&o);
#pragma unroll
  for (int e = 0; e < VE; ++e) t[e] = static_cast<T>(acc[e]);
  reinterpret_cast<uint4*>(out)[i] = o;
}
}  // namespace ddp

extern "C" int ddp_sum_partials(const void* part, void* out, int dtype, int splits, int64_t n, void* stream) {
  const int es = dtype == DDP_F32 ? 4 : 2;
  const int ve = 16 / es;
  if (splits < 1 || n < 1) return set_error(-4, "sum_partials: empty input");
  if (n % ve || (reinterpret_cast<uintptr_t>(part) & 15) || (reinterpret_cast<uintptr_t>(out) & 15))
    return set_error(-2, "sum_partials: length must be a multiple of 16 bytes and the pointers 16-byte aligned");
  const long long n_vec = n / ve;
  const unsigned grid = static_cast<unsigned>((n_vec + 255) / 256);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case DDP_F32:
      sum_partials_kernel<float><<<grid, 256, 0, st>>>((const float*)part, (float*)out, splits, n_vec, n_vec);
      break;
    case DDP_BF16:
      sum_partials_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)part, (__nv_bfloat16*)out, splits,
                                                                n_vec, n_vec);
      break;
    case DDP_F16:
      sum_partials_kernel<__half><<<grid, 256, 0, st>>>((const __half*)part, (__half*)out, splits, n_vec, n_vec);
      break;
    default: return set_error(-4, "sum_partials: unsupported dtype");
  }
  return cuda_check(cudaGetLastError(), "sum_partials launch");
}

// ---------------------------------------------------------------------------------------------------------
// Allocation-free variant.
//
// ddp_colsum needs a scratch buffer from its caller, and the usual caller gets one from an allocator.  Some
// callers have no allocator to ask: a serving process that gave every byte of device memory to its own pool
// at start-up and has none left for cudaMalloc to hand out; a latency-bound step loop that must not run
// cudaMalloc/cudaFree because both stall every stream on the device while they work; a plug-in that runs
// while the host framework holds its allocator lock and therefore may not re-enter it.  For them the partial
// sums live in the fixed array below, which is part of the library image: every GPU that loads
// libddp_b200.so sets aside DDP_COLSUM_NOALLOC_SCRATCH_BYTES for it until the process exits, used or not.
//
// The array is shared by every call on a device and is never cleared: stage 1 writes each partial-sum slot
// that stage 2 reads, so no slot has to start at zero and stale contents from an earlier call are harmless.
// Two calls whose kernels may be in flight on the same GPU at the same time (different streams, or another
// host thread) would write the same slots and race; callers must order such calls themselves.
#define DDP_COLSUM_NOALLOC_MAX_COLS 16384
#define DDP_COLSUM_NOALLOC_SPLITS 64
#define DDP_COLSUM_NOALLOC_SCRATCH_BYTES \
  (static_cast<size_t>(DDP_COLSUM_NOALLOC_MAX_COLS) * DDP_COLSUM_NOALLOC_SPLITS * sizeof(float)) /* 4 MiB */

static __device__ __align__(256) float g_colsum_noalloc_scratch[DDP_COLSUM_NOALLOC_MAX_COLS * DDP_COLSUM_NOALLOC_SPLITS];

// In host code `g_colsum_noalloc_scratch` names a host-side shadow of the symbol, not device memory, so it
// must never be passed to a kernel; the device address comes from cudaGetSymbolAddress, asked once per
// device while that device is current.
static int colsum_noalloc_scratch(float** out) {
  // Written without a lock: two host threads that race on the first call for a device both ask the runtime and
  // both store the same address (the symbol has one address per device), so the race is benign.
  static float* cached[64] = {};
  int dev = -1;
  if (int e = cuda_check(cudaGetDevice(&dev), "cudaGetDevice")) return e;
  if (dev < 0 || dev >= 64) return set_error(-4, "colsum_noalloc: device ordinal out of range");
  if (cached[dev] == nullptr) {
    void* p = nullptr;
    if (int e = cuda_check(cudaGetSymbolAddress(&p, g_colsum_noalloc_scratch), "cudaGetSymbolAddress(colsum scratch)"))
      return e;
    cached[dev] = static_cast<float*>(p);
  }
  *out = cached[dev];
  return 0;
}

extern "C" size_t ddp_colsum_noalloc_max_cols(void) { return DDP_COLSUM_NOALLOC_MAX_COLS; }

extern "C" int ddp_colsum_noalloc(const ddp_tensor* x, void* out, void* stream) {
  if (x->cols > DDP_COLSUM_NOALLOC_MAX_COLS)
    return set_error(-5, "ddp_colsum_noalloc: more than 16384 columns do not fit the built-in scratch array; "
                         "use ddp_colsum with caller-provided scratch");
  float* scratch = nullptr;
  if (int e = colsum_noalloc_scratch(&scratch)) return e;
  return dispatch_colsum(x, out, scratch, DDP_COLSUM_NOALLOC_SCRATCH_BYTES, static_cast<cudaStream_t>(stream));
}
