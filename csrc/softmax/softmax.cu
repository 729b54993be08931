// Row-local kernels of the attention layer (reference: distributed_dot_product/module.py:65-67 and its
// autograd): fused scale + boolean mask + softmax over the T/N x T score slab, its backward, and the
// column sum used for bias gradients.  The whole row is local to a rank, so no communication is needed.
//
// Two code paths per kernel:
//   * register-resident: the row lives in registers as packed 16-byte vectors -> one HBM read + one write
//   * streaming: arbitrary length / alignment, online max+sum then a second (L2-assisted) read.
// Work item = (outer batch, row); the head loop is innermost so a mask row is re-read from L1/L2, not HBM.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cfloat>
#include <cstdlib>
#include <cstdint>

#include "../../include/ddp_b200.h"
#include "../common/host_util.h"
#include "../common/ptx.cuh"

namespace ddp {

struct SmParams {
  const void* x;  // fwd: scores; bwd: y (probabilities)
  const void* dy; // bwd only
  void* out;      // fwd: y; bwd: dx
  const uint8_t* mask;
  const float* delta;  // bwd: optional precomputed rowsum(dy*y), [b1][b0][rows]
  long long x_ld, x_sb0, x_sb1;
  long long dy_ld, dy_sb0, dy_sb1;
  long long o_ld, o_sb0, o_sb1;
  long long m_ld, m_sb1;
  int rows, cols, b0, b1;
  float scale;
};

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f<__half>(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f<__half>(float v) { return __float2half_rn(v); }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Block-wide reductions through a small shared array; every thread gets the result.
template <bool kMax>
__device__ __forceinline__ float block_reduce(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = kMax ? warp_max(v) : warp_sum(v);
  __syncthreads();  // protect `red` from the previous use
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : (kMax ? -INFINITY : 0.f);
  r = kMax ? warp_max(r) : warp_sum(r);
  return r;
}

constexpr float kLog2e = 1.4426950408889634f;

// 2^x on the SFU without the range fix-up code of exp2f (ex2(-inf) = 0, ex2(NaN) = NaN).
__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ------------------------------------------------------------------ forward, register resident
template <typename T> struct Packed2 { using type = __nv_bfloat162; };
template <> struct Packed2<__half> { using type = __half2; };

// Row maximum of the (mask-applied) packed vectors without unpacking: HMNMX2 on 16-bit pairs.
template <typename T>
__device__ __forceinline__ float vec_max(const uint4& q, float mx) {
  if constexpr (sizeof(T) == 4) {
    const float* e = reinterpret_cast<const float*>(&q);
    return fmaxf(fmaxf(fmaxf(mx, e[0]), fmaxf(e[1], e[2])), e[3]);
  } else {
    using P2 = typename Packed2<T>::type;
    const P2* e = reinterpret_cast<const P2*>(&q);
    const P2 m = __hmax2(__hmax2(e[0], e[1]), __hmax2(e[2], e[3]));
    return fmaxf(mx, fmaxf(to_f(m.x), to_f(m.y)));
  }
}

// Mask bits of this thread's VPT vectors (bit j*VE+k = element k of vector j), built once per (batch, row) and
// reused for every head: the mask loads leave the per-row critical path (they were its main stall).
template <int VPT, int VE, int THREADS>
__device__ __forceinline__ void load_mask_bits(const uint8_t* mrow, int nvec, uint32_t (&bits)[(VPT * VE + 31) / 32]) {
#pragma unroll
  for (int w = 0; w < (VPT * VE + 31) / 32; ++w) bits[w] = 0u;
  if (mrow == nullptr) return;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * THREADS;
    if (v < nvec) {
      unsigned long long mw;
      if constexpr (VE == 8) {
        mw = *reinterpret_cast<const unsigned long long*>(mrow + static_cast<long long>(v) * VE);
      } else {
        mw = *reinterpret_cast<const uint32_t*>(mrow + static_cast<long long>(v) * VE);
      }
      uint32_t b = 0;
#pragma unroll
      for (int k = 0; k < VE; ++k) b |= ((mw >> (8 * k)) & 0xffull) ? (1u << k) : 0u;
      bits[(i * VE) / 32] |= b << ((i * VE) % 32);
    }
  }
}

template <typename T, int VPT, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) softmax_fwd_reg(const SmParams p) {
  constexpr int VE = 16 / sizeof(T);
  __shared__ float red[32];
  const int nvec = p.cols / VE;
  const int items = p.b1 * p.rows;
  const float s2 = p.scale * kLog2e;
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int b1i = item / p.rows, row = item - b1i * p.rows;
    const uint8_t* mrow = p.mask ? p.mask + b1i * p.m_sb1 + static_cast<long long>(row) * p.m_ld : nullptr;
    uint32_t mbits[(VPT * VE + 31) / 32];
    load_mask_bits<VPT, VE, THREADS>(mrow, nvec, mbits);
    for (int h = 0; h < p.b0; ++h) {
      const T* xr = reinterpret_cast<const T*>(p.x) + b1i * p.x_sb1 + h * p.x_sb0 + static_cast<long long>(row) * p.x_ld;
      T* yr = reinterpret_cast<T*>(p.out) + b1i * p.o_sb1 + h * p.o_sb0 + static_cast<long long>(row) * p.o_ld;
      uint4 q[VPT];
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec) q[i] = *reinterpret_cast<const uint4*>(xr + static_cast<long long>(v) * VE);
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec) {
          T* e = reinterpret_cast<T*>(&q[i]);
          const uint32_t mb = (mbits[(i * VE) / 32] >> ((i * VE) % 32)) & ((1u << VE) - 1u);
          if (mb != 0u) {
#pragma unroll
            for (int k = 0; k < VE; ++k)
              if ((mb >> k) & 1u) e[k] = from_f<T>(-INFINITY);
          }
          mx = vec_max<T>(q[i], mx);  // max of raw scores; scale > 0 commutes with max
        }
      }
      mx = block_reduce<true>(mx, red) * p.scale;
      const float m2 = mx * kLog2e;
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec) {
          T* e = reinterpret_cast<T*>(&q[i]);
#pragma unroll
          for (int k = 0; k < VE; ++k) {
            // one FFMA + one MUFU.EX2; fully masked row: -inf - -inf = NaN, as the reference
            const float ex = ex2_fast(fmaf(to_f(e[k]), s2, -m2));
            sum += ex;
            e[k] = from_f<T>(ex);
          }
        }
      }
      sum = block_reduce<false>(sum, red);
      const float inv = 1.f / sum;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec) {
          T* e = reinterpret_cast<T*>(&q[i]);
#pragma unroll
          for (int k = 0; k < VE; ++k) e[k] = from_f<T>(to_f(e[k]) * inv);
          *reinterpret_cast<uint4*>(yr + static_cast<long long>(v) * VE) = q[i];
        }
      }
    }
  }
}

// ------------------------------------------------------------------ forward, long rows (software pipelined)
// Rows of 32-128 KB need most of the register file of an SM, so only one CTA fits and the register-resident
// kernel above alternates "load the row" and "compute the row" with nothing to overlap them.  Here the NEXT
// row streams into shared memory through the TMA engine (cp.async.bulk, mbarrier completion) while the
// current one is being reduced, exponentiated and written: per row the SM pays max(load, compute + store)
// instead of their sum.
template <typename T, int VPT, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) softmax_fwd_pipe(const SmParams p) {
  constexpr int VE = 16 / sizeof(T);
  extern __shared__ __align__(128) uint8_t sm_raw[];
  __shared__ float red[32];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t row_smem = smem_u32(sm_raw);
  const int nvec = p.cols / VE;
  const uint32_t row_bytes = static_cast<uint32_t>(p.cols) * sizeof(T);
  const int items = p.b1 * p.rows;
  const float s2 = p.scale * kLog2e;
  // this CTA's rows, in order: item = blockIdx.x + k * gridDim.x, heads innermost (mask row reuse)
  const int my_items = (items - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const long long my_rows = static_cast<long long>(blockIdx.x < static_cast<unsigned>(items) ? my_items : 0) * p.b0;
  auto row_ptr = [&](long long r, const void* base, long long ld, long long sb0, long long sb1) {
    const int it = static_cast<int>(r / p.b0), h = static_cast<int>(r - static_cast<long long>(it) * p.b0);
    const int item = blockIdx.x + it * gridDim.x;
    const int b1i = item / p.rows, row = item - b1i * p.rows;
    return reinterpret_cast<const T*>(base) + b1i * sb1 + h * sb0 + static_cast<long long>(row) * ld;
  };
  auto issue_row = [&](long long r) {  // one thread
    const T* src = row_ptr(r, p.x, p.x_ld, p.x_sb0, p.x_sb1);
    mbar_arrive_expect_tx(&bar, row_bytes);
    for (uint32_t off = 0; off < row_bytes; off += 32768u) {
      const uint32_t n = row_bytes - off < 32768u ? row_bytes - off : 32768u;
      bulk_load_1d(row_smem + off, reinterpret_cast<const uint8_t*>(src) + off, n, &bar);
    }
  };
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0 && my_rows > 0) issue_row(0);
  uint32_t phase = 0;
  uint32_t mbits[(VPT * VE + 31) / 32];
  for (long long r = 0; r < my_rows; ++r) {
    const int it = static_cast<int>(r / p.b0);
    const int item = blockIdx.x + it * gridDim.x;
    const int b1i = item / p.rows, row = item - b1i * p.rows;
    if (r % p.b0 == 0) {  // new (batch, row): its mask bits serve all heads
      const uint8_t* mrow = p.mask ? p.mask + b1i * p.m_sb1 + static_cast<long long>(row) * p.m_ld : nullptr;
      load_mask_bits<VPT, VE, THREADS>(mrow, nvec, mbits);
    }
    T* yr = const_cast<T*>(row_ptr(r, p.out, p.o_ld, p.o_sb0, p.o_sb1));
    mbar_wait(&bar, phase, 6);
    phase ^= 1;
    uint4 q[VPT];
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * THREADS;
      if (v < nvec) q[i] = ld_shared_v4(row_smem + static_cast<uint32_t>(v) * 16u);
    }
    __syncthreads();  // the whole row is in registers: the buffer may take the next row
    if (threadIdx.x == 0 && r + 1 < my_rows) issue_row(r + 1);
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * THREADS;
      if (v < nvec) {
        T* e = reinterpret_cast<T*>(&q[i]);
        const uint32_t mb = (mbits[(i * VE) / 32] >> ((i * VE) % 32)) & ((1u << VE) - 1u);
        if (mb != 0u) {
#pragma unroll
          for (int k = 0; k < VE; ++k)
            if ((mb >> k) & 1u) e[k] = from_f<T>(-INFINITY);
        }
        mx = vec_max<T>(q[i], mx);
      }
    }
    mx = block_reduce<true>(mx, red) * p.scale;
    const float m2 = mx * kLog2e;
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * THREADS;
      if (v < nvec) {
        T* e = reinterpret_cast<T*>(&q[i]);
#pragma unroll
        for (int k = 0; k < VE; ++k) {
          const float ex = ex2_fast(fmaf(to_f(e[k]), s2, -m2));
          sum += ex;
          e[k] = from_f<T>(ex);
        }
      }
    }
    sum = block_reduce<false>(sum, red);
    const float inv = 1.f / sum;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * THREADS;
      if (v < nvec) {
        T* e = reinterpret_cast<T*>(&q[i]);
#pragma unroll
        for (int k = 0; k < VE; ++k) e[k] = from_f<T>(to_f(e[k]) * inv);
        *reinterpret_cast<uint4*>(yr + static_cast<long long>(v) * VE) = q[i];
      }
    }
  }
}

// ------------------------------------------------------------------ forward, long rows split over a cluster
// A 64-128 KB row fills the register file of an SM, so at most one row per SM is in flight and its phases
// (load, max, exp, sum, scale, store) run back to back; the exp phase alone is SFU-bound (8 ex2/clk/SM).
// Splitting every row over a thread-block cluster of CS CTAs of 128 threads puts 4 independent row slices on
// each SM, whose phases overlap; the two row-wide reductions go through distributed shared memory.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(const float* local_ptr, uint32_t cta) {
  uint32_t raddr;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(local_ptr)), "r"(cta));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(raddr) : "memory");
  return v;
}

template <typename T, int CS>
__global__ void __launch_bounds__(128, 4) softmax_fwd_cluster(const SmParams p) {
  constexpr int VE = 16 / sizeof(T);
  constexpr int THREADS = 128;
  constexpr int VPT = 16;
  __shared__ float red[32];
  __shared__ float part[2][2];  // [row parity][0 = max, 1 = sum] of this CTA's slice
  const uint32_t crank = cluster_ctarank();
  const int cluster_id = blockIdx.x / CS;
  const int n_clusters = gridDim.x / CS;
  const int nvec = p.cols / VE;
  const int nvec_cta = nvec / CS;          // host guarantees nvec % CS == 0 and nvec_cta <= THREADS * VPT
  const int v0 = crank * nvec_cta;
  const int items = p.b1 * p.rows;
  const float s2 = p.scale * kLog2e;
  int parity = 0;
  for (int item = cluster_id; item < items; item += n_clusters) {
    const int b1i = item / p.rows, row = item - b1i * p.rows;
    const uint8_t* mrow = p.mask ? p.mask + b1i * p.m_sb1 + static_cast<long long>(row) * p.m_ld : nullptr;
    for (int h = 0; h < p.b0; ++h, parity ^= 1) {
      const T* xr = reinterpret_cast<const T*>(p.x) + b1i * p.x_sb1 + h * p.x_sb0 + static_cast<long long>(row) * p.x_ld;
      T* yr = reinterpret_cast<T*>(p.out) + b1i * p.o_sb1 + h * p.o_sb0 + static_cast<long long>(row) * p.o_ld;
      uint4 q[VPT];
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec_cta) q[i] = *reinterpret_cast<const uint4*>(xr + static_cast<long long>(v0 + v) * VE);
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec_cta) {
          T* e = reinterpret_cast<T*>(&q[i]);
          if (mrow) {
            unsigned long long mw;
            if constexpr (VE == 8) {
              mw = *reinterpret_cast<const unsigned long long*>(mrow + static_cast<long long>(v0 + v) * VE);
            } else {
              mw = *reinterpret_cast<const uint32_t*>(mrow + static_cast<long long>(v0 + v) * VE);
            }
            if (mw != 0ull) {
#pragma unroll
              for (int k = 0; k < VE; ++k)
                if ((mw >> (8 * k)) & 0xffull) e[k] = from_f<T>(-INFINITY);
            }
          }
          mx = vec_max<T>(q[i], mx);
        }
      }
      mx = block_reduce<true>(mx, red);
      if (threadIdx.x == 0) part[parity][0] = mx;
      cluster_sync_all();
      float gmx = -INFINITY;
#pragma unroll
      for (int c = 0; c < CS; ++c) gmx = fmaxf(gmx, ld_dsmem_f32(&part[parity][0], c));
      const float m2 = gmx * p.scale * kLog2e;
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec_cta) {
          T* e = reinterpret_cast<T*>(&q[i]);
#pragma unroll
          for (int k = 0; k < VE; ++k) {
            const float ex = ex2_fast(fmaf(to_f(e[k]), s2, -m2));
            sum += ex;
            e[k] = from_f<T>(ex);
          }
        }
      }
      sum = block_reduce<false>(sum, red);
      if (threadIdx.x == 0) part[parity][1] = sum;
      cluster_sync_all();
      float gsum = 0.f;
#pragma unroll
      for (int c = 0; c < CS; ++c) gsum += ld_dsmem_f32(&part[parity][1], c);
      const float inv = 1.f / gsum;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * THREADS;
        if (v < nvec_cta) {
          T* e = reinterpret_cast<T*>(&q[i]);
#pragma unroll
          for (int k = 0; k < VE; ++k) e[k] = from_f<T>(to_f(e[k]) * inv);
          *reinterpret_cast<uint4*>(yr + static_cast<long long>(v0 + v) * VE) = q[i];
        }
      }
    }
  }
  cluster_sync_all();  // no CTA may exit while a peer can still read its shared memory
}

template <typename T, int CS>
static int launch_fwd_cluster(const SmParams& p, int items, int sms, cudaStream_t st) {
  int clusters = (sms * 4) / CS;
  if (clusters > items) clusters = items;
  if (clusters < 1) clusters = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(clusters * CS));
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cuda_check(cudaLaunchKernelEx(&cfg, softmax_fwd_cluster<T, CS>, p), "softmax_fwd_cluster launch");
}

// ------------------------------------------------------------------ forward, streaming (any shape)
template <typename T>
__global__ void __launch_bounds__(512) softmax_fwd_stream(const SmParams p) {
  __shared__ float red[32];
  const int items = p.b1 * p.rows;
  const float s2 = p.scale * kLog2e;
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int b1i = item / p.rows, row = item - b1i * p.rows;
    const uint8_t* mrow = p.mask ? p.mask + b1i * p.m_sb1 + static_cast<long long>(row) * p.m_ld : nullptr;
    for (int h = 0; h < p.b0; ++h) {
      const T* xr = reinterpret_cast<const T*>(p.x) + b1i * p.x_sb1 + h * p.x_sb0 + static_cast<long long>(row) * p.x_ld;
      T* yr = reinterpret_cast<T*>(p.out) + b1i * p.o_sb1 + h * p.o_sb0 + static_cast<long long>(row) * p.o_ld;
      float mx = -INFINITY;
      for (int c = threadIdx.x; c < p.cols; c += blockDim.x) {
        const float v = (mrow && mrow[c]) ? -INFINITY : to_f(xr[c]) * p.scale;
        mx = fmaxf(mx, v);
      }
      mx = block_reduce<true>(mx, red);
      const float m2 = mx * kLog2e;
      float sum = 0.f;
      for (int c = threadIdx.x; c < p.cols; c += blockDim.x) {
        const float v = (mrow && mrow[c]) ? -INFINITY : to_f(xr[c]) * s2;
        sum += exp2f(v - m2);
      }
      sum = block_reduce<false>(sum, red);
      const float inv = 1.f / sum;
      for (int c = threadIdx.x; c < p.cols; c += blockDim.x) {
        const float v = (mrow && mrow[c]) ? -INFINITY : to_f(xr[c]) * s2;
        yr[c] = from_f<T>(exp2f(v - m2) * inv);
      }
      __syncthreads();  // in-place safety across heads is per row, nothing shared; keeps `red` reuse simple
    }
  }
}

// ------------------------------------------------------------------ backward, register resident
template <typename T, int VPT>
__global__ void __launch_bounds__(512) softmax_bwd_reg(const SmParams p) {
  constexpr int VE = 16 / sizeof(T);
  __shared__ float red[32];
  const int nvec = p.cols / VE;
  const long long items = static_cast<long long>(p.b1) * p.b0 * p.rows;
  for (long long item = blockIdx.x; item < items; item += gridDim.x) {
    const int row = static_cast<int>(item % p.rows);
    const int bh = static_cast<int>(item / p.rows);
    const int b1i = bh / p.b0, h = bh - b1i * p.b0;
    const T* yr = reinterpret_cast<const T*>(p.x) + b1i * p.x_sb1 + h * p.x_sb0 + static_cast<long long>(row) * p.x_ld;
    const T* gr = reinterpret_cast<const T*>(p.dy) + b1i * p.dy_sb1 + h * p.dy_sb0 + static_cast<long long>(row) * p.dy_ld;
    T* dr = reinterpret_cast<T*>(p.out) + b1i * p.o_sb1 + h * p.o_sb0 + static_cast<long long>(row) * p.o_ld;
    uint4 qy[VPT], qg[VPT];
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
        qy[i] = *reinterpret_cast<const uint4*>(yr + static_cast<long long>(v) * VE);
        qg[i] = *reinterpret_cast<const uint4*>(gr + static_cast<long long>(v) * VE);
      }
    }
    float dot;
    if (p.delta) {
      dot = p.delta[item];
    } else {
      dot = 0.f;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = threadIdx.x + i * blockDim.x;
        if (v < nvec) {
          const T* a = reinterpret_cast<const T*>(&qy[i]);
          const T* b = reinterpret_cast<const T*>(&qg[i]);
#pragma unroll
          for (int k = 0; k < VE; ++k) dot += to_f(a[k]) * to_f(b[k]);
        }
      }
      dot = block_reduce<false>(dot, red);
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
        const T* a = reinterpret_cast<const T*>(&qy[i]);
        T* b = reinterpret_cast<T*>(&qg[i]);
#pragma unroll
        for (int k = 0; k < VE; ++k) b[k] = from_f<T>(p.scale * to_f(a[k]) * (to_f(b[k]) - dot));
        *reinterpret_cast<uint4*>(dr + static_cast<long long>(v) * VE) = qg[i];
      }
    }
  }
}

// ------------------------------------------------------------------ backward, streaming
template <typename T, bool kVec>
__global__ void __launch_bounds__(512) softmax_bwd_stream(const SmParams p) {
  constexpr int VE = kVec ? 16 / sizeof(T) : 1;
  __shared__ float red[32];
  const int nvec = p.cols / VE;
  const long long items = static_cast<long long>(p.b1) * p.b0 * p.rows;
  for (long long item = blockIdx.x; item < items; item += gridDim.x) {
    const int row = static_cast<int>(item % p.rows);
    const int bh = static_cast<int>(item / p.rows);
    const int b1i = bh / p.b0, h = bh - b1i * p.b0;
    const T* yr = reinterpret_cast<const T*>(p.x) + b1i * p.x_sb1 + h * p.x_sb0 + static_cast<long long>(row) * p.x_ld;
    const T* gr = reinterpret_cast<const T*>(p.dy) + b1i * p.dy_sb1 + h * p.dy_sb0 + static_cast<long long>(row) * p.dy_ld;
    T* dr = reinterpret_cast<T*>(p.out) + b1i * p.o_sb1 + h * p.o_sb0 + static_cast<long long>(row) * p.o_ld;
    float dot;
    if (p.delta) {
      dot = p.delta[item];
    } else {
      dot = 0.f;
      for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
        if constexpr (kVec) {
          const uint4 a4 = *reinterpret_cast<const uint4*>(yr + static_cast<long long>(v) * VE);
          const uint4 b4 = *reinterpret_cast<const uint4*>(gr + static_cast<long long>(v) * VE);
          const T* a = reinterpret_cast<const T*>(&a4);
          const T* b = reinterpret_cast<const T*>(&b4);
#pragma unroll
          for (int k = 0; k < VE; ++k) dot += to_f(a[k]) * to_f(b[k]);
        } else {
          dot += to_f(yr[v]) * to_f(gr[v]);
        }
      }
      dot = block_reduce<false>(dot, red);
    }
    for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
      if constexpr (kVec) {
        const uint4 a4 = *reinterpret_cast<const uint4*>(yr + static_cast<long long>(v) * VE);
        uint4 b4 = *reinterpret_cast<const uint4*>(gr + static_cast<long long>(v) * VE);
        const T* a = reinterpret_cast<const T*>(&a4);
        T* b = reinterpret_cast<T*>(&b4);
#pragma unroll
        for (int k = 0; k < VE; ++k) b[k] = from_f<T>(p.scale * to_f(a[k]) * (to_f(b[k]) - dot));
        *reinterpret_cast<uint4*>(dr + static_cast<long long>(v) * VE) = b4;
      } else {
        dr[v] = from_f<T>(p.scale * to_f(yr[v]) * (to_f(gr[v]) - dot));
      }
    }
  }
}

// ------------------------------------------------------------------ rowdot: delta[r] = sum_c a[r][c]*b[r][c]
template <typename T>
__global__ void rowdot_kernel(const T* a, const T* b, float* out, long long rows, int cols, long long a_ld, long long b_ld) {
  const long long r = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  float s = 0.f;
  for (int c = lane; c < cols; c += 32) s += to_f(a[r * a_ld + c]) * to_f(b[r * b_ld + c]);
  s = warp_sum(s);
  if (lane == 0) out[r] = s;
}

static bool aligned16(const void* p, long long ld, long long sb0, long long sb1, int es) {
  return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (ld * es) % 16 == 0 && (sb0 * es) % 16 == 0 && (sb1 * es) % 16 == 0;
}

template <typename T>
static int launch_fwd(const SmParams& p, bool vec_ok, cudaStream_t st) {
  constexpr int VE = 16 / sizeof(T);
  const int items = p.b1 * p.rows;
  const int sms = ddp_num_sms();
  const int nvec = p.cols / VE;
  if (vec_ok && nvec <= 1024 * 8 && p.scale > 0.f) {
    // (vectors per thread, threads, CTAs per SM): as many independent rows in flight per SM as registers allow
    if (nvec <= 256) {
      softmax_fwd_reg<T, 2, 128, 8><<<min(items, sms * 8), 128, 0, st>>>(p);
    } else if (nvec <= 512) {
      softmax_fwd_reg<T, 4, 128, 8><<<min(items, sms * 8), 128, 0, st>>>(p);
    } else if (nvec <= 1024) {
      softmax_fwd_reg<T, 8, 128, 6><<<min(items, sms * 6), 128, 0, st>>>(p);
    } else if (nvec <= 2048) {
      softmax_fwd_reg<T, 8, 256, 3><<<min(items, sms * 3), 256, 0, st>>>(p);
    } else if (nvec % 2 == 0 && nvec / 2 <= 2048 && getenv("DDP_SOFTMAX_CLUSTER")) {  // experimental, slower
      return launch_fwd_cluster<T, 2>(p, items, sms, st);   // rows of 32-64 KB: 2 CTAs per row
    } else if (nvec % 4 == 0 && nvec / 4 <= 2048 && getenv("DDP_SOFTMAX_CLUSTER")) {
      return launch_fwd_cluster<T, 4>(p, items, sms, st);   // rows of 64-128 KB: 4 CTAs per row
    } else {
      // long rows: next row streams into shared memory (TMA bulk copy) while the current one is processed
      const int smem = p.cols * static_cast<int>(sizeof(T));
      static bool configured[2][64] = {};
      int dev = 0;
      cudaGetDevice(&dev);
      if (nvec <= 4096) {
        // rows up to 64 KB: 256 threads x 16 vectors, two CTAs (= two rows in different phases) per SM
        auto kern = softmax_fwd_pipe<T, 16, 256, 2>;
        if (dev < 64 && !configured[0][dev]) {
          cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
          configured[0][dev] = true;
        }
        kern<<<min(items, sms * 2), 256, smem, st>>>(p);
      } else {
        auto kern = softmax_fwd_pipe<T, 16, 512, 1>;
        if (dev < 64 && !configured[1][dev]) {
          cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072);
          configured[1][dev] = true;
        }
        kern<<<min(items, sms), 512, smem, st>>>(p);
      }
    }
  } else {
    softmax_fwd_stream<T><<<min(items, sms * 4), p.cols >= 4096 ? 512 : 256, 0, st>>>(p);
  }
  return cuda_check(cudaGetLastError(), "softmax_fwd launch");
}

template <typename T>
static int launch_bwd(const SmParams& p, bool vec_ok, cudaStream_t st) {
  constexpr int VE = 16 / sizeof(T);
  const long long items = static_cast<long long>(p.b1) * p.b0 * p.rows;
  const int sms = ddp_num_sms();
  const int nvec = p.cols / VE;
  const int cap = static_cast<int>(items < (1 << 30) ? items : (1 << 30));
  if (vec_ok && nvec <= 512 * 8) {
    if (nvec <= 256) {
      softmax_bwd_reg<T, 2><<<min(cap, sms * 8), 128, 0, st>>>(p);
    } else if (nvec <= 1024) {
      softmax_bwd_reg<T, 4><<<min(cap, sms * 4), 256, 0, st>>>(p);
    } else {
      softmax_bwd_reg<T, 8><<<min(cap, sms * 2), 512, 0, st>>>(p);
    }
  } else if (vec_ok) {
    softmax_bwd_stream<T, true><<<min(cap, sms * 2), 512, 0, st>>>(p);
  } else {
    softmax_bwd_stream<T, false><<<min(cap, sms * 4), 256, 0, st>>>(p);
  }
  return cuda_check(cudaGetLastError(), "softmax_bwd launch");
}

}  // namespace ddp

using namespace ddp;

static int es_of(int dt) { return dt == DDP_F32 ? 4 : 2; }

extern "C" int ddp_softmax_fwd(const ddp_tensor* x, const ddp_tensor* y, const uint8_t* mask, int64_t mask_ld,
                               int64_t mask_sb1, float scale, void* stream) {
  if (x->dtype != y->dtype) return set_error(-4, "softmax: dtype mismatch");
  if (x->rows != y->rows || x->cols != y->cols || x->b0 != y->b0 || x->b1 != y->b1)
    return set_error(-4, "softmax: shape mismatch");
  if (x->rows * x->b1 > INT32_MAX) return set_error(-4, "softmax: too many rows");
  SmParams p = {};
  p.x = x->ptr; p.out = y->ptr; p.mask = mask;
  p.x_ld = x->ld; p.x_sb0 = x->sb0; p.x_sb1 = x->sb1;
  p.o_ld = y->ld; p.o_sb0 = y->sb0; p.o_sb1 = y->sb1;
  p.m_ld = mask_ld; p.m_sb1 = mask_sb1;
  p.rows = (int)x->rows; p.cols = (int)x->cols; p.b0 = (int)x->b0; p.b1 = (int)x->b1;
  p.scale = scale;
  const int es = es_of(x->dtype);
  const int ve = 16 / es;
  bool vec_ok = aligned16(x->ptr, x->ld, x->sb0, x->sb1, es) && aligned16(y->ptr, y->ld, y->sb0, y->sb1, es) &&
                (x->cols % ve == 0);
  if (mask) vec_ok = vec_ok && ((reinterpret_cast<uintptr_t>(mask) % ve) == 0) && (mask_ld % ve == 0) && (mask_sb1 % ve == 0);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (x->dtype) {
    case DDP_F32: return launch_fwd<float>(p, vec_ok, st);
    case DDP_BF16: return launch_fwd<__nv_bfloat16>(p, vec_ok, st);
    default: return launch_fwd<__half>(p, vec_ok, st);
  }
}

extern "C" int ddp_softmax_bwd(const ddp_tensor* y, const ddp_tensor* dy, const ddp_tensor* dx, const float* delta,
                               float scale, void* stream) {
  if (y->dtype != dy->dtype || y->dtype != dx->dtype) return set_error(-4, "softmax_bwd: dtype mismatch");
  if (y->rows != dy->rows || y->cols != dy->cols || y->b0 != dy->b0 || y->b1 != dy->b1 || y->rows != dx->rows ||
      y->cols != dx->cols || y->b0 != dx->b0 || y->b1 != dx->b1)
    return set_error(-4, "softmax_bwd: shape mismatch");
  SmParams p = {};
  p.x = y->ptr; p.dy = dy->ptr; p.out = dx->ptr; p.delta = delta;
  p.x_ld = y->ld; p.x_sb0 = y->sb0; p.x_sb1 = y->sb1;
  p.dy_ld = dy->ld; p.dy_sb0 = dy->sb0; p.dy_sb1 = dy->sb1;
  p.o_ld = dx->ld; p.o_sb0 = dx->sb0; p.o_sb1 = dx->sb1;
  p.rows = (int)y->rows; p.cols = (int)y->cols; p.b0 = (int)y->b0; p.b1 = (int)y->b1;
  p.scale = scale;
  const int es = es_of(y->dtype);
  const bool vec_ok = aligned16(y->ptr, y->ld, y->sb0, y->sb1, es) && aligned16(dy->ptr, dy->ld, dy->sb0, dy->sb1, es) &&
                      aligned16(dx->ptr, dx->ld, dx->sb0, dx->sb1, es) && (y->cols % (16 / es) == 0);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (y->dtype) {
    case DDP_F32: return launch_bwd<float>(p, vec_ok, st);
    case DDP_BF16: return launch_bwd<__nv_bfloat16>(p, vec_ok, st);
    default: return launch_bwd<__half>(p, vec_ok, st);
  }
}

extern "C" int ddp_rowdot(const ddp_tensor* a, const ddp_tensor* b, float* out, void* stream) {
  if (a->dtype != b->dtype || a->rows != b->rows || a->cols != b->cols) return set_error(-4, "rowdot: mismatch");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long rows = a->rows;
  const int wpb = 8;
  const unsigned grid = (unsigned)((rows + wpb - 1) / wpb);
  switch (a->dtype) {
    case DDP_F32:
      rowdot_kernel<float><<<grid, wpb * 32, 0, st>>>((const float*)a->ptr, (const float*)b->ptr, out, rows, (int)a->cols, a->ld, b->ld);
      break;
    case DDP_BF16:
      rowdot_kernel<__nv_bfloat16><<<grid, wpb * 32, 0, st>>>((const __nv_bfloat16*)a->ptr, (const __nv_bfloat16*)b->ptr, out, rows, (int)a->cols, a->ld, b->ld);
      break;
    default:
      rowdot_kernel<__half><<<grid, wpb * 32, 0, st>>>((const __half*)a->ptr, (const __half*)b->ptr, out, rows, (int)a->cols, a->ld, b->ld);
  }
  return cuda_check(cudaGetLastError(), "rowdot launch");
}

// ---------------------------------------------------------------------------------------------------------
// Helpers of the fused softmax-forward path (ddp_gemm_ex with DDP_EPI_SOFTMAX_STATS / _APPLY).
namespace ddp {

// One thread packs 32 mask bytes (two 16-byte loads when aligned) into one word; bits past the row end read
// as "masked".
__global__ void pack_mask_kernel(const uint8_t* __restrict__ mask, long long rows_total, long long rows, long long cols,
                                 long long m_ld, long long m_sb1, uint32_t* __restrict__ bits, long long ld_words) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= rows_total * ld_words) return;
  const long long r = idx / ld_words, w = idx - r * ld_words;
  const long long b1i = r / rows, row = r - b1i * rows;
  const uint8_t* src = mask + b1i * m_sb1 + row * m_ld + w * 32;
  uint32_t word = 0;
  if (w * 32 + 32 <= cols && (reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const uint4 a = *reinterpret_cast<const uint4*>(src);
    const uint4 b = *reinterpret_cast<const uint4*>(src + 16);
    const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if ((v[i] >> (8 * k)) & 0xffu) word |= 1u << (4 * i + k);
  } else {
    for (int k = 0; k < 32; ++k) {
      const long long c = w * 32 + k;
      if (c >= cols || src[k] != 0) word |= 1u << k;
    }
  }
  bits[idx] = word;
}

// One thread per row: fold the per-tile (max, sum) pairs.
__global__ void softmax_combine_kernel(const float2* __restrict__ stats, long long n_rows, int nt,
                                       float2* __restrict__ rowstat) {
  const long long r = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  const float2* s = stats + r * nt;
  float m = -INFINITY;
  for (int t = 0; t < nt; ++t) m = fmaxf(m, s[t].x);
  float l = 0.f;
  for (int t = 0; t < nt; ++t) {
    const float2 v = s[t];
    if (v.x != -INFINITY) l += v.y * exp2f(v.x - m);
  }
  // rowstat.x = max + log2(sum): P = exp2(t - rowstat.x) needs no separate division.  A fully masked row
  // gets NaN -> NaN probabilities, like softmax over a row of -inf in the reference.
  rowstat[r] = (m == -INFINITY) ? make_float2(__int_as_float(0x7fc00000), 0.f) : make_float2(m + log2f(l), 1.f / l);
}

}  // namespace ddp

extern "C" int ddp_pack_mask(const uint8_t* mask, int64_t b1, int64_t rows, int64_t cols, int64_t mask_ld,
                             int64_t mask_sb1, uint32_t* bits, int64_t ld_words, void* stream) {
  const long long total_words = b1 * rows * ld_words;
  const int threads = 256;
  const long long blocks = (total_words + threads - 1) / threads;
  if (blocks > 0x7fffffffLL) return ddp::set_error(-4, "pack_mask: mask too large");
  ddp::pack_mask_kernel<<<static_cast<unsigned>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      mask, b1 * rows, rows, cols, mask_ld, mask_sb1, bits, ld_words);
  return ddp::cuda_check(cudaGetLastError(), "pack_mask launch");
}

extern "C" int ddp_softmax_combine(const void* stats, int64_t n_rows, int32_t nt, void* rowstat, void* stream) {
  const int threads = 128;
  const long long blocks = (n_rows + threads - 1) / threads;
  ddp::softmax_combine_kernel<<<static_cast<unsigned>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const float2*>(stats), n_rows, nt, static_cast<float2*>(rowstat));
  return ddp::cuda_check(cudaGetLastError(), "softmax_combine launch");
}

extern "C" int ddp_debug_timeout_site_rowops(void) {
  unsigned int v = 0, zero = 0;
  if (cudaMemcpyFromSymbol(&v, ddp::g_ddp_timeout_site, sizeof v) != cudaSuccess) return -1;
  if (v) cudaMemcpyToSymbol(ddp::g_ddp_timeout_site, &zero, sizeof zero);
  return static_cast<int>(v);
}
