// Persistent warp-specialised tcgen05 GEMM for sm_100a with the collective fused in.
//
//   warp 0   TMA producer   (one elected lane; waits on per-chunk peer flags before touching gathered rows)
//   warp 1   MMA issuer     (one elected lane; tcgen05.mma, accumulators in TMEM, double buffered)
//   warp 2   TMEM alloc/free, then all-gather pusher
//   warp 3   all-gather pusher
//   warp 4-7 epilogue       (tcgen05.ld -> regs -> smem transpose -> coalesced global / peer stores;
//                            reduce-scatter owner reduction for OP_TN)
//
// Three products share the main loop (reference: multiplication/functions.py:44-212):
//   OP_NT  C[M, seg*N + n] = A[M,K] . Bseg[N,K]^T      A K-major,  B K-major
//   OP_NN  C[M, n]        = sum_seg A[M, seg*K + k] . Bseg[k, n]   A K-major,  B MN-major
//   OP_TN  P_dest[m, n]   = A[k, dest*M + m]^T . B[k, n]           A MN-major, B MN-major
// "seg" walks the ranks in rank-rotated order (own block first for NT/NN, last for TN).
#pragma once
#include "../common/ptx.cuh"
#include "../../include/ddp_b200.h"

namespace ddp {

constexpr int kBlockM = 128;
constexpr int kThreads = 256;
constexpr int kStageABytes = kBlockM * 128;  // 128 rows x 128 B (either majorness)

struct alignas(64) GemmKParams {
  CUtensorMap tmA;   // A operand
  CUtensorMap tmB;   // B operand, local tensor
  CUtensorMap tmBg;  // B operand, gathered symmetric buffer [world][Bt][rows][cols]
  CUtensorMap tmC;   // output, box = 32 rows x 128 bytes (TMA store path of the epilogue)
  // per-segment problem extents
  int M, N, K;
  int b0, b1;  // batch extents; Bt = b0*b1, linear batch b = b1i*b0 + b0i
  int world, rank, nseg;
  int m_blocks, n_blocks, k_blocks;
  int a_seg_c0, a_seg_c1;  // A tensor-map coordinate increments per segment
  // output
  void* C;
  long long ldc, c_sb0, c_sb1;  // element strides
  int c_vec_ok;                 // 16-byte vector stores legal for every tile
  int c_tma_ok;                 // tmC is valid: full-width chunks leave through TMA stores
  float alpha;
  const void* bias;
  // collective
  uint32_t seq;
  int phase_mask;
  int chunk_rows, n_chunks;
  char* ws[DDP_MAX_WORLD];
  uint32_t* flags[DDP_MAX_WORLD];
  uint32_t* done[DDP_MAX_WORLD];
  uint32_t* exit_counter;
  uint32_t* status;  // sticky error word in local memory (0 = ok; 1..5 = which wait timed out)
  // all-gather push source (the local `right` operand), byte strides
  const char* push_src;
  long long push_ld, push_sb0, push_sb1;
  int push_rows, push_row_bytes;
  int dbg;
};

struct TileCoord {
  int seg, b, b0i, b1i, m0, n0;
};

template <int OP>
__device__ __forceinline__ TileCoord decode_tile(const GemmKParams& p, int t) {
  const int mn = p.m_blocks * p.n_blocks;
  const int per_seg = p.b0 * p.b1 * mn;
  const int si = t / per_seg;
  int r = t - si * per_seg;
  TileCoord c;
  c.b = r / mn;
  r -= c.b * mn;
  const int nb = r / p.m_blocks;
  const int mb = r - nb * p.m_blocks;
  c.b1i = c.b / p.b0;
  c.b0i = c.b - c.b1i * p.b0;
  c.m0 = mb * kBlockM;
  if (OP == DDP_OP_NT) {
    c.seg = (p.rank + si) % p.world;
  } else if (OP == DDP_OP_TN) {
    c.seg = (p.rank + 1 + si) % p.world;
  } else {
    c.seg = 0;
  }
  return c;
}

template <int BN>
__device__ __forceinline__ int tile_n0(const GemmKParams& p, int t) {
  const int mn = p.m_blocks * p.n_blocks;
  const int r = t % mn;
  return (r / p.m_blocks) * BN;
}

template <typename OutT>
__device__ __forceinline__ OutT cvt_out(float v);
template <>
__device__ __forceinline__ float cvt_out<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
template <>
__device__ __forceinline__ __half cvt_out<__half>(float v) {
  return __float2half_rn(v);
}

// Epilogue staging tile of one warp: 32 rows x 128 bytes; 16-byte vector j of row r lives at slot
// j ^ (r & 7) - exactly TMA's SWIZZLE_128B, so the same tile can leave through a TMA store or through
// coalesced generic stores.  NF = 128 / sizeof(StoreT) values per row.
template <typename StoreT, int NF>
__device__ __forceinline__ void stage_row(char* sbuf, int row, const float (&v)[NF]) {
  const int sw = row & 7;
  if constexpr (sizeof(StoreT) == 4) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 q = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      *reinterpret_cast<float4*>(sbuf + row * 128 + ((j ^ sw) * 16)) = q;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      StoreT h[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) h[e] = cvt_out<StoreT>(v[8 * j + e]);
      *reinterpret_cast<uint4*>(sbuf + row * 128 + ((j ^ sw) * 16)) = *reinterpret_cast<uint4*>(h);
    }
  }
}

// Generic (predicated) store of the staged tile: dst points at element (row 0, col 0) of the chunk.
template <typename StoreT>
__device__ __forceinline__ void store_rows_generic(const char* sbuf, StoreT* dst, long long ld, int rows_valid,
                                                   int cols_valid, bool vec_ok, int lane) {
  constexpr int VE = 16 / sizeof(StoreT);
  if (vec_ok) {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = it * 4 + (lane >> 3);
      const int j = lane & 7;
      const uint4 q = *reinterpret_cast<const uint4*>(sbuf + r * 128 + ((j ^ (r & 7)) * 16));
      if (r < rows_valid) {
        StoreT* g = dst + static_cast<long long>(r) * ld + j * VE;
        if ((j + 1) * VE <= cols_valid) {
          st_v4(g, q);
        } else {
          const StoreT* e = reinterpret_cast<const StoreT*>(&q);
#pragma unroll
          for (int x = 0; x < VE; ++x)
            if (j * VE + x < cols_valid) g[x] = e[x];
        }
      }
    }
  } else {
    for (int r = 0; r < rows_valid; ++r) {
      for (int c = lane; c < cols_valid; c += 32) {
        const int j = c / VE, x = c % VE;
        dst[static_cast<long long>(r) * ld + c] =
            *reinterpret_cast<const StoreT*>(sbuf + r * 128 + ((j ^ (r & 7)) * 16) + x * sizeof(StoreT));
      }
    }
  }
}

template <int OP, int DT, int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1) gemm_kernel(const __grid_constant__ GemmKParams p) {
  using InT = typename std::conditional<DT == DDP_F32, float,
                                        typename std::conditional<DT == DDP_BF16, __nv_bfloat16, __half>::type>::type;
  constexpr int ELEM = (DT == DDP_F32) ? 4 : 2;
  constexpr int BK = 128 / ELEM;   // elements per 128-byte swizzle row: one k-block
  constexpr int UK = 32 / ELEM;    // K per tcgen05.mma
  constexpr int CH = 128 / ELEM;   // MN elements per 128-byte chunk (MN-major operands)
  constexpr bool A_MN = (OP == DDP_OP_TN);
  constexpr bool B_MN = (OP != DDP_OP_NT);
  constexpr int kStageBBytes = BN * 128;
  constexpr int kStageBytes = kStageABytes + kStageBBytes;
  constexpr uint32_t kTmemCols = 2 * BN;
  constexpr uint32_t kIdesc = umma_idesc(DT, A_MN ? 1u : 0u, B_MN ? 1u : 0u, kBlockM, BN);

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_ab = smem;                                   // STAGES x (A | B)
  char* smem_stage = reinterpret_cast<char*>(smem + STAGES * kStageBytes);  // 4 warps x 2 x 4 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stage + 8 * 4096);
  uint64_t* full_bar = bars;                 // [STAGES]
  uint64_t* empty_bar = bars + STAGES;       // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;   // [2]
  uint64_t* tempty_bar = tfull_bar + 2;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int Bt = p.b0 * p.b1;
  const bool dist = p.world > 1;
  const bool do_produce = (p.phase_mask & 1) != 0;
  const bool do_consume = (p.phase_mask & 2) != 0;

  // The GEMM main loop belongs to the consume phase for the gather ops and to the produce phase for TN.
  const bool run_gemm = (OP == DDP_OP_TN) ? do_produce : do_consume;
  const int tiles_per_seg = Bt * p.m_blocks * p.n_blocks;
  const int n_tile_seg = (OP == DDP_OP_NN) ? 1 : p.nseg;
  const int total_tiles = run_gemm ? tiles_per_seg * n_tile_seg : 0;
  const int k_iters = (OP == DDP_OP_NN) ? p.nseg * p.k_blocks : p.k_blocks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    if (dist && OP != DDP_OP_TN) tma_prefetch_desc(&p.tmBg);
    if (p.c_tma_ok) tma_prefetch_desc(&p.tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ======================================================= TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t* my_flags = p.flags[p.rank];
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile<OP>(p, t);
        const int n0 = tile_n0<BN>(p, t);
        if (OP == DDP_OP_NT && dist && tc.seg != p.rank) {
          // gathered rows [n0, n0+BN) of rank seg must have landed in our buffer
          const int c_lo = n0 / p.chunk_rows;
          const int hi_row = min(n0 + BN, p.N) - 1;
          const int c_hi = hi_row / p.chunk_rows;
          for (int c = c_lo; c <= c_hi; ++c)
            wait_flag_ge(my_flags + (static_cast<size_t>(tc.seg) * Bt + tc.b) * p.n_chunks + c, p.seq, p.status, 1);
          fence_proxy_async();
        }
        int waited_seg = -1, waited_chunk = -1;
        for (int it = 0; it < k_iters; ++it) {
          int seg = tc.seg, kb = it;
          if (OP == DDP_OP_NN) {
            const int si = it / p.k_blocks;
            kb = it - si * p.k_blocks;
            seg = (p.rank + si) % p.world;
          }
          const int k0 = kb * BK;
          if (OP == DDP_OP_NN && dist && seg != p.rank) {
            const int c_lo = k0 / p.chunk_rows;
            const int c_hi = (min(k0 + BK, p.K) - 1) / p.chunk_rows;
            if (seg != waited_seg) {
              waited_seg = seg;
              waited_chunk = -1;
            }
            if (c_hi > waited_chunk) {
              for (int c = max(c_lo, waited_chunk + 1); c <= c_hi; ++c)
                wait_flag_ge(my_flags + (static_cast<size_t>(seg) * Bt + tc.b) * p.n_chunks + c, p.seq, p.status, 2);
              waited_chunk = c_hi;
              fence_proxy_async();
            }
          }
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_ab + stage * kStageBytes;
          uint8_t* sb = sa + kStageABytes;
          mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
          // ---- A
          if constexpr (!A_MN) {
            tma_load_5d(sa, &p.tmA, &full_bar[stage], k0 + seg * p.a_seg_c0, seg * p.a_seg_c1, tc.m0, tc.b0i,
                        tc.b1i);
          } else {
#pragma unroll
            for (int j = 0; j < kBlockM / CH; ++j)
              tma_load_5d(sa + j * (BK * 128), &p.tmA, &full_bar[stage], tc.seg * p.M + tc.m0 + j * CH, 0, k0,
                          tc.b0i, tc.b1i);
          }
          // ---- B
          const bool remote = (OP != DDP_OP_TN) && dist && seg != p.rank;
          const CUtensorMap* mb = remote ? &p.tmBg : &p.tmB;
          const int c3 = remote ? tc.b : tc.b0i;
          const int c4 = remote ? seg : tc.b1i;
          if constexpr (!B_MN) {
            tma_load_5d(sb, mb, &full_bar[stage], k0, 0, n0, c3, c4);
          } else {
#pragma unroll
            for (int j = 0; j < BN / CH; ++j)
              tma_load_5d(sb + j * (BK * 128), mb, &full_bar[stage], n0 + j * CH, 0, k0, c3, c4);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const bool swap = (p.dbg & 1) != 0;
      // MN-major canonical layouts: 16-bit types use 8-row K groups (1024 B apart) under the 16B-chunk
      // swizzle; tf32 must use the 32B-chunk swizzle whose K groups are 4 rows (512 B apart).
      constexpr uint32_t kMnType = (DT == DDP_F32) ? 1u : 2u;
      const uint32_t kMnGroup = (DT == DDP_F32 && !(p.dbg & 2)) ? 512u : 1024u;
      const uint32_t mn_lbo = swap ? kMnGroup : static_cast<uint32_t>(BK * 128);
      const uint32_t mn_sbo = swap ? static_cast<uint32_t>(BK * 128) : kMnGroup;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int it = 0; it < k_iters; ++it) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_ab + stage * kStageBytes);
          const uint32_t sb = sa + kStageABytes;
#pragma unroll
          for (int kk = 0; kk < BK / UK; ++kk) {
            const uint64_t adesc = A_MN ? umma_smem_desc(sa + kk * UK * 128, mn_lbo, mn_sbo, kMnType)
                                        : umma_smem_desc(sa + kk * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? umma_smem_desc(sb + kk * UK * 128, mn_lbo, mn_sbo, kMnType)
                                        : umma_smem_desc(sb + kk * 32, 16, 1024);
            umma_ss<DT == DDP_F32>(d_tmem, adesc, bdesc, kIdesc, (it | kk) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (it == k_iters - 1) umma_commit(&tfull_bar[acc]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp < 4) {
    // ======================================================= all-gather pushers (NT / NN)
    if (OP != DDP_OP_TN && dist && do_produce) {
      const int gw = blockIdx.x * 2 + (warp - 2);
      const int gstride = gridDim.x * 2;
      const int items = Bt * p.n_chunks;
      const int vpr = p.push_row_bytes / 16;
      if (gw < items) {
        // The peers must have finished call seq-2 before we overwrite this parity of their workspace.
        if (lane > 0 && lane < p.world) {
          const int q = (p.rank + lane) % p.world;
          wait_flag_ge(p.done[p.rank] + q, p.seq - 2u, p.status, 3);
        }
        __syncwarp();
      }
      for (int item = gw; item < items; item += gstride) {
        const int b = item / p.n_chunks;
        const int c = item - b * p.n_chunks;
        const int b1i = b / p.b0, b0i = b - b1i * p.b0;
        const int r0 = c * p.chunk_rows;
        const int nrows = min(p.chunk_rows, p.push_rows - r0);
        const char* src = p.push_src + b1i * p.push_sb1 + b0i * p.push_sb0 + static_cast<long long>(r0) * p.push_ld;
        const size_t dst_off =
            ((static_cast<size_t>(p.rank) * Bt + b) * p.push_rows + r0) * static_cast<size_t>(p.push_row_bytes);
        const int nvec = nrows * vpr;
        for (int v0 = 0; v0 < nvec; v0 += 32 * 4) {
          uint4 q[4];
          int rr[4], jj[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int v = v0 + u * 32 + lane;
            rr[u] = v / vpr;
            jj[u] = v - rr[u] * vpr;
            if (v < nvec) q[u] = ld_nc_v4(src + static_cast<long long>(rr[u]) * p.push_ld + jj[u] * 16);
          }
          for (int i = 1; i < p.world; ++i) {
            const int peer = (p.rank - i + p.world) % p.world;  // rank-1 consumes our rows first
            char* dst = p.ws[peer] + dst_off;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int v = v0 + u * 32 + lane;
              if (v < nvec) st_v4(dst + static_cast<size_t>(rr[u]) * p.push_row_bytes + jj[u] * 16, q[u]);
            }
          }
        }
        __threadfence_system();
        __syncwarp();
        if (lane > 0 && lane < p.world) {
          const int peer = (p.rank - lane + p.world) % p.world;
          st_release_sys(p.flags[peer] + (static_cast<size_t>(p.rank) * Bt + b) * p.n_chunks + c, p.seq);
        }
        __syncwarp();
      }
    }
  } else {
    // ======================================================= epilogue warps
    using OutT = InT;
    const int q = warp - 4;  // TMEM lane quadrant
    char* sbuf0 = smem_stage + q * 8192;  // two 4 KB staging buffers per warp
    int sb = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t done_ok = 0;  // bitmask of peers whose done counter we have already checked (TN)
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const TileCoord tc = decode_tile<OP>(p, t);
      const int n0 = tile_n0<BN>(p, t);
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const int row0 = tc.m0 + q * 32;                       // first row of this warp's fragment
      const int rows_valid = max(0, min(32, p.M - row0));
      const bool tn_remote = (OP == DDP_OP_TN) && dist;
      if (tn_remote && tc.seg != p.rank && !(done_ok & (1u << tc.seg))) {
        if (lane == 0) wait_flag_ge(p.done[p.rank] + tc.seg, p.seq - 2u, p.status, 4);
        __syncwarp();
        done_ok |= 1u << tc.seg;
      }
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      // Drains the accumulator in chunks of 128 output bytes per row (64 bf16/fp16 or 32 fp32 columns).
      auto drain = [&](auto tag, bool to_peer) {
        using StoreT = decltype(tag);
        constexpr int CC = 128 / sizeof(StoreT);
        constexpr int NLD = CC / 32;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += CC) {
          const int col0 = n0 + c0;
          const int cols_valid = min(CC, p.N - col0);
          if (cols_valid <= 0 || rows_valid == 0) break;  // warp-uniform
          uint32_t r[NLD][32];
#pragma unroll
          for (int i = 0; i < NLD; ++i) tmem_ld_32x32(taddr + c0 + i * 32, r[i]);
          tmem_ld_wait();
          float v[CC];
#pragma unroll
          for (int i = 0; i < NLD; ++i)
#pragma unroll
            for (int e = 0; e < 32; ++e) v[i * 32 + e] = __uint_as_float(r[i][e]) * p.alpha;
          if (!to_peer && p.bias != nullptr) {
            const StoreT* bias = reinterpret_cast<const StoreT*>(p.bias);
#pragma unroll
            for (int e = 0; e < CC; ++e)
              if (e < cols_valid) v[e] += static_cast<float>(bias[col0 + e]);
          }
          char* sbuf = sbuf0 + sb * 4096;
          if (lane == 0) bulk_wait_group_read<1>();  // the TMA store that last read this buffer is done
          __syncwarp();
          stage_row<StoreT, CC>(sbuf, lane, v);
          if (to_peer) {
            // fp32 partial tile -> owner's slot [src = rank][b][M][N]
            __syncwarp();
            StoreT* dst = reinterpret_cast<StoreT*>(p.ws[tc.seg]) +
                          ((static_cast<size_t>(p.rank) * Bt + tc.b) * p.M + row0) * p.N + col0;
            store_rows_generic<StoreT>(sbuf, dst, p.N, rows_valid, cols_valid, (p.N & 3) == 0, lane);
            __syncwarp();
          } else {
            const long long col_g = (OP == DDP_OP_NT ? static_cast<long long>(tc.seg) * p.N : 0) + col0;
            if (p.c_tma_ok && cols_valid == CC) {
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_5d(&p.tmC, sbuf, static_cast<int>(col_g), 0, row0, tc.b0i, tc.b1i);
                bulk_commit_group();
              }
            } else {
              __syncwarp();
              StoreT* dst = reinterpret_cast<StoreT*>(p.C) + tc.b1i * p.c_sb1 + tc.b0i * p.c_sb0 +
                            static_cast<long long>(row0) * p.ldc + col_g;
              store_rows_generic<StoreT>(sbuf, dst, p.ldc, rows_valid, cols_valid, p.c_vec_ok != 0, lane);
              __syncwarp();
            }
          }
          sb ^= 1;
        }
      };
      if (tn_remote) {
        drain(float{}, true);
      } else {
        drain(OutT{}, false);
      }
      // accumulator drained -> hand the TMEM buffer back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (tn_remote) {
        // publish this warp's 32 rows of the partial tile to the owner
        __threadfence_system();
        __syncwarp();
        if (lane == 0) {
          const int mb = tc.m0 / kBlockM, nb = n0 / BN;
          const size_t fidx =
              (((static_cast<size_t>(p.rank) * Bt + tc.b) * p.m_blocks + mb) * p.n_blocks + nb) * 4 + q;
          st_release_sys(p.flags[tc.seg] + fidx, p.seq);
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (lane == 0) bulk_wait_group_read<0>();  // staging smem must outlive the TMA stores reading it
    // ---------------------------------------------------- reduce-scatter: owner-side reduction
    if (OP == DDP_OP_TN && dist && do_consume) {
      const int items = Bt * p.m_blocks * p.n_blocks * 4;
      const uint32_t* my_flags = p.flags[p.rank];
      for (int item = blockIdx.x * 4 + q; item < items; item += gridDim.x * 4) {
        const int quad = item & 3;
        int r = item >> 2;
        const int nb = r % p.n_blocks;
        r /= p.n_blocks;
        const int mb = r % p.m_blocks;
        const int b = r / p.m_blocks;
        if (lane < p.world) {
          const size_t fidx = (((static_cast<size_t>(lane) * Bt + b) * p.m_blocks + mb) * p.n_blocks + nb) * 4 + quad;
          wait_flag_ge(my_flags + fidx, p.seq, p.status, 5);
        }
        __syncwarp();
        const int b1i = b / p.b0, b0i = b - b1i * p.b0;
        const int row0 = mb * kBlockM + quad * 32;
        const int rows_valid = max(0, min(32, p.M - row0));
        const int col0 = nb * BN;
        const int cols_valid = max(0, min(BN, p.N - col0));
        const float* part = reinterpret_cast<const float*>(p.ws[p.rank]);
        OutT* out = reinterpret_cast<OutT*>(p.C) + b1i * p.c_sb1 + b0i * p.c_sb0;
        for (int rr = 0; rr < rows_valid; ++rr) {
          for (int cc = lane; cc < cols_valid; cc += 32) {
            float s = 0.f;
            for (int src = 0; src < p.world; ++src)
              s += part[((static_cast<size_t>(src) * Bt + b) * p.M + row0 + rr) * p.N + col0 + cc];
            out[static_cast<long long>(row0 + rr) * p.ldc + col0 + cc] = cvt_out<OutT>(s);
          }
        }
      }
    }
  }

  // ---------------------------------------------------------- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
  if (dist && do_consume && threadIdx.x == 0) {
    // Last CTA out tells every peer that this rank no longer reads workspace parity (seq & 1).
    __threadfence();
    const uint32_t prev = atomicAdd(p.exit_counter, 1u);
    if (prev == gridDim.x - 1) {
      *p.exit_counter = 0;
      __threadfence_system();
      for (int i = 0; i < p.world; ++i) st_release_sys(p.done[i] + p.rank, p.seq);
    }
  }
}

}  // namespace ddp
