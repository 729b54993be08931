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
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <type_traits>
#include "../common/ptx.cuh"
#include "../../include/ddp_b200.h"

namespace ddp {

constexpr int kBlockM = 128;
constexpr int kThreadsBase = 128;  // producer, MMA, 2 pusher warps; plus 32 x EW epilogue threads
constexpr int kStageABytes = kBlockM * 128;  // 128 rows x 128 B (either majorness)

struct alignas(64) GemmKParams {
  CUtensorMap tmA;   // A operand
  CUtensorMap tmB;   // B operand, local tensor
  CUtensorMap tmBg;  // B operand, gathered symmetric buffer [world][Bt][rows][cols]
  CUtensorMap tmC;   // output, box = 32 rows x 128 bytes (TMA store path of the epilogue)
  CUtensorMap tmP;   // EPI == 1: probabilities P, same shape / box as C
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
  // EPI == 1 (softmax backward fused into the dP = dO V^T epilogue): C = epi_scale * P * (acc - delta[row])
  const float* delta;  // [Bt][M] fp32, rowsum(dO * O)
  float epi_scale;
  // EPI == 2 / 3 (softmax forward fused into the K Q^T product, two passes over the tiles, no score slab):
  //   2: per (row, column tile) online-softmax statistics (max, sum) of t = acc * epi_scale (log2 domain)
  //   3: C = exp2(t - rowstat.x) * rowstat.y   with rowstat = (row max, 1 / row sum) from the combine kernel
  const uint32_t* mask_bits;   // bit-packed mask [b1][M][mb_ld words], bit set = masked out; may be null
  long long mb_ld, mb_sb1;     // in 32-bit words
  float2* stats;               // EPI 2 out: [Bt][M][stats_nt]
  const float2* rowstat;       // EPI 3 in:  [Bt][M]
  int stats_nt;
  // collective
  uint32_t seq;
  int phase_mask;
  int chunk_rows, n_chunks;
  char* ws[DDP_MAX_WORLD];
  char* mc_ws;  // multicast alias of the workspace (null: unicast pushes)
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

// Tile order: segments outermost (rank-rotated).  kHeadFastest (fused softmax epilogues) makes the inner
// batch index b0 (heads) vary fastest so the 16 tiles that share one mask tile run back to back (L2 reuse).
template <int OP, int BN, bool kHeadFastest = false>
__device__ __forceinline__ TileCoord decode_tile(const GemmKParams& p, int t) {
  const int mn = p.m_blocks * p.n_blocks;
  const int per_seg = p.b0 * p.b1 * mn;
  const int si = t / per_seg;
  int r = t - si * per_seg;
  TileCoord c;
  int nb, mb;
  if (kHeadFastest) {
    const int per_b1 = p.b0 * mn;
    c.b1i = r / per_b1;
    r -= c.b1i * per_b1;
    const int per_nb = p.m_blocks * p.b0;
    nb = r / per_nb;
    r -= nb * per_nb;
    mb = r / p.b0;
    c.b0i = r - mb * p.b0;
    c.b = c.b1i * p.b0 + c.b0i;
  } else {
    c.b = r / mn;
    r -= c.b * mn;
    nb = r / p.m_blocks;
    mb = r - nb * p.m_blocks;
    c.b1i = c.b / p.b0;
    c.b0i = c.b - c.b1i * p.b0;
  }
  c.m0 = mb * kBlockM;
  c.n0 = nb * BN;
  if (OP == DDP_OP_NT) {
    c.seg = (p.rank + si) % p.world;
  } else if (OP == DDP_OP_TN) {
    c.seg = (p.rank + 1 + si) % p.world;
  } else {
    c.seg = 0;
  }
  return c;
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

// Two fp32 values -> one packed 32-bit register of 16-bit floats (single F2FP, no stack traffic).
template <typename T>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  if constexpr (std::is_same<T, __half>::value) {
    const __half2 t = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&t);
  } else {
    const __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&t);
  }
}

// Epilogue staging tile of one warp: 32 rows x 128 bytes; 16-byte vector j of row r lives at slot
// j ^ (r & 7) - exactly TMA's SWIZZLE_128B, so the same tile can leave through a TMA store or through
// coalesced generic stores.  NF = 128 / sizeof(StoreT) values per row.
template <typename StoreT>
__device__ __forceinline__ void stage_row32(uint32_t sbuf, int row, const float (&v)[32], int part) {
  const int sw = row & 7;
  if constexpr (sizeof(StoreT) == 4) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      uint4 q;
      q.x = __float_as_uint(v[4 * j]); q.y = __float_as_uint(v[4 * j + 1]);
      q.z = __float_as_uint(v[4 * j + 2]); q.w = __float_as_uint(v[4 * j + 3]);
      st_shared_v4(sbuf + row * 128 + ((j ^ sw) * 16), q);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint4 q;
      q.x = pack2<StoreT>(v[8 * j + 0], v[8 * j + 1]);
      q.y = pack2<StoreT>(v[8 * j + 2], v[8 * j + 3]);
      q.z = pack2<StoreT>(v[8 * j + 4], v[8 * j + 5]);
      q.w = pack2<StoreT>(v[8 * j + 6], v[8 * j + 7]);
      st_shared_v4(sbuf + row * 128 + (((part * 4 + j) ^ sw) * 16), q);
    }
  }
}

// Generic (predicated) store of the staged tile: dst points at element (row 0, col 0) of the chunk.
template <typename StoreT>
__device__ __forceinline__ void store_rows_generic(uint32_t sbuf, StoreT* dst, long long ld, int rows_valid,
                                                   int cols_valid, bool vec_ok, int lane) {
  constexpr int VE = 16 / sizeof(StoreT);
  if (vec_ok) {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = it * 4 + (lane >> 3);
      const int j = lane & 7;
      const uint4 q = ld_shared_v4(sbuf + r * 128 + ((j ^ (r & 7)) * 16));
      if (r < rows_valid) {
        StoreT* g = dst + static_cast<long long>(r) * ld + j * VE;
        if ((j + 1) * VE <= cols_valid) {
          st_v4(g, q);
        } else {
          const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
          for (int x = 0; x < VE; ++x) {
            if (j * VE + x < cols_valid) {
              if constexpr (sizeof(StoreT) == 2) {
                const unsigned short hbits = static_cast<unsigned short>(w[x / 2] >> (16 * (x & 1)));
                g[x] = *reinterpret_cast<const StoreT*>(&hbits);
              } else {
                g[x] = *reinterpret_cast<const StoreT*>(&w[x]);
              }
            }
          }
        }
      }
    }
  } else {
    for (int r = 0; r < rows_valid; ++r) {
      for (int c = lane; c < cols_valid; c += 32) {
        const int j = c / VE, x = c % VE;
        const uint4 q = ld_shared_v4(sbuf + r * 128 + ((j ^ (r & 7)) * 16));
        // select the element with shifts instead of indexing the register vector
        constexpr int PER_WORD = 4 / sizeof(StoreT);
        const int wi = x / PER_WORD;
        uint32_t word = (wi & 2) ? ((wi & 1) ? q.w : q.z) : ((wi & 1) ? q.y : q.x);
        if constexpr (sizeof(StoreT) == 2) {
          const unsigned short hbits = static_cast<unsigned short>(word >> (16 * (x & 1)));
          dst[static_cast<long long>(r) * ld + c] = *reinterpret_cast<const StoreT*>(&hbits);
        } else {
          dst[static_cast<long long>(r) * ld + c] = *reinterpret_cast<const StoreT*>(&word);
        }
      }
    }
  }
}

template <int OP, int DT, int BN, int STAGES, int EPI = 0, int EW = 4>
__global__ void __launch_bounds__(kThreadsBase + 32 * EW, 1) gemm_kernel(const __grid_constant__ GemmKParams p) {
  static_assert(EW == 4 || EW == 8, "4 or 8 epilogue warps");
  static_assert(EW == 4 || OP == DDP_OP_NT, "8 epilogue warps: nt only");
  static_assert(EW == 4 || BN >= 128, "8 epilogue warps need two 128-byte chunks per row and tile");
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
  char* smem_stage = reinterpret_cast<char*>(smem + STAGES * kStageBytes);  // EW warps x 2 x 4 KB
  // EPI == 1: instead of staging buffers every warp owns a ring of kPBufs 4 KB buffers: P lands there by TMA
  // (two chunks ahead), is turned into dS in place, and leaves from the same buffer by TMA store.
  constexpr int kPBufs = 4;
  char* smem_pst = smem_stage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stage + (EPI == 1 ? EW * kPBufs : 2 * EW) * 4096);
  uint64_t* full_bar = bars;                 // [STAGES]
  uint64_t* empty_bar = bars + STAGES;       // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;   // [2]
  uint64_t* tempty_bar = tfull_bar + 2;      // [2]
  uint64_t* p_bar = tempty_bar + 2;          // [EW warps][kPBufs]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_bar + 8 * kPBufs);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler, too
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
    if (EPI == 1) tma_prefetch_desc(&p.tmP);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], EW);
    }
    for (int s = 0; s < 8 * kPBufs; ++s) mbar_init(&p_bar[s], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ======================================================= TMA producer
    // The whole warp walks the tile / k-block sequence (uniform control flow): all 32 lanes poll the peer
    // flags a tile or rank segment depends on in parallel (one round trip instead of one per flag), lane 0
    // alone talks to the mbarriers and issues the TMA loads.
    {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t* my_flags = p.flags[p.rank];
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile<OP, BN, (EPI >= 2)>(p, t);
        const int n0 = tc.n0;
        bool need_fence = false;
        if (OP == DDP_OP_NT && dist && tc.seg != p.rank) {
          // gathered rows [n0, n0+BN) of rank seg must have landed in our buffer
          const int c_lo = n0 / p.chunk_rows;
          const int c_hi = (min(n0 + BN, p.N) - 1) / p.chunk_rows;
          const uint32_t* f = my_flags + (static_cast<size_t>(tc.seg) * Bt + tc.b) * p.n_chunks;
          for (int c = c_lo + lane; c <= c_hi; c += 32) wait_flag_ge(f + c, p.seq, p.status, 1);
          __syncwarp();
          need_fence = true;
        }
        int polled_seg = -1;
        for (int it = 0; it < k_iters; ++it) {
          int seg = tc.seg, kb = it;
          if (OP == DDP_OP_NN) {
            const int si = it / p.k_blocks;
            kb = it - si * p.k_blocks;
            seg = (p.rank + si) % p.world;
          }
          const int k0 = kb * BK;
          if (OP == DDP_OP_NN && dist && seg != p.rank && seg != polled_seg) {
            // first k-block of a remote rank segment: all of its chunks for this batch, lane-parallel
            const uint32_t* f = my_flags + (static_cast<size_t>(seg) * Bt + tc.b) * p.n_chunks;
            for (int c = lane; c < p.n_chunks; c += 32) wait_flag_ge(f + c, p.seq, p.status, 2);
            __syncwarp();
            polled_seg = seg;
            need_fence = true;
          }
          if (lane == 0) {
            if (need_fence) fence_proxy_async();  // peer (generic-proxy) writes -> our TMA (async-proxy) reads
            mbar_wait(&empty_bar[stage], phase ^ 1, 1);
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
          }
          need_fence = false;
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer
    // The whole warp walks the tile loop (waits included) and one elected lane issues: warp-uniform operands keep
    // the MMA descriptors in uniform registers.  Issued from inside `if (lane == 0)` every tcgen05.mma came out
    // wrapped in an elect / R2UR / branch loop of ~14 dependent instructions (~100 clk per MMA: hidden behind
    // N = 256 MMAs, but not behind the N = 64 / 128 ones of narrow problems).
    {
      const bool leader = elect_one();
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
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1, 2);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int it = 0; it < k_iters; ++it) {
          mbar_wait(&full_bar[stage], phase, 3);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_ab + stage * kStageBytes);
          const uint32_t sb = sa + kStageABytes;
          if (leader) {
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
          }
          __syncwarp();
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
          if (p.mc_ws != nullptr) {
            // one multicast store: the switch writes the vector into every rank's gather buffer
            char* dst = p.mc_ws + dst_off;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int v = v0 + u * 32 + lane;
              if (v < nvec) multimem_st_v4(dst + static_cast<size_t>(rr[u]) * p.push_row_bytes + jj[u] * 16, q[u]);
            }
          } else
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
    const int ew = warp - 4;
    const int q = ew & 3;    // TMEM lane quadrant this warp may read (= warp id mod 4)
    const int ch = ew >> 2;  // with 8 epilogue warps, two warps share a quadrant and split the columns
    constexpr int CW = BN / (EW / 4);  // columns of a tile handled by one warp
    char* sbuf0 = smem_stage + ew * 8192;  // two 4 KB staging buffers per warp
    int sb = 0;
    // EPI == 1: P chunks arrive through TMA kPBufs-1 chunks ahead of their use (enough bytes in flight to
    // cover DRAM latency), in the running order (tile, chunk) of this warp; g counts consumed chunks,
    // buffer/barrier = g % kPBufs, parity = (g / kPBufs) & 1.
    char* pbuf0 = smem_pst + ew * (kPBufs * 4096);
    int g = 0;
    auto issue_p = [&](int gi) {  // lane 0 only
      constexpr int CCp = 128 / sizeof(OutT);
      constexpr int CPT = CW / CCp;
      const int ti = gi / CPT, ci = gi - ti * CPT;
      const int tt = blockIdx.x + ti * gridDim.x;
      if (tt >= total_tiles) return;
      const TileCoord c2 = decode_tile<OP, BN, (EPI >= 2)>(p, tt);
      const int col = c2.seg * p.N + c2.n0 + ch * CW + ci * CCp;
      uint64_t* bar = &p_bar[ew * kPBufs + (gi % kPBufs)];
      mbar_arrive_expect_tx(bar, 4096);
      tma_load_5d(pbuf0 + (gi % kPBufs) * 4096, &p.tmP, bar, col, 0, c2.m0 + q * 32, c2.b0i, c2.b1i);
    };
    if (EPI == 1 && lane == 0) {
      issue_p(0);
      issue_p(1);
    }
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t done_ok = 0;  // bitmask of peers whose done counter we have already checked (TN)
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const TileCoord tc = decode_tile<OP, BN, (EPI >= 2)>(p, t);
      const int n0 = tc.n0;
      const int row0 = tc.m0 + q * 32;                       // first row of this warp's fragment
      const int rows_valid = max(0, min(32, p.M - row0));
      // Per-tile operands of the fused epilogues are fetched BEFORE waiting for the accumulator, so their
      // latency hides behind the MMA: delta (EPI 1); mask words of this thread's row and row stats (EPI 2/3).
      float row_delta = 0.f;
      if (EPI == 1 && row0 + lane < p.M) row_delta = p.delta[static_cast<size_t>(tc.b) * p.M + row0 + lane];
      uint32_t mk0 = 0, mk1 = 0, mk2 = 0, mk3 = 0, mk4 = 0, mk5 = 0, mk6 = 0, mk7 = 0;
      float2 rstat = make_float2(0.f, 0.f);
      if constexpr (EPI >= 2) {
        const bool row_ok = row0 + lane < p.M;
        if (p.mask_bits != nullptr && row_ok) {
          const long long col_g0 = static_cast<long long>(tc.seg) * p.N + n0;
          const uint32_t* mrow =
              p.mask_bits + tc.b1i * p.mb_sb1 + static_cast<long long>(row0 + lane) * p.mb_ld + (col_g0 >> 5);
          const int nw = min(BN / 32, (p.N - n0 + 31) >> 5);  // words that exist for this tile
          if (0 < nw) mk0 = __ldg(mrow + 0);
          if (1 < nw) mk1 = __ldg(mrow + 1);
          if (BN > 64) {
            if (2 < nw) mk2 = __ldg(mrow + 2);
            if (3 < nw) mk3 = __ldg(mrow + 3);
          }
          if (BN > 128) {
            if (4 < nw) mk4 = __ldg(mrow + 4);
            if (5 < nw) mk5 = __ldg(mrow + 5);
            if (6 < nw) mk6 = __ldg(mrow + 6);
            if (7 < nw) mk7 = __ldg(mrow + 7);
          }
        }
        if (EPI == 3 && row_ok) rstat = p.rowstat[static_cast<size_t>(tc.b) * p.M + row0 + lane];
      }
      mbar_wait(&tfull_bar[acc], acc_phase, 4);
      tc_fence_after();
      const bool tn_remote = (OP == DDP_OP_TN) && dist;
      if (tn_remote && tc.seg != p.rank && !(done_ok & (1u << tc.seg))) {
        if (lane == 0) wait_flag_ge(p.done[p.rank] + tc.seg, p.seq - 2u, p.status, 4);
        __syncwarp();
        done_ok |= 1u << tc.seg;
      }
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      if constexpr (EPI == 2) {
        // ---- statistics pass: nothing is stored but (max, sum) of this row over the tile's columns
        float m_run = -INFINITY, l_run = 0.f;
#pragma unroll 1
        for (int c0 = ch * CW; c0 < (ch + 1) * CW; c0 += 32) {
          if (n0 + c0 >= p.N) break;
          uint32_t r[32];
          tmem_ld_32x32(taddr + c0, r);
          tmem_ld_wait();
          tmem_ld_pin(r);
          const uint32_t mb = pick8(mk0, mk1, mk2, mk3, mk4, mk5, mk6, mk7, c0 >> 5);
          const int nvalid = min(32, p.N - (n0 + c0));
          float tv[32];
          float cmax = -INFINITY;
          if (nvalid == 32 && !__any_sync(0xffffffffu, mb != 0u)) {
            // common case (no mask bit anywhere in these 32 x 32 scores): max of the raw accumulators
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              tv[e] = __uint_as_float(r[e]);
              cmax = fmaxf(cmax, tv[e]);
            }
            cmax *= p.epi_scale;  // epi_scale > 0
            const float m_new = fmaxf(m_run, cmax);
            float acc_l = 0.f;
#pragma unroll
            for (int e = 0; e < 32; ++e) acc_l += ex2_approx(fmaf(tv[e], p.epi_scale, -m_new));
            l_run = l_run * ex2_approx(m_run - m_new) + acc_l;
            m_run = m_new;
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const bool dead = ((mb >> e) & 1u) || e >= nvalid;
              tv[e] = dead ? -INFINITY : __uint_as_float(r[e]) * p.epi_scale;
              cmax = fmaxf(cmax, tv[e]);
            }
            const float m_new = fmaxf(m_run, cmax);
            if (m_new != -INFINITY) {
              float acc_l = 0.f;
#pragma unroll
              for (int e = 0; e < 32; ++e) acc_l += ex2_approx(tv[e] - m_new);
              l_run = l_run * ex2_approx(m_run - m_new) + acc_l;
              m_run = m_new;
            }
          }
        }
        if (row0 + lane < p.M) {
          const int ti = (tc.seg * p.n_blocks + n0 / BN) * (EW / 4) + ch;
          p.stats[(static_cast<size_t>(tc.b) * p.M + row0 + lane) * p.stats_nt + ti] = make_float2(m_run, l_run);
        }
      }
      // Drains the accumulator in chunks of 128 output bytes per row (64 bf16/fp16 or 32 fp32 columns).
      auto drain = [&](auto tag, bool to_peer) {
        using StoreT = decltype(tag);
        constexpr int CC = 128 / sizeof(StoreT);
        constexpr int NLD = CC / 32;
#pragma unroll 1
        for (int c0 = ch * CW; c0 < (ch + 1) * CW; c0 += CC) {
          const int col0 = n0 + c0;
          const int cols_valid = min(CC, p.N - col0);
          const bool active = cols_valid > 0 && rows_valid > 0;  // warp-uniform
          const uint32_t pbuf = smem_u32(pbuf0) + (g % kPBufs) * 4096;  // EPI 1: landing = staging buffer
          if constexpr (EPI == 1) {
            if (!to_peer) {
              if (lane == 0) {
                // buffer (g+2) % kPBufs was last read by the TMA store of chunk g-2: all but the newest
                // store group must have finished reading before P for chunk g+2 may land there
                bulk_wait_group_read<1>();
                issue_p(g + 2);
              }
              mbar_wait(&p_bar[ew * kPBufs + (g % kPBufs)], (g / kPBufs) & 1, 5);  // every chunk is loaded, used or not
              ++g;
            }
          }
          if (!active) {
            if (EPI == 1) continue;
            break;
          }
          // TMEM -> registers.  With two epilogue warps per lane quadrant (EW == 8) only ONE 32-column load
          // per warp is kept in flight: more than two outstanding 32x32b.x32 loads on a quadrant returned
          // zeros for the oldest one on B200 (observed, intermittent), so those loads are issued per part.
          uint32_t r[NLD][32];
          if constexpr (EW == 4) {
            if constexpr (NLD == 2) {
              tmem_ld_32x64(taddr + c0, reinterpret_cast<uint32_t(&)[64]>(r));
            } else {
              tmem_ld_32x32(taddr + c0, r[0]);
            }
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < NLD; ++i) tmem_ld_pin(r[i]);
          }
          if (p.dbg & 4) continue;  // harness only: time the TMEM drain alone
          // shared-space address of this chunk's staging buffer (EPI 1: the P landing buffer itself)
          const uint32_t sbuf = (EPI == 1 && !to_peer) ? pbuf : smem_u32(sbuf0) + sb * 4096;
          if (!(EPI == 1 && !to_peer)) {
            if (lane == 0) bulk_wait_group_read<1>();  // the TMA store that last read this buffer is done
          }
          __syncwarp();
          // 32 columns at a time: registers -> epilogue math -> this row's slice of the staging tile
#pragma unroll
          for (int i = 0; i < NLD; ++i) {
            if constexpr (EW == 8) {
              tmem_ld_32x32(taddr + c0 + i * 32, r[i]);
              tmem_ld_wait();
              tmem_ld_pin(r[i]);
            }
            float v[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(r[i][e]) * p.alpha;
            if (EPI == 3 && !(p.dbg & 32)) {  // dbg 32 (harness): store the raw scores instead
              // P = exp2(scale*log2e*s - (row max + log2 row sum)); masked entries are exactly 0 and a fully
              // masked row (rstat.x = NaN) is NaN, as softmax over a row of -inf is in the reference
              const uint32_t cw = pick8(mk0, mk1, mk2, mk3, mk4, mk5, mk6, mk7, (c0 >> 5) + i);
              const float neg_x = -rstat.x;
              if (!__any_sync(0xffffffffu, cw != 0u)) {
#pragma unroll
                for (int e = 0; e < 32; ++e) v[e] = ex2_approx(fmaf(v[e], p.epi_scale, neg_x));
              } else {
                const float dead_t = (rstat.x == rstat.x) ? -INFINITY : rstat.x;
#pragma unroll
                for (int e = 0; e < 32; ++e)
                  v[e] = ex2_approx(((cw >> e) & 1u) ? dead_t : fmaf(v[e], p.epi_scale, neg_x));
              }
            }
            if constexpr (EPI == 1 && sizeof(StoreT) == 2) {
              if (!to_peer) {
                // dS = P * (scale*dP - scale*delta) with P kept packed: one fp32 FMA per element, then the
                // product in packed 16-bit math straight into the staging tile (P never gets unpacked).
                using P2 = typename std::conditional<std::is_same<StoreT, __half>::value, __half2, __nv_bfloat162>::type;
                const float nsd = -p.epi_scale * row_delta;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const int jj = i * 4 + j;
                  const uint32_t slot = lane * 128 + ((jj ^ (lane & 7)) * 16);
                  const uint4 pv = ld_shared_v4(pbuf + slot);
                  const P2* pp = reinterpret_cast<const P2*>(&pv);
                  uint4 ov;
                  P2* op = reinterpret_cast<P2*>(&ov);
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    P2 t;
                    if constexpr (std::is_same<StoreT, __half>::value) {
                      t = __floats2half2_rn(fmaf(v[j * 8 + 2 * e], p.epi_scale, nsd),
                                            fmaf(v[j * 8 + 2 * e + 1], p.epi_scale, nsd));
                    } else {
                      t = __floats2bfloat162_rn(fmaf(v[j * 8 + 2 * e], p.epi_scale, nsd),
                                                fmaf(v[j * 8 + 2 * e + 1], p.epi_scale, nsd));
                    }
                    op[e] = __hmul2(t, pp[e]);
                  }
                  st_shared_v4(sbuf + slot, ov);
                }
                continue;  // this part is staged
              }
            }
            if constexpr (EPI == 1 && sizeof(StoreT) == 4) {
              if (!to_peer) {
                // dS = scale * P * (dP - delta): P row of this thread from the swizzled landing buffer
                constexpr int VE = 16 / sizeof(StoreT);
                constexpr int VPP = 32 / VE;  // 16-byte vectors per 32 columns
#pragma unroll
                for (int j = 0; j < VPP; ++j) {
                  const int jj = i * VPP + j;
                  const uint4 pv = ld_shared_v4(pbuf + lane * 128 + ((jj ^ (lane & 7)) * 16));
                  const StoreT* pe = reinterpret_cast<const StoreT*>(&pv);
#pragma unroll
                  for (int e = 0; e < VE; ++e)
                    v[j * VE + e] = p.epi_scale * static_cast<float>(pe[e]) * (v[j * VE + e] - row_delta);
                }
              }
            }
            if (!to_peer && p.bias != nullptr) {
              const StoreT* bias = reinterpret_cast<const StoreT*>(p.bias);
#pragma unroll
              for (int e = 0; e < 32; ++e)
                if (i * 32 + e < cols_valid) v[e] += static_cast<float>(bias[col0 + i * 32 + e]);
            }
            stage_row32<StoreT>(sbuf, lane, v, i);
          }
          if (p.dbg & 8) {  // harness only: no store
            __syncwarp();
            sb ^= 1;
            continue;
          }
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
      if (EPI == 2) {
        // statistics only
      } else if (tn_remote) {
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
        const size_t src_stride = static_cast<size_t>(Bt) * p.M * p.N;  // floats between two sources' slots
        const float* base = part + (static_cast<size_t>(b) * p.M + row0) * p.N + col0;
        if ((p.N & 3) == 0) {
          // float4 per lane; the loads of all sources are issued before the first add (fixed summation order)
          const int c4n = cols_valid >> 2;
          const int total = rows_valid * c4n;
          for (int idx = lane; idx < total; idx += 32) {
            const int rr = idx / c4n, c4 = idx - rr * c4n;
            const float* src0 = base + static_cast<size_t>(rr) * p.N + c4 * 4;
            float4 v[DDP_MAX_WORLD];
#pragma unroll
            for (int src = 0; src < DDP_MAX_WORLD; ++src)
              if (src < p.world) v[src] = *reinterpret_cast<const float4*>(src0 + src * src_stride);
            float4 acc4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int src = 0; src < DDP_MAX_WORLD; ++src)
              if (src < p.world) {
                acc4.x += v[src].x; acc4.y += v[src].y; acc4.z += v[src].z; acc4.w += v[src].w;
              }
            OutT* o = out + static_cast<long long>(row0 + rr) * p.ldc + col0 + c4 * 4;
            o[0] = cvt_out<OutT>(acc4.x); o[1] = cvt_out<OutT>(acc4.y);
            o[2] = cvt_out<OutT>(acc4.z); o[3] = cvt_out<OutT>(acc4.w);
          }
        } else {
          for (int rr = 0; rr < rows_valid; ++rr) {
            for (int cc = lane; cc < cols_valid; cc += 32) {
              float s = 0.f;
              for (int src = 0; src < p.world; ++src)
                s += base[src * src_stride + static_cast<size_t>(rr) * p.N + cc];
              out[static_cast<long long>(row0 + rr) * p.ldc + col0 + cc] = cvt_out<OutT>(s);
            }
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
