// Slab-free attention backward, second generation (see attn_bwd_sm100.cuh for the math and attn_fwd2_sm100.cuh for
// what the pipeline trace taught): the recomputed tiles never leave TENSOR MEMORY.
//
//     row side     dA_i = sum_j dS_ij B_j                                         (reference: ops.py:35 via module.py:61)
//     column side  dV_j = sum_i P_ij dO_i ,  dB_j = sum_i dS_ij A_i               (reference: ops.py:53, ops.py:34)
//
// Both kernels: one 128-row (128-column) block per work item instead of a pair - twice as many items, so the
// last wave of the flagship shape is 99 % full instead of 87 % - and TWO streams that take the 64-wide steps of
// the item in turn (stream = step parity).  Each stream owns its own S / dP accumulators and its own 16-bit dS
// (P^T, dS^T) tile IN TENSOR MEMORY (tcgen05.st), which enters the accumulating product as the TMEM A operand.
// Per stream the issuer runs
//     [S_s(j+2), dP_s(j+2)]   as soon as the stream's warps hold S_s(j), dP_s(j) in registers
//     [acc += dS_s(j) . X_j]  as soon as they have written dS_s(j)
// so the next tiles of a stream are computed while its warps are still busy with the current ones, and the other
// stream's exponentials fill the rest.  No staging tile in shared memory and no shared-memory traffic for the
// product operand.  One whole-warp MMA issuer with an elected lane (uniform-register operands), mask words
// through the shared-memory ring.
#pragma once
#include "attn_bwd_sm100.cuh"

namespace ddp {

constexpr int kAb2Stages = 8;
constexpr int kAb2RowsStage = 2 * kAbHalfTile + 1024;  // B_j | V_j (64 rows each) | mask words 4 groups x 2 x 32

// ============================================================================================================
// Row side.  Work item = (batch, 128-row block, head); A and dO stay in shared memory, (B_j, V_j) stream by in
// 64-row steps; step j belongs to stream j & 1.
//   TMEM: stream s: S [128 s, +64)  dP [128 s + 64, +64);  dA [256, 320);  dS_s [320 + 32 s, +32)
//         dS of the thread (row, 32-column half h): the 16 columns at 320 + 32 s + 16 h
//   warp 0 producer, warp 1 MMA issuer, warp 2 TMEM alloc + all-gather pusher, warp 3 idle,
//   warps 4-19 elementwise: warp = (stream s, column half h, lane quadrant q), thread = one row x 32 columns
// ============================================================================================================
template <int DT>
__global__ void __launch_bounds__(kAtThreads, 1) attn_bwd2_rows_kernel(const __grid_constant__ AttnBwdParams p) {
  using T = typename std::conditional<DT == DDP_BF16, __nv_bfloat16, __half>::type;
  constexpr int STAGES = kAb2Stages;
  static_assert((STAGES & (STAGES - 1)) == 0, "ring depth must be a power of two");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                 // A   (128 rows x 128 bytes)
  uint8_t* sG = sA + kAtTile;         // dO
  uint8_t* sOut = sG + kAtTile;       // dA staging: 4 quadrants x 4 KB
  uint8_t* sKV = sOut + kAtTile;      // STAGES x 17 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + STAGES * kAb2RowsStage);
  uint64_t* kv_full = bars;
  uint64_t* kv_empty = bars + STAGES;
  uint64_t* a_full = bars + 2 * STAGES;
  uint64_t* a_empty = a_full + 1;
  uint64_t* s_full = a_empty + 1;    // [2] S_s and dP_s landed in TMEM
  uint64_t* s_empty = s_full + 2;    // [2] all 8 warps of the stream hold them in registers
  uint64_t* p_full = s_empty + 2;    // [2] dS_s written by all 8 warps of the stream
  uint64_t* o_done = p_full + 2;     // [2] dA += dS_s B_j finished (dS_s may be overwritten)
  uint64_t* o_final = o_done + 2;    // [1] the last product of the work item into dA finished
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_final + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int Bt = p.b0 * p.b1;
  const bool dist = p.world > 1;
  const bool do_produce = (p.phase_mask & 1) != 0;
  const bool do_consume = (p.phase_mask & 2) != 0;
  const int m_blocks = (p.M + kAtBM - 1) / kAtBM;
  const int items = do_consume ? p.b1 * m_blocks * p.b0 : 0;
  const int nj = p.world * p.n_steps;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmG);
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmV);
    if (dist) {
      tma_prefetch_desc(&p.tmBg);
      tma_prefetch_desc(&p.tmVg);
    }
    if (p.o0_tma_ok) tma_prefetch_desc(&p.tmOut0);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
    }
    mbar_init(a_full, 1);
    mbar_init(a_empty, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&s_empty[s], 8);
      mbar_init(&p_full[s], 8);
      mbar_init(&o_done[s], 1);
    }
    mbar_init(o_final, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto decode = [&](int it, int& b1i, int& b0i, int& m0) {  // heads fastest: concurrent CTAs share mask rows in L2
    b0i = it % p.b0;
    const int r = it / p.b0;
    const int mb = r % m_blocks;
    b1i = r / m_blocks;
    m0 = mb * kAtBM;
  };

  if (warp < 4) {
    setmaxnreg_dec<kAtRegsLow>();
    if (warp == 0) {
      // ===================================================== TMA producer
      // (whole warp walks the loop and waits; one elected lane issues the copies, so that their operands are
      //  warp-uniform: issued from inside `if (lane == 0)` every TMA instruction is wrapped in an elect / R2UR loop,
      //  and the four small mask copies per step alone cost 0.1 ms of this kernel)
      const bool leader = elect_one();
      uint32_t g = 0;  // running step count: ring slot g % STAGES
      uint32_t a_ph = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        int b1i, b0i, m0;
        decode(it, b1i, b0i, m0);
        const int b = b1i * p.b0 + b0i;
        mbar_wait(a_empty, a_ph ^ 1, 61);
        if (leader) {
          mbar_arrive_expect_tx(a_full, 2 * kAtTile);
          tma_load_5d(sA, &p.tmA, a_full, 0, 0, m0, b0i, b1i);
          tma_load_5d(sG, &p.tmG, a_full, 0, 0, m0, b0i, b1i);
        }
        __syncwarp();
        a_ph ^= 1;
        for (int si = 0; si < p.world; ++si) {
          const int seg = (p.rank + si) % p.world;
          const bool remote = dist && seg != p.rank;
          bool need_fence = false;
          if (remote) {
            attn_wait_gathered(p, Bt, seg, b, lane);
            need_fence = true;
          }
          for (int nb = 0; nb < p.n_steps; ++nb, ++g) {
            {
              const uint32_t slot = g & (STAGES - 1);
              if (need_fence) fence_proxy_async();
              mbar_wait(&kv_empty[slot], ((g / STAGES) & 1) ^ 1, 62);
              uint8_t* sb = sKV + slot * kAb2RowsStage;
              const int n0 = nb * kAbStep;
              // mask words of the step: per 32-row group the 2 column words x 32 rows are contiguous in the packed mask
              int mask_bytes = 0, mask_groups = 0;
              const long long w0 = (static_cast<long long>(seg) * p.Tn + n0) >> 5;
              if (p.mask_bits != nullptr && !(p.dbg & 256)) {
                mask_bytes = static_cast<int>(min(2ll, p.mb_ld - w0)) * 128;
                mask_groups = min(4, (p.M - m0 + 31) / 32);
              }
              if (leader) {
                mbar_arrive_expect_tx(&kv_full[slot], 2 * kAbHalfTile + mask_bytes * mask_groups);
                for (int gq = 0; gq < mask_groups; ++gq) {
                  const uint32_t* src =
                      p.mask_bits + b1i * p.mb_sb1 + ((static_cast<long long>(m0 >> 5) + gq) * p.mb_ld + w0) * 32;
                  bulk_load_1d(smem_u32(sb + 2 * kAbHalfTile + gq * 256), src, mask_bytes, &kv_full[slot]);
                }
                if (remote) {
                  tma_load_5d(sb, &p.tmBg, &kv_full[slot], 0, 0, n0, b, seg);
                  tma_load_5d(sb + kAbHalfTile, &p.tmVg, &kv_full[slot], 0, 0, n0, b, seg);
                } else {
                  tma_load_5d(sb, &p.tmB, &kv_full[slot], 0, 0, n0, b0i, b1i);
                  tma_load_5d(sb + kAbHalfTile, &p.tmV, &kv_full[slot], 0, 0, n0, b0i, b1i);
                }
              }
              __syncwarp();
            }
            need_fence = false;
          }
        }
      }
    } else if (warp == 1) {
      // ===================================================== MMA issuer (whole warp, one elected lane issues)
      constexpr uint32_t kIdS = umma_idesc(DT, 0u, 0u, kAtBM, kAbStep);  // S, dP: A K-major, B K-major, N = 64
      constexpr uint32_t kIdA = umma_idesc(DT, 0u, 1u, kAtBM, 64);       // dA: A = dS (TMEM), B = B_j MN-major
      const bool leader = elect_one();
      const uint32_t sKV_u = smem_u32(sKV);
      const uint64_t dA0 = umma_smem_desc(smem_u32(sA), 16, 1024);
      const uint64_t dG0 = umma_smem_desc(smem_u32(sG), 16, 1024);
      const uint32_t tA = tmem_base + 256;
      uint32_t a_ph = 0, pf_ph0 = 0, pf_ph1 = 0;
      uint32_t g0 = 0;  // running step count at the start of the item
      auto issue_sp = [&](uint32_t g, int s) {  // S_s, dP_s of the step with running count g
        const uint32_t slot = g & (STAGES - 1);
        mbar_wait(&kv_full[slot], (g / STAGES) & 1, 63);
        tc_fence_after();
        const uint64_t db = umma_smem_desc(sKV_u + slot * kAb2RowsStage, 16, 1024);
        const uint64_t dv = umma_smem_desc(sKV_u + slot * kAb2RowsStage + kAbHalfTile, 16, 1024);
        const uint32_t tS = tmem_base + s * 128;
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS, dA0 + 2 * kk, db + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS + 64, dG0 + 2 * kk, dv + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
          umma_commit(&s_full[s]);
        }
        __syncwarp();
      };
      uint32_t se_ph0 = 0, se_ph1 = 0;
      auto issue_da = [&](uint32_t g, int j, bool last) {  // dA += dS_s(j) B_j, s = j & 1
        const int s = j & 1;
        mbar_wait(&p_full[s], s ? pf_ph1 : pf_ph0, 65);
        if (s) pf_ph1 ^= 1; else pf_ph0 ^= 1;
        tc_fence_after();
        // the two 16-column dS runs of the stream (column halves h = 0, 1) are K slices 0-1 / 2-3;
        // B_j as [K = j][N = d], MN-major
        const uint64_t db = umma_smem_desc(sKV_u + (g & (STAGES - 1)) * kAb2RowsStage, kAbHalfTile, 1024, 2);
        const uint32_t tD = tmem_base + 320 + s * 32;
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_ts(tA, tD + kk * 8, db + 128 * kk, kIdA, (j | kk) != 0 ? 1u : 0u);
          umma_commit(&o_done[s]);
          if (last) umma_commit(o_final);
          umma_commit(&kv_empty[g & (STAGES - 1)]);
        }
        __syncwarp();
      };
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        mbar_wait(a_full, a_ph, 64);
        a_ph ^= 1;
        tc_fence_after();
        issue_sp(g0, 0);
        if (nj > 1) issue_sp(g0 + 1, 1);
        for (int j = 0; j < nj; ++j) {
          const int s = j & 1;
          // the stream's warps hold S_s(j), dP_s(j): their accumulators are free for step j + 2
          mbar_wait(&s_empty[s], s ? se_ph1 : se_ph0, 69);
          if (s) se_ph1 ^= 1; else se_ph0 ^= 1;
          if (j + 2 < nj) issue_sp(g0 + j + 2, s);
          if (j > 0) issue_da(g0 + j - 1, j - 1, false);
        }
        issue_da(g0 + nj - 1, nj - 1, true);
        if (leader) umma_commit(a_empty);
        __syncwarp();
        g0 += static_cast<uint32_t>(nj);
      }
    } else if (warp == 2) {
      if (dist && do_produce) attn_push_gather(p, Bt, lane);
    }
  } else {
    setmaxnreg_inc<kAtRegsHigh>();
    // ======================================================= elementwise warps (thread = one row x 32 columns)
    const int idx = warp - 4;
    const int q = idx & 3;         // TMEM lane quadrant (= warp id mod 4)
    const int h = (idx >> 2) & 1;  // 32-column half of the stream's 64-column steps
    const int s = idx >> 3;        // stream: steps j = s, s + 2, ...
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t tS = lane_base + s * 128 + h * 32;        // scores
    const uint32_t tP = lane_base + s * 128 + 64 + h * 32;   // dP
    const uint32_t tD = lane_base + 320 + s * 32 + h * 16;   // this thread's 16 columns of dS_s
    const bool has_mask = p.mask_bits != nullptr && !(p.dbg & 512);
    const uint32_t mask_off = smem_u32(sKV) + 2 * kAbHalfTile + q * 256 + h * 128 + lane * 4;
    const uint32_t quad_bar = 1 + q;  // named barrier of the four warps of a lane quadrant (epilogue)
    uint32_t sf_ph = 0, of_ph = 0, od_ph = 0;
    uint32_t g0 = 0;  // running step count at the start of the item (ring slot of step j: (g0 + j) % STAGES)
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int b1i, b0i, m0;
      decode(it, b1i, b0i, m0);
      const int b = b1i * p.b0 + b0i;
      const int row0 = m0 + q * 32;
      const int row_c = min(row0 + lane, p.M - 1);
      const float neg_lse = p.neg_lse[static_cast<size_t>(b) * p.stat_ld + row_c];
      const float nds = p.nds[static_cast<size_t>(b) * p.stat_ld + row_c];
      int cur_nb = s % p.n_steps;  // 64-column block of step j inside its rank segment
      bool sf_ok = false;
      for (int j = s; j < nj; j += 2) {
        const int nvalid = p.Tn - (cur_nb * kAbStep + h * 32);  // live columns of this thread's half (<= 0: none)
        cur_nb += 2;
        while (cur_nb >= p.n_steps) cur_nb -= p.n_steps;
        // Barrier waits cost ~100 clk of latency even when the phase is already complete, and the step is short:
        // every wait is probed early (try_wait issued where its latency overlaps other work) and only blocks,
        // at the point of use, if the probe failed.
        if (!sf_ok) mbar_wait(&s_full[s], sf_ph, 67);
        sf_ph ^= 1;
        tc_fence_after();
        uint32_t rs[32], rp[32];
        tmem_ld_32x32(tS, rs);
        tmem_ld_32x32(tP, rp);
        // the mask words came with the step's B / V tiles, which the score MMAs behind s_full have read: no wait
        uint32_t mk = has_mask ? ld_shared_u32(mask_off + ((g0 + j) & (STAGES - 1)) * kAb2RowsStage) : 0u;
        if (nvalid < 32) mk = (nvalid <= 0) ? 0xffffffffu : (mk | (0xffffffffu << nvalid));
        const bool od_ok = mbar_try_wait(&o_done[s], od_ph ^ 1);
        tmem_ld_wait();
        tmem_ld_pin(rs);
        tmem_ld_pin(rp);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[s]);  // the accumulators may take the stream's next step
        // dS = exp2(scale_log2 S - L) * (scale dP - scale delta), packed 16-bit pairs
        uint32_t pk[16];
        if (mk == 0u) {
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float x0, x1, t0, t1;
            ffma2(x0, x1, __uint_as_float(rs[e]), __uint_as_float(rs[e + 1]), p.scale_log2, neg_lse);
            ffma2(t0, t1, __uint_as_float(rp[e]), __uint_as_float(rp[e + 1]), p.scale, nds);
            x0 = ex2_approx(x0);
            x1 = ex2_approx(x1);
            fmul2(x0, x1, x0, x1, t0, t1);
            pk[e / 2] = pack2<T>(x0, x1);
          }
        } else {
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            const float v0 = ((mk >> e) & 1u) ? 0.f
                                              : ex2_approx(fmaf(__uint_as_float(rs[e]), p.scale_log2, neg_lse)) *
                                                    fmaf(__uint_as_float(rp[e]), p.scale, nds);
            const float v1 = ((mk >> (e + 1)) & 1u)
                                 ? 0.f
                                 : ex2_approx(fmaf(__uint_as_float(rs[e + 1]), p.scale_log2, neg_lse)) *
                                       fmaf(__uint_as_float(rp[e + 1]), p.scale, nds);
            pk[e / 2] = pack2<T>(v0, v1);
          }
        }
        // the stream's previous product has read dS_s (issued a whole stream period ago; passes at once on the
        // stream's first step of the kernel)
        if (!od_ok) mbar_wait(&o_done[s], od_ph ^ 1, 70);
        od_ph ^= 1;
        tc_fence_after();
        tmem_st_32x16(tD, pk);
        sf_ok = (j + 2 < nj) && mbar_try_wait(&s_full[s], sf_ph);  // probe of the stream's next tiles
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[s]);
      }
      g0 += static_cast<uint32_t>(nj);
      // ---- row block finished: dA -> 16-bit -> global; this warp converts columns [16 (2 s + h), +16) of its rows
      mbar_wait(o_final, of_ph, 68);
      of_ph ^= 1;
      tc_fence_after();
      const uint32_t sbuf = smem_u32(sOut) + q * 4096;
      {
        const int part = s * 2 + h;
        uint32_t ra[16];
        tmem_ld_32x16(lane_base + 256 + part * 16, ra);
        tmem_ld_wait();
        tmem_ld_pin16(ra);
        uint4 v0, v1;
        v0.x = pack2<T>(__uint_as_float(ra[0]), __uint_as_float(ra[1]));
        v0.y = pack2<T>(__uint_as_float(ra[2]), __uint_as_float(ra[3]));
        v0.z = pack2<T>(__uint_as_float(ra[4]), __uint_as_float(ra[5]));
        v0.w = pack2<T>(__uint_as_float(ra[6]), __uint_as_float(ra[7]));
        v1.x = pack2<T>(__uint_as_float(ra[8]), __uint_as_float(ra[9]));
        v1.y = pack2<T>(__uint_as_float(ra[10]), __uint_as_float(ra[11]));
        v1.z = pack2<T>(__uint_as_float(ra[12]), __uint_as_float(ra[13]));
        v1.w = pack2<T>(__uint_as_float(ra[14]), __uint_as_float(ra[15]));
        const uint32_t sw = lane & 7;
        st_shared_v4(sbuf + lane * 128 + (((part * 2) ^ sw) * 16), v0);
        st_shared_v4(sbuf + lane * 128 + (((part * 2 + 1) ^ sw) * 16), v1);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      named_bar_sync(quad_bar, 128);  // the staging tile of the quadrant is complete; all four warps have read dA
      const int rows_valid = max(0, min(32, p.M - row0));
      if (h == 0 && s == 0 && rows_valid > 0) {
        if (p.o0_tma_ok) {
          if (lane == 0) {
            tma_store_5d(&p.tmOut0, sbuf, 0, 0, row0, b0i, b1i);
            bulk_commit_group();
            bulk_wait_group_read<0>();
          }
          __syncwarp();
        } else {
          T* dst = reinterpret_cast<T*>(p.out0) + b1i * p.o0_sb1 + b0i * p.o0_sb0 + static_cast<long long>(row0) * p.ld0;
          store_rows_generic<T>(sbuf, dst, p.ld0, rows_valid, p.d, p.o0_vec_ok != 0, lane);
          __syncwarp();
        }
      }
      named_bar_sync(quad_bar, 128);  // the staging tile may be written again (next work item)
    }
  }

  // ---------------------------------------------------------- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
  if (dist && do_consume && threadIdx.x == 0) {
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
