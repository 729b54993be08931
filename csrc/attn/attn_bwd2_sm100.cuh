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
  uint64_t* st_full = o_final + 1;   // [4] the quadrant's fp32 partial sums are staged (4 warps)
  uint64_t* st_free = st_full + 4;   // [4] the store warp's copies have read them
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(st_free + 4);

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
    if (p.dbg & 1024) {  // row blocks fastest: concurrent CTAs stream the SAME (B_j, V_j) tiles of one head
      const int mb = it % m_blocks;
      const int r = it / m_blocks;
      b0i = r % p.b0;
      b1i = r / p.b0;
      m0 = mb * kAtBM;
      return;
    }
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
          if (remote && !p.gather_ready) {
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
      if (dist && do_produce && !p.gather_ready) attn_push_gather(p, Bt, lane);
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


// ============================================================================================================
// Column side.  Work item = (rank segment, batch, 128-column block, head); B and V of the block stay in shared
// memory, (A_i, dO_i, -L_i, -scale delta_i, transposed mask words) stream by in 64-row steps; step j belongs to
// stream j & 1.  The tiles are computed TRANSPOSED (S^T = B A_i^T: TMEM lanes = columns j, TMEM columns = rows i),
// so P^T and dS^T are K-major A operands for dV += P^T dO_i and dB += dS^T A_i.
//   TMEM: stream s: S^T [128 s, +64)  dP^T [128 s + 64, +64);  dV [256, 320)  dB [320, 384)
//         P^T_s [384 + 32 s, +32)  dS^T_s [448 + 32 s, +32)  (16 columns per 32-row half h)
//   warps as in the row-side kernel; thread = one column j x 32 rows i
// phase_mask bits: 1 = push the gathered operands, 2 = main loop, 4 = owner-side reduction (production: all).
// ============================================================================================================
constexpr int kAb2ColsStage = 2 * kAbHalfTile + 2048;
constexpr int kAb2ColsStages = 7;
constexpr int kAb2ColsOut = 4 * kAtTile;  // staging: 4 quadrants x (dV | dB) x 8 KB of fp32 partial sums  // A_i | dO_i | 64 x -L | 64 x -scale delta | mask 4 x 2 x 32 words

template <int DT>
__global__ void __launch_bounds__(kAtThreads, 1) attn_bwd2_cols_kernel(const __grid_constant__ AttnBwdParams p) {
  using T = typename std::conditional<DT == DDP_BF16, __nv_bfloat16, __half>::type;
  constexpr int STAGES = kAb2ColsStages;  // not a power of two: slots by % STAGES

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sB = smem;                 // B block (128 rows x 128 bytes)
  uint8_t* sV = sB + kAtTile;         // V block
  uint8_t* sOut = sV + kAtTile;       // dV | dB staging (16-bit results: 2 x 4 x 4 KB; fp32 partial sums: 4 x 2 x 8 KB)
  uint8_t* sKV = sOut + kAb2ColsOut;  // STAGES x 18 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + STAGES * kAb2ColsStage);
  uint64_t* kv_full = bars;
  uint64_t* kv_empty = bars + STAGES;
  uint64_t* a_full = bars + 2 * STAGES;
  uint64_t* a_empty = a_full + 1;
  uint64_t* s_full = a_empty + 1;    // [2] S^T_s and dP^T_s landed in TMEM
  uint64_t* s_empty = s_full + 2;    // [2] all 8 warps of the stream hold them in registers
  uint64_t* p_full = s_empty + 2;    // [2] P^T_s and dS^T_s written by all 8 warps of the stream
  uint64_t* o_done = p_full + 2;     // [2] the two products of the stream's step finished
  uint64_t* o_final = o_done + 2;    // [1] the last products of the work item finished
  uint64_t* st_full = o_final + 1;   // [4] the quadrant's fp32 partial sums are staged (4 warps)
  uint64_t* st_free = st_full + 4;   // [4] the store warp's copies have read them
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(st_free + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int Bt = p.b0 * p.b1;
  const bool dist = p.world > 1;
  const bool do_push = (p.phase_mask & 1) != 0;
  const bool do_main = (p.phase_mask & 2) != 0;
  const bool do_reduce = (p.phase_mask & 4) != 0;
  const int n_mb = (p.Tn + kAtBM - 1) / kAtBM;  // 128-column blocks per segment
  const int per_seg = p.b1 * n_mb * p.b0;
  const int items = do_main ? p.world * per_seg : 0;
  const int nj = p.n_steps;

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
    if (p.o1_tma_ok) tma_prefetch_desc(&p.tmOut1);
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
    for (int s = 0; s < 4; ++s) {
      mbar_init(&st_full[s], 4);
      mbar_init(&st_free[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // item -> (segment, outer batch, column block, head); segments rank-rotated with the own one FIRST: its B / V
  // blocks are local, so the all-gather pushes of the peers (single pusher warps, ~0.2 ms for 2 x 16 MB) hide behind
  // it.  (With the own segment last every CTA sat waiting for gathered rows at the start of the kernel.)
  auto decode = [&](int it, int& seg, int& b1i, int& b0i, int& c0) {
    const int si = it / per_seg;
    int r = it - si * per_seg;
    seg = (p.rank + si) % p.world;
    b0i = r % p.b0;
    r /= p.b0;
    const int mb = r % n_mb;
    b1i = r / n_mb;
    c0 = mb * kAtBM;
  };

  if (warp < 4) {
    setmaxnreg_dec<kAtRegsLow>();
    if (warp == 0) {
      // ===================================================== TMA producer (whole warp waits, elected lane issues)
      const bool leader = elect_one();
      uint32_t g = 0;
      uint32_t a_ph = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        int seg, b1i, b0i, c0;
        decode(it, seg, b1i, b0i, c0);
        const int b = b1i * p.b0 + b0i;
        const bool remote = dist && seg != p.rank;
        if (remote && !p.gather_ready) {
          attn_wait_gathered(p, Bt, seg, b, lane);
          fence_proxy_async();
        }
        mbar_wait(a_empty, a_ph ^ 1, 71);
        if (leader) {
          mbar_arrive_expect_tx(a_full, 2 * kAtTile);
          if (remote) {
            tma_load_5d(sB, &p.tmBg, a_full, 0, 0, c0, b, seg);
            tma_load_5d(sV, &p.tmVg, a_full, 0, 0, c0, b, seg);
          } else {
            tma_load_5d(sB, &p.tmB, a_full, 0, 0, c0, b0i, b1i);
            tma_load_5d(sV, &p.tmV, a_full, 0, 0, c0, b0i, b1i);
          }
        }
        __syncwarp();
        a_ph ^= 1;
        // transposed mask words of a step: per 32-column group the 2 row words x 32 columns are contiguous
        int mask_groups = 0;
        const long long cg0 = static_cast<long long>(seg) * p.Tn + c0;
        if (p.mask_bits != nullptr) mask_groups = min(4, (p.Tn - c0 + 31) / 32);
        for (int nb = 0; nb < nj; ++nb, ++g) {
          const uint32_t slot = g % STAGES;
          mbar_wait(&kv_empty[slot], ((g / STAGES) & 1) ^ 1, 72);
          uint8_t* sb = sKV + slot * kAb2ColsStage;
          const int i0 = nb * kAbStep;
          const long long w0 = i0 >> 5;
          const int mask_bytes = mask_groups ? static_cast<int>(min(2ll, p.mb_ld - w0)) * 128 : 0;
          if (leader) {
            mbar_arrive_expect_tx(&kv_full[slot], 2 * kAbHalfTile + 512 + mask_bytes * mask_groups);
            tma_load_5d(sb, &p.tmA, &kv_full[slot], 0, 0, i0, b0i, b1i);
            tma_load_5d(sb + kAbHalfTile, &p.tmG, &kv_full[slot], 0, 0, i0, b0i, b1i);
            const size_t so = static_cast<size_t>(b) * p.stat_ld + i0;
            bulk_load_1d(smem_u32(sb + 2 * kAbHalfTile), p.neg_lse + so, 256, &kv_full[slot]);
            bulk_load_1d(smem_u32(sb + 2 * kAbHalfTile + 256), p.nds + so, 256, &kv_full[slot]);
            for (int gq = 0; gq < mask_groups; ++gq) {
              const uint32_t* src = p.mask_bits + b1i * p.mb_sb1 + (((cg0 >> 5) + gq) * p.mb_ld + w0) * 32;
              bulk_load_1d(smem_u32(sb + 2 * kAbHalfTile + 1024 + gq * 256), src, mask_bytes, &kv_full[slot]);
            }
          }
          __syncwarp();
        }
      }
    } else if (warp == 1) {
      // ===================================================== MMA issuer (whole warp, one elected lane issues)
      constexpr uint32_t kIdS = umma_idesc(DT, 0u, 0u, kAtBM, kAbStep);  // S^T, dP^T: both operands K-major, N = 64
      constexpr uint32_t kIdO = umma_idesc(DT, 0u, 1u, kAtBM, 64);       // dV, dB: A from TMEM, B MN-major
      const bool leader = elect_one();
      const uint32_t sKV_u = smem_u32(sKV);
      const uint64_t dB0 = umma_smem_desc(smem_u32(sB), 16, 1024);
      const uint64_t dV0 = umma_smem_desc(smem_u32(sV), 16, 1024);
      const uint32_t tV = tmem_base + 256, tB = tmem_base + 320;
      uint32_t a_ph = 0, pf_ph0 = 0, pf_ph1 = 0, se_ph0 = 0, se_ph1 = 0;
      uint32_t g0 = 0;
      auto issue_sp = [&](uint32_t g, int s) {  // S^T_s, dP^T_s of the step with running count g
        const uint32_t slot = g % STAGES;
        mbar_wait(&kv_full[slot], (g / STAGES) & 1, 73);
        tc_fence_after();
        const uint64_t da = umma_smem_desc(sKV_u + slot * kAb2ColsStage, 16, 1024);
        const uint64_t dg = umma_smem_desc(sKV_u + slot * kAb2ColsStage + kAbHalfTile, 16, 1024);
        const uint32_t tS = tmem_base + s * 128;
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS, dB0 + 2 * kk, da + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS + 64, dV0 + 2 * kk, dg + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
          umma_commit(&s_full[s]);
        }
        __syncwarp();
      };
      auto issue_acc = [&](uint32_t g, int j, bool last) {  // dV += P^T_s dO_i, dB += dS^T_s A_i, s = j & 1
        const int s = j & 1;
        mbar_wait(&p_full[s], s ? pf_ph1 : pf_ph0, 75);
        if (s) pf_ph1 ^= 1; else pf_ph0 ^= 1;
        tc_fence_after();
        const uint32_t slot = g % STAGES;
        const uint64_t da = umma_smem_desc(sKV_u + slot * kAb2ColsStage, kAbHalfTile, 1024, 2);  // A_i [K = i][N = d]
        const uint64_t dg = umma_smem_desc(sKV_u + slot * kAb2ColsStage + kAbHalfTile, kAbHalfTile, 1024, 2);
        const uint32_t tPt = tmem_base + 384 + s * 32, tDt = tmem_base + 448 + s * 32;
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_ts(tV, tPt + kk * 8, dg + 128 * kk, kIdO, (j | kk) != 0 ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_ts(tB, tDt + kk * 8, da + 128 * kk, kIdO, (j | kk) != 0 ? 1u : 0u);
          umma_commit(&o_done[s]);
          if (last) umma_commit(o_final);
          umma_commit(&kv_empty[slot]);
        }
        __syncwarp();
      };
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        mbar_wait(a_full, a_ph, 74);
        a_ph ^= 1;
        tc_fence_after();
        issue_sp(g0, 0);
        if (nj > 1) issue_sp(g0 + 1, 1);
        for (int j = 0; j < nj; ++j) {
          const int s = j & 1;
          mbar_wait(&s_empty[s], s ? se_ph1 : se_ph0, 79);
          if (s) se_ph1 ^= 1; else se_ph0 ^= 1;
          if (j + 2 < nj) issue_sp(g0 + j + 2, s);
          if (j > 0) issue_acc(g0 + j - 1, j - 1, false);
        }
        issue_acc(g0 + nj - 1, nj - 1, true);
        if (leader) umma_commit(a_empty);
        __syncwarp();
        g0 += static_cast<uint32_t>(nj);
      }
    } else if (warp == 2) {
      if (dist && do_push && !p.gather_ready) attn_push_gather(p, Bt, lane);
    } else if (dist && !(p.dbg & 2048) && lane < 4) {
      // ===================================================== store warp: lane q moves the blocks of quadrant q
      const int q = lane;
      uint32_t sf_ph = 0;
      uint32_t done_ok = 0;  // peers whose done counter this lane has already checked
      const uint32_t sbufq = smem_u32(sOut) + q * 16384;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        int seg, b1i, b0i, c0;
        decode(it, seg, b1i, b0i, c0);
        const int b = b1i * p.b0 + b0i;
        const int col0 = c0 + q * 32;
        const int rows_valid = max(0, min(32, p.Tn - col0));
        if (seg != p.rank && !(done_ok & (1u << seg))) {  // the owner has finished reading its slots of two calls ago
          wait_flag_ge(p.done[p.rank] + seg, p.seq - 2u, p.status, 4);
          done_ok |= 1u << seg;
        }
        mbar_wait(&st_full[q], sf_ph, 82);
        sf_ph ^= 1;
        if (rows_valid > 0) {
#pragma unroll
          for (int kind = 0; kind < 2; ++kind) {
            const uint32_t row_bytes = static_cast<uint32_t>(kind ? p.d : p.dv) * 4u;
            char* dst = p.ws[(p.dbg & 16384) ? p.rank : seg] + (kind ? p.ws_p1_off : p.ws_p0_off) +  // 16384: timing only
                        ((static_cast<size_t>(p.rank) * Bt + b) * p.Tn + col0) * static_cast<size_t>(row_bytes);
            bulk_store_1d(dst, sbufq + kind * 8192, static_cast<uint32_t>(rows_valid) * row_bytes);
          }
          bulk_commit_group();
          bulk_wait_group_read<0>();
        }
        mbar_arrive(&st_free[q]);
        if (rows_valid > 0) bulk_wait_group<0>();
        fence_proxy_async();
        const size_t fidx = static_cast<size_t>(p.rs_flag_off) +
                            (((static_cast<size_t>(p.rank) * Bt + b) * n_mb + c0 / kAtBM) * 4 + q) * 4;
        st_release_sys(p.flags[seg] + fidx, p.seq);  // one flag per (source, batch, block, quadrant)
      }
    }
  } else {
    setmaxnreg_inc<kAtRegsHigh>();
    // ======================================================= elementwise warps (thread = one column j x 32 rows i)
    const int idx = warp - 4;
    const int q = idx & 3;
    const int h = (idx >> 2) & 1;  // 32-row half of the stream's 64-row steps
    const int s = idx >> 3;        // stream: steps j = s, s + 2, ...
    const int part = s * 2 + h;    // epilogue: this warp converts columns [16 part, +16) of dV and dB
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t tS = lane_base + s * 128 + h * 32;
    const uint32_t tP = lane_base + s * 128 + 64 + h * 32;
    const uint32_t tPt = lane_base + 384 + s * 32 + h * 16;
    const uint32_t tDt = lane_base + 448 + s * 32 + h * 16;
    const bool has_mask = p.mask_bits != nullptr;
    const uint32_t stage0 = smem_u32(sKV) + 2 * kAbHalfTile;
    const uint32_t stat_off = stage0 + h * 128;
    const uint32_t mask_off = stage0 + 1024 + q * 256 + h * 128 + lane * 4;
    const uint32_t quad_bar = 1 + q;
    uint32_t sf_ph = 0, of_ph = 0, od_ph = 0;
    uint32_t g0 = 0;
    uint32_t sfree_ph = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int seg, b1i, b0i, c0;
      decode(it, seg, b1i, b0i, c0);
      const int b = b1i * p.b0 + b0i;
      const int col0 = c0 + q * 32;  // first column (row of B / V within the segment) of this warp
      const int col = col0 + lane;
      bool sf_ok = false;
      for (int j = s; j < nj; j += 2) {
        const int nvalid = p.M - (j * kAbStep + h * 32);  // live rows of this thread's half
        const uint32_t slot_off = ((g0 + j) % STAGES) * kAb2ColsStage;
        if (!sf_ok) mbar_wait(&s_full[s], sf_ph, 77);
        sf_ph ^= 1;
        tc_fence_after();
        uint32_t rs[32], rp[32];
        tmem_ld_32x32(tS, rs);
        tmem_ld_32x32(tP, rp);
        // mask words and row statistics came with the step's tiles, which the MMAs behind s_full have read
        uint32_t mk = has_mask ? ld_shared_u32(mask_off + slot_off) : 0u;
        if (nvalid < 32) mk = (nvalid <= 0) ? 0xffffffffu : (mk | (0xffffffffu << nvalid));
        const bool od_ok = mbar_try_wait(&o_done[s], od_ph ^ 1);
        tmem_ld_wait();
        tmem_ld_pin(rs);
        tmem_ld_pin(rp);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[s]);
        // P^T = exp2(scale_log2 S^T - L_i), dS^T = P^T (scale dP^T - scale delta_i): packed pairs along i
        uint32_t pp[16], dd[16];
        const uint32_t sst = stat_off + slot_off;
        // (one branch per step on "any dead element", not one per group of four: -30 instructions per step)
        auto chunk = [&](int c, uint32_t m4) {
          const uint4 nl = ld_shared_v4(sst + c * 16);        // -L_i (same address in every lane: broadcast)
          const uint4 nd = ld_shared_v4(sst + 256 + c * 16);  // -scale delta_i
          float x0, x1, x2, x3, t0, t1, t2, t3;
          ffma2p(x0, x1, __uint_as_float(rs[4 * c]), __uint_as_float(rs[4 * c + 1]), p.scale_log2,
                 __uint_as_float(nl.x), __uint_as_float(nl.y));
          ffma2p(x2, x3, __uint_as_float(rs[4 * c + 2]), __uint_as_float(rs[4 * c + 3]), p.scale_log2,
                 __uint_as_float(nl.z), __uint_as_float(nl.w));
          ffma2p(t0, t1, __uint_as_float(rp[4 * c]), __uint_as_float(rp[4 * c + 1]), p.scale, __uint_as_float(nd.x),
                 __uint_as_float(nd.y));
          ffma2p(t2, t3, __uint_as_float(rp[4 * c + 2]), __uint_as_float(rp[4 * c + 3]), p.scale, __uint_as_float(nd.z),
                 __uint_as_float(nd.w));
          x0 = ex2_approx(x0);
          x1 = ex2_approx(x1);
          x2 = ex2_approx(x2);
          x3 = ex2_approx(x3);
          // dead elements: P = dS = 0 (also keeps a NaN of a fully masked row's L out)
          if (m4 & 1u) x0 = 0.f, t0 = 0.f;
          if (m4 & 2u) x1 = 0.f, t1 = 0.f;
          if (m4 & 4u) x2 = 0.f, t2 = 0.f;
          if (m4 & 8u) x3 = 0.f, t3 = 0.f;
          fmul2(t0, t1, t0, t1, x0, x1);
          fmul2(t2, t3, t2, t3, x2, x3);
          pp[2 * c] = pack2<T>(x0, x1);
          pp[2 * c + 1] = pack2<T>(x2, x3);
          dd[2 * c] = pack2<T>(t0, t1);
          dd[2 * c + 1] = pack2<T>(t2, t3);
        };
        if (mk == 0u) {
#pragma unroll
          for (int c = 0; c < 8; ++c) chunk(c, 0u);
        } else {
#pragma unroll
          for (int c = 0; c < 8; ++c) chunk(c, mk >> (4 * c));
        }
        if (!od_ok) mbar_wait(&o_done[s], od_ph ^ 1, 78);  // the stream's previous products have read P^T_s / dS^T_s
        od_ph ^= 1;
        tc_fence_after();
        tmem_st_32x16(tPt, pp);
        tmem_st_32x16(tDt, dd);
        sf_ok = (j + 2 < nj) && mbar_try_wait(&s_full[s], sf_ph);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[s]);
      }
      g0 += static_cast<uint32_t>(nj);
      // ---- column block finished: dV, dB
      mbar_wait(o_final, of_ph, 80);
      of_ph ^= 1;
      tc_fence_after();
      const int rows_valid = max(0, min(32, p.Tn - col0));
      if (!dist) {
        // final 16-bit results straight from the accumulators (TMA store through a staging tile per quadrant)
#pragma unroll
        for (int kind = 0; kind < 2; ++kind) {
          const uint32_t sbuf = smem_u32(sOut) + kind * kAtTile + q * 4096;
          uint32_t ra[16];
          tmem_ld_32x16(lane_base + 256 + kind * 64 + part * 16, ra);
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
        named_bar_sync(quad_bar, 128);
        if (part == 0 && rows_valid > 0) {
#pragma unroll
          for (int kind = 0; kind < 2; ++kind) {
            const uint32_t sbuf = smem_u32(sOut) + kind * kAtTile + q * 4096;
            const bool tma_ok = kind ? (p.o1_tma_ok != 0) : (p.o0_tma_ok != 0);
            if (tma_ok) {
              if (lane == 0) {
                tma_store_5d(kind ? &p.tmOut1 : &p.tmOut0, sbuf, 0, 0, col0, b0i, b1i);
                bulk_commit_group();
              }
            } else {
              T* base = reinterpret_cast<T*>(kind ? p.out1 : p.out0);
              const long long ld = kind ? p.ld1 : p.ld0;
              T* dst = base + b1i * (kind ? p.o1_sb1 : p.o0_sb1) + b0i * (kind ? p.o1_sb0 : p.o0_sb0) +
                       static_cast<long long>(col0) * ld;
              store_rows_generic<T>(sbuf, dst, ld, rows_valid, kind ? p.d : p.dv,
                                    (kind ? p.o1_vec_ok : p.o0_vec_ok) != 0, lane);
            }
          }
          if (lane == 0) bulk_wait_group_read<0>();
          __syncwarp();
        }
        named_bar_sync(quad_bar, 128);  // the staging tiles belong to the next work item again
      } else if (p.dbg & 2048) {
        named_bar_sync(quad_bar, 128);  // ablation (timing only): no partial sums leave, no reduction
      } else {
        // fp32 partial sums -> the owner's slot [src = this rank][b][Tn][cols].  The 32 columns of a lane quadrant
        // are 32 consecutive rows of that slot, i.e. ONE contiguous block per result: the four warps of the
        // quadrant stage their 16-column parts of dV and dB in shared memory and hand the blocks to the store warp
        // (warp 3), which moves them with bulk copies (TMA engine, shared -> peer memory), waits for them and
        // publishes the flag.  Nothing here waits for a copy, a system-scope fence or another warp: those cost
        // ~13 us per work item when the elementwise warps did them (all CTAs reach this point together).
        mbar_wait(&st_free[q], sfree_ph ^ 1, 81);  // the copies of the previous work item have read the staging blocks
        sfree_ph ^= 1;
        const uint32_t sbufq = smem_u32(sOut) + q * 16384;
#pragma unroll
        for (int kind = 0; kind < 2; ++kind) {
          const int ncols = kind ? p.d : p.dv;
          const uint32_t row_bytes = static_cast<uint32_t>(ncols) * 4u;
          uint32_t ra[16];
          tmem_ld_32x16(lane_base + 256 + kind * 64 + part * 16, ra);
          tmem_ld_wait();
          tmem_ld_pin16(ra);
#pragma unroll
          for (int c = 0; c < 4; ++c)
            if (part * 16 + c * 4 < ncols)
              st_shared_v4(sbufq + kind * 8192 + lane * row_bytes + part * 64 + c * 16,
                           make_uint4(ra[4 * c], ra[4 * c + 1], ra[4 * c + 2], ra[4 * c + 3]));
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&st_full[q]);
        // the next work item's first accumulating MMA follows a step of EACH stream, i.e. every warp's reads above,
        // unless the item has a single step
        if (nj < 2) named_bar_sync(quad_bar, 128);
      }
    }
    // ---------------------------------------------------- reduce-scatter: owner-side reduction
    if (dist && do_reduce && !(p.dbg & (2048 | 4096))) {  // 4096: ablation (timing only), partials leave but nobody reduces
      const int r_items = Bt * n_mb * 16;
      const uint32_t* my_flags = p.flags[p.rank] + p.rs_flag_off;
      for (int item = blockIdx.x * 16 + idx; item < r_items; item += gridDim.x * 16) {
        const int pt = item & 3, qq = (item >> 2) & 3;
        int r = item >> 4;
        const int mb = r % n_mb;
        const int b = r / n_mb;
        if (lane < p.world) {
          const size_t fidx = (((static_cast<size_t>(lane) * Bt + b) * n_mb + mb) * 4 + qq) * 4;
          wait_flag_ge(my_flags + fidx, p.seq, p.status, 5);
        }
        __syncwarp();
        const int b1i = b / p.b0, b0i = b - b1i * p.b0;
        // 8 rows x all columns per warp item: 16 lanes read one row's 16-byte vectors (coalesced 256-byte rows),
        // four rows per lane with the four loads of a source in flight together
        const int vec = lane & 15, rsub = lane >> 4;
        const int row0 = mb * kAtBM + qq * 32 + pt * 8 + rsub;
#pragma unroll
        for (int kind = 0; kind < 2; ++kind) {
          const int ncols = kind ? p.d : p.dv;
          const bool vec_on = vec * 4 < ncols;
          const float* part_base = reinterpret_cast<const float*>(p.ws[p.rank] + (kind ? p.ws_p1_off : p.ws_p0_off));
          const size_t src_stride = static_cast<size_t>(Bt) * p.Tn * ncols;
          const float* src0 = part_base + (static_cast<size_t>(b) * p.Tn + row0) * ncols + vec * 4;
          float4 acc[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int src = 0; src < p.world; ++src) {  // fixed order: deterministic
            float4 t[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
              t[i] = (vec_on && row0 + 2 * i < p.Tn)
                         ? *reinterpret_cast<const float4*>(src0 + src * src_stride + static_cast<size_t>(2 * i) * ncols)
                         : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              acc[i].x += t[i].x; acc[i].y += t[i].y; acc[i].z += t[i].z; acc[i].w += t[i].w;
            }
          }
          T* out = reinterpret_cast<T*>(kind ? p.out1 : p.out0) + b1i * (kind ? p.o1_sb1 : p.o0_sb1) +
                   b0i * (kind ? p.o1_sb0 : p.o0_sb0) + vec * 4;
          const long long ld = kind ? p.ld1 : p.ld0;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (vec_on && row0 + 2 * i < p.Tn) {
              uint2 o;
              o.x = pack2<T>(acc[i].x, acc[i].y);
              o.y = pack2<T>(acc[i].z, acc[i].w);
              *reinterpret_cast<uint2*>(out + static_cast<long long>(row0 + 2 * i) * ld) = o;
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
    tmem_dealloc(tmem_base, 512);
  }
  if (dist && do_reduce && threadIdx.x == 0) {
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
