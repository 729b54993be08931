// Slab-free attention backward for sm_100a (the autograd of module.py:61-69 without the [T/N x T] slabs).
//
// With A = this rank's rows (the reference's keys), B / V = the column-side operands of all ranks (queries /
// values), O = softmax(scale A B^T, mask) V, and from the forward pass the row log-sum-exp L (log2 domain),
// the backward recomputes the probabilities tile by tile instead of reading a saved slab:
//
//     P_ij  = exp2(scale_log2 * A_i.B_j - L_i)            dP_ij = dO_i . V_j
//     dS_ij = scale * P_ij * (dP_ij - delta_i)            delta_i = rowsum(dO_i * O_i)
//     dA_i  = sum_j dS_ij B_j      (row side:    kernel attn_bwd_rows_kernel, accumulates over column blocks)
//     dB_j  = sum_i dS_ij A_i      (column side: kernel attn_bwd_cols_kernel, accumulates over row blocks,
//     dV_j  = sum_i P_ij dO_i       partial sums of the ranks are reduce-scattered inside the kernel)
//
// Two kernels instead of one with atomics: every result is accumulated in TMEM in a fixed order, so the
// backward is deterministic, like the reference.  Both kernels share the structure of the forward kernel
// (attn_fwd_sm100.cuh): warp 0 TMA producer, warp 1 event-driven MMA issuer, warps 2-3 all-gather pushers,
// 16 elementwise warps (thread = one row x 32 columns of a 64-column step; the tiles have no row statistics to
// agree on, L is known, so the threads are independent).
#pragma once
#include "attn_fwd_sm100.cuh"

namespace ddp {

constexpr int kAbStep = 64;                // streamed rows per step
constexpr int kAbHalfTile = 64 * 128;      // 64 rows x 128 bytes

struct alignas(64) AttnBwdParams {
  // rows kernel: tmA = A (box 128 rows), tmG = dO (box 128 rows), tmB / tmV = streamed B, V (box 64 rows)
  // cols kernel: tmB / tmV = stationary B, V (box 128 rows), tmA / tmG = streamed A, dO (box 64 rows)
  CUtensorMap tmA, tmG, tmB, tmV, tmBg, tmVg;
  CUtensorMap tmOut0;  // rows kernel: dA;  cols kernel: dV   (box 64 x 32 rows, TMA store)
  CUtensorMap tmOut1;  // cols kernel: dB
  int M, Tn, d, dv;
  int b0, b1;
  int world, rank;
  int m_pairs;   // rows kernel: pairs of 128-row blocks;  cols kernel: pairs of 128-column blocks per segment
  int n_steps;   // rows kernel: 64-column steps per rank segment;  cols kernel: 64-row steps
  void* out0;
  void* out1;
  long long ld0, o0_sb0, o0_sb1, ld1, o1_sb0, o1_sb1;
  int o0_vec_ok, o0_tma_ok, o1_vec_ok, o1_tma_ok;
  const float* neg_lse;  // [Bt][stat_ld]: -L_i
  const float* nds;      // [Bt][stat_ld]: -scale * delta_i
  long long stat_ld;     // row stride of the two arrays (multiple of 64, the arrays are padded)
  const uint32_t* mask_bits;  // rows kernel: [b1][M][ld] bits along the columns;  cols kernel: transposed,
  long long mb_ld, mb_sb1;    //              [b1][T][ld] bits along the rows
  float scale, scale_log2;
  // collective
  uint32_t seq;
  int phase_mask;
  int chunk_rows, n_chunks;
  char* ws[DDP_MAX_WORLD];
  char* mc_ws;  // multicast alias of the workspace (null: unicast pushes)
  uint32_t* flags[DDP_MAX_WORLD];
  uint32_t* done[DDP_MAX_WORLD];
  uint32_t* exit_counter;
  uint32_t* status;
  long long ws_v_off;   // gathered V region
  long long ws_p0_off;  // cols kernel: fp32 partial slots of dV, [src][Bt][Tn][dv]
  long long ws_p1_off;  // cols kernel: fp32 partial slots of dB, [src][Bt][Tn][d]
  long long rs_flag_off;  // cols kernel: first reduce-scatter flag (after the gather flags)
  const char* push_src[2];
  long long push_ld[2], push_sb0[2], push_sb1[2];
  int push_rows, push_row_bytes[2];
  int gather_ready;  // the gathered B / V are already complete in the workspace the gather maps point at
  int dbg;
};

// All-gather pusher shared by the backward kernels: warp 2 of every CTA copies this rank's B and V in chunks
// into every peer's workspace and publish one flag per (operand, batch entry, chunk).
__device__ __forceinline__ void attn_push_gather(const AttnBwdParams& p, int Bt, int lane) {
  const int gw = blockIdx.x;
  const int gstride = gridDim.x;
  const int per_which = Bt * p.n_chunks;
  const int n_items = 2 * per_which;
  if (gw < n_items) {
    // The peers must have finished call seq-2 before we overwrite this parity of their workspace.
    if (lane > 0 && lane < p.world) {
      const int q = (p.rank + lane) % p.world;
      wait_flag_ge(p.done[p.rank] + q, p.seq - 2u, p.status, 3);
    }
    __syncwarp();
  }
  for (int item = gw; item < n_items; item += gstride) {
    const int which = item / per_which;
    const int rem = item - which * per_which;
    const int b = rem / p.n_chunks;
    const int c = rem - b * p.n_chunks;
    const int b1i = b / p.b0, b0i = b - b1i * p.b0;
    const int r0 = c * p.chunk_rows;
    const int nrows = min(p.chunk_rows, p.push_rows - r0);
    const int rb = p.push_row_bytes[which];
    const int vpr = rb / 16;
    const long long ld = p.push_ld[which];
    const char* src =
        p.push_src[which] + b1i * p.push_sb1[which] + b0i * p.push_sb0[which] + static_cast<long long>(r0) * ld;
    const size_t dst_off = (which ? static_cast<size_t>(p.ws_v_off) : 0) +
                           ((static_cast<size_t>(p.rank) * Bt + b) * p.push_rows + r0) * static_cast<size_t>(rb);
    const int nvec = nrows * vpr;
    for (int v0 = 0; v0 < nvec; v0 += 32 * 4) {
      uint4 qv[4];
      int rr[4], jj[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int v = v0 + u * 32 + lane;
        rr[u] = v / vpr;
        jj[u] = v - rr[u] * vpr;
        if (v < nvec) qv[u] = ld_nc_v4(src + static_cast<long long>(rr[u]) * ld + jj[u] * 16);
      }
      if (p.mc_ws != nullptr) {
        // one multicast store: the switch writes the vector into every rank's gather buffer
        char* dst = p.mc_ws + dst_off;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int v = v0 + u * 32 + lane;
          if (v < nvec) multimem_st_v4(dst + static_cast<size_t>(rr[u]) * rb + jj[u] * 16, qv[u]);
        }
      } else
      for (int i = 1; i < p.world; ++i) {
        const int peer = (p.rank - i + p.world) % p.world;
        char* dst = p.ws[peer] + dst_off;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int v = v0 + u * 32 + lane;
          if (v < nvec) st_v4(dst + static_cast<size_t>(rr[u]) * rb + jj[u] * 16, qv[u]);
        }
      }
    }
    __threadfence_system();
    __syncwarp();
    if (lane > 0 && lane < p.world) {
      const int peer = (p.rank - lane + p.world) % p.world;
      st_release_sys(p.flags[peer] + ((static_cast<size_t>(which) * p.world + p.rank) * Bt + b) * p.n_chunks + c,
                     p.seq);
    }
    __syncwarp();
  }
}

// Lane-parallel wait for every chunk of both gathered operands of rank `seg`, batch entry b (producer warp).
__device__ __forceinline__ void attn_wait_gathered(const AttnBwdParams& p, int Bt, int seg, int b, int lane) {
  const uint32_t* my_flags = p.flags[p.rank];
  for (int c = lane; c < 2 * p.n_chunks; c += 32) {
    const int which = c / p.n_chunks, cc = c - which * p.n_chunks;
    const uint32_t* f = my_flags + ((static_cast<size_t>(which) * p.world + seg) * Bt + b) * p.n_chunks + cc;
    wait_flag_ge(f, p.seq, p.status, 6);
  }
  __syncwarp();
}

// The MMA issuer's idle-spin watchdog (same policy as mbar_wait).  Returns true when the loop must be abandoned.
__device__ __forceinline__ bool attn_issuer_watchdog(bool progress, uint32_t& spins, uint64_t& t0, uint32_t site) {
  if (progress) {
    spins = 0;
    t0 = 0;
    return false;
  }
  if ((++spins & 0xFFFu) != 0) return false;
  const uint64_t now = global_timer_ns();
  if (t0 == 0) {
    t0 = now;
  } else if (now - t0 > 4000000000ull) {
    atomicMax(&g_ddp_timeout_site, site);
    return true;
  }
  return *reinterpret_cast<volatile unsigned int*>(&g_ddp_timeout_site) != 0;
}

// ============================================================================================================
// Row side: dA_i = sum_j dS_ij B_j for this rank's rows.  Work item = (batch, pair of 128-row blocks, head);
// A_w and dO_w (w = 0, 1) stay in shared memory, (B_j, V_j) stream by in 64-row steps.
//   TMEM: S_w [0,128)  dP_w [128,256)  dA_w [256,384)   (64 columns each)
// ============================================================================================================
template <int DT, int STAGES>
__global__ void __launch_bounds__(kAtThreads, 1) attn_bwd_rows_kernel(const __grid_constant__ AttnBwdParams p) {
  using T = typename std::conditional<DT == DDP_BF16, __nv_bfloat16, __half>::type;
  constexpr int kStage = 2 * kAbHalfTile;  // B_j | V_j, 64 rows each

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                     // A_0 A_1
  uint8_t* sG = sA + 2 * kAtTile;         // dO_0 dO_1
  uint8_t* sDS = sG + 2 * kAtTile;        // dS_0 dS_1 (128 rows x 64 columns, 16-bit)
  uint8_t* sKV = sDS + 2 * kAtTile;       // STAGES x 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + STAGES * kStage);
  uint64_t* kv_full = bars;
  uint64_t* kv_empty = bars + STAGES;
  uint64_t* a_full = bars + 2 * STAGES;
  uint64_t* a_empty = a_full + 1;
  uint64_t* s_full = a_empty + 1;    // [2] S_w and dP_w landed in TMEM
  uint64_t* s_empty = s_full + 2;    // [2] all 8 warps of row block w hold their S / dP values in registers
  uint64_t* p_full = s_empty + 2;    // [2] dS_w staged
  uint64_t* o_done = p_full + 2;     // [2] dA_w += dS_w B_j finished
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler, too
  const int lane = threadIdx.x & 31;
  const int Bt = p.b0 * p.b1;
  const bool dist = p.world > 1;
  const bool do_produce = (p.phase_mask & 1) != 0;
  const bool do_consume = (p.phase_mask & 2) != 0;
  const int items = do_consume ? p.b1 * p.m_pairs * p.b0 : 0;
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
      mbar_init(&kv_empty[s], 2);  // one arrival per MMA issuer
    }
    mbar_init(a_full, 1);
    mbar_init(a_empty, 2);
    for (int w = 0; w < 2; ++w) {
      mbar_init(&s_full[w], 1);
      mbar_init(&s_empty[w], 8);
      mbar_init(&p_full[w], 8);
      mbar_init(&o_done[w], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto decode = [&](int it, int& b1i, int& b0i, int& m0) {
    b0i = it % p.b0;
    const int r = it / p.b0;
    const int pair = r % p.m_pairs;
    b1i = r / p.m_pairs;
    m0 = pair * (2 * kAtBM);
  };

  if (warp < 4) {
    setmaxnreg_dec<kAtRegsLow>();
    if (warp == 0) {
      // ===================================================== TMA producer
      int stage = 0;
      uint32_t ph = 0, a_ph = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        int b1i, b0i, m0;
        decode(it, b1i, b0i, m0);
        const int b = b1i * p.b0 + b0i;
        if (lane == 0) {
          mbar_wait(a_empty, a_ph ^ 1, 21);
          mbar_arrive_expect_tx(a_full, 4 * kAtTile);
          tma_load_5d(sA, &p.tmA, a_full, 0, 0, m0, b0i, b1i);
          tma_load_5d(sA + kAtTile, &p.tmA, a_full, 0, 0, m0 + kAtBM, b0i, b1i);
          tma_load_5d(sG, &p.tmG, a_full, 0, 0, m0, b0i, b1i);
          tma_load_5d(sG + kAtTile, &p.tmG, a_full, 0, 0, m0 + kAtBM, b0i, b1i);
        }
        a_ph ^= 1;
        for (int si = 0; si < p.world; ++si) {
          const int seg = (p.rank + si) % p.world;
          const bool remote = dist && seg != p.rank;
          bool need_fence = false;
          if (remote) {
            attn_wait_gathered(p, Bt, seg, b, lane);
            need_fence = true;
          }
          for (int nb = 0; nb < p.n_steps; ++nb) {
            if (lane == 0) {
              if (need_fence) fence_proxy_async();
              mbar_wait(&kv_empty[stage], ph ^ 1, 22);
              uint8_t* sb = sKV + stage * kStage;
              mbar_arrive_expect_tx(&kv_full[stage], kStage);
              const int n0 = nb * kAbStep;
              if (remote) {
                tma_load_5d(sb, &p.tmBg, &kv_full[stage], 0, 0, n0, b, seg);
                tma_load_5d(sb + kAbHalfTile, &p.tmVg, &kv_full[stage], 0, 0, n0, b, seg);
              } else {
                tma_load_5d(sb, &p.tmB, &kv_full[stage], 0, 0, n0, b0i, b1i);
                tma_load_5d(sb + kAbHalfTile, &p.tmV, &kv_full[stage], 0, 0, n0, b0i, b1i);
              }
            }
            need_fence = false;
            if (++stage == STAGES) {
              stage = 0;
              ph ^= 1;
            }
          }
        }
      }
    } else if (warp == 1 || warp == 3) {
      // ===================================================== MMA issuers: warp 1 -> w = 0, warp 3 -> w = 1
      // Static schedule in the order the events arrive in steady state - [S, dP of step j+1 as soon as the
      // elementwise warps hold those of step j] [dA += dS_j B_j once dS_j is staged] - walked by the WHOLE warp
      // (blocking waits included) with one elected lane issuing.  Warp-uniform operands keep the MMA descriptors
      // in uniform registers; issued from inside `if (lane == 0)` every tcgen05.mma was wrapped in an
      // elect / R2UR / branch loop of ~14 dependent instructions, ~100 clk per MMA and 12 MMAs per step here.
      {
        constexpr uint32_t kIdS = umma_idesc(DT, 0u, 0u, kAtBM, kAbStep);  // S, dP: A K-major, B K-major, N = 64
        constexpr uint32_t kIdA = umma_idesc(DT, 0u, 1u, kAtBM, 64);       // dA: A = dS K-major, B = B_j MN-major
        const bool leader = elect_one();
        const int w = (warp == 1) ? 0 : 1;
        uint32_t a_ph = 0, se_ph = 0, pf_ph = 0;
        const uint32_t sKV_u = smem_u32(sKV);
        const uint32_t tS_w = tmem_base + w * 64, tP_w = tmem_base + 128 + w * 64, tA_w = tmem_base + 256 + w * 64;
        const uint64_t dA0 = umma_smem_desc(smem_u32(sA) + w * kAtTile, 16, 1024);   // the operands that do not move
        const uint64_t dG0 = umma_smem_desc(smem_u32(sG) + w * kAtTile, 16, 1024);
        const uint64_t dD0 = umma_smem_desc(smem_u32(sDS) + w * kAtTile, 16, 1024);
        uint32_t slot = 0, slot_ph = 0;  // ring position of the next S / dP step to issue
        auto issue_s = [&]() {
          mbar_wait(&kv_full[slot], slot_ph, 44);
          mbar_wait(&s_empty[w], se_ph ^ 1, 45);
          se_ph ^= 1;
          tc_fence_after();
          const uint64_t db = umma_smem_desc(sKV_u + slot * kStage, 16, 1024);
          const uint64_t dv = umma_smem_desc(sKV_u + slot * kStage + kAbHalfTile, 16, 1024);
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS_w, dA0 + 2 * kk, db + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tP_w, dG0 + 2 * kk, dv + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
            umma_commit(&s_full[w]);
          }
          __syncwarp();
          if (++slot == STAGES) {
            slot = 0;
            slot_ph ^= 1;
          }
        };
        uint32_t pslot = 0;  // ring position of the next dA product
        for (int it = blockIdx.x; it < items; it += gridDim.x) {
          mbar_wait(a_full, a_ph, 24);
          a_ph ^= 1;
          tc_fence_after();
          issue_s();
          for (int j = 0; j < nj; ++j) {
            if (j + 1 < nj) issue_s();
            mbar_wait(&p_full[w], pf_ph, 46);
            pf_ph ^= 1;
            tc_fence_after();
            // B_j as [K = j][N = d], MN-major
            const uint64_t db = umma_smem_desc(sKV_u + pslot * kStage, kAbHalfTile, 1024, 2);
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                umma_ss<false>(tA_w, dD0 + 2 * kk, db + 128 * kk, kIdA, (j | kk) != 0 ? 1u : 0u);
              umma_commit(&o_done[w]);
              umma_commit(&kv_empty[pslot]);  // one of the two arrivals that free the ring slot
            }
            __syncwarp();
            if (++pslot == STAGES) pslot = 0;
          }
          if (leader) umma_commit(a_empty);
          __syncwarp();
        }
      }
    } else {
      if (dist && do_produce) attn_push_gather(p, Bt, lane);  // warp 2
    }
  } else {
    setmaxnreg_inc<kAtRegsHigh>();
    // ======================================================= elementwise warps (thread = one row x 32 columns)
    const int idx = warp - 4;
    const int q = idx & 3;         // TMEM lane quadrant (= warp id mod 4)
    const int h = (idx >> 2) & 1;  // 32-column half of every 64-column step
    const int w = idx >> 3;        // row block of the pair
    const int row_in = q * 32 + lane;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t tS = lane_base + w * 64 + h * 32;
    const uint32_t tP = lane_base + 128 + w * 64 + h * 32;
    const uint32_t tA = lane_base + 256 + w * 64 + h * 32;
    const uint32_t sDSw = smem_u32(sDS) + w * kAtTile;
    const uint32_t sRow = sDSw + row_in * 128;
    const uint32_t sw = lane & 7;
    const uint32_t pair_bar = 1 + w * 4 + q;
    uint32_t sf_ph = 0, od_ph = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int b1i, b0i, m0;
      decode(it, b1i, b0i, m0);
      const int b = b1i * p.b0 + b0i;
      const int row0 = m0 + w * kAtBM + q * 32;
      const int row = row0 + lane;
      const int row_c = min(row, p.M - 1);
      const float neg_lse = p.neg_lse[static_cast<size_t>(b) * p.stat_ld + row_c];
      const float nds = p.nds[static_cast<size_t>(b) * p.stat_ld + row_c];
      const uint32_t* mrow = nullptr;
      if (p.mask_bits != nullptr)  // interleaved layout: the 32 rows of a warp read consecutive words
        mrow = p.mask_bits + b1i * p.mb_sb1 + static_cast<long long>(row_c >> 5) * p.mb_ld * 32 + (row_c & 31);
      // mask word of this row for the 32 columns of step jj, fetched one step ahead (see the forward kernel)
      int pf_nb = 0, pf_si = 0;  // cursor of the prefetch (no integer divisions in the step loop)
      auto mask_word_next = [&]() -> uint32_t {
        uint32_t o = 0u;
        if (mrow == nullptr || pf_si >= p.world) return o;
        int seg = p.rank + pf_si;
        if (seg >= p.world) seg -= p.world;
        const int n0 = pf_nb * kAbStep + h * 32;
        if (p.Tn - n0 > 0) o = __ldg(mrow + ((static_cast<long long>(seg) * p.Tn + n0) >> 5) * 32);
        if (++pf_nb == p.n_steps) {
          pf_nb = 0;
          ++pf_si;
        }
        return o;
      };
      uint32_t mk_next = mask_word_next();
      int cur_nb = 0;
      for (int j = 0; j < nj; ++j) {
        {
          const int n0 = cur_nb * kAbStep + h * 32;
          if (++cur_nb == p.n_steps) cur_nb = 0;
          const int nvalid = max(0, min(32, p.Tn - n0));
          uint32_t mk = mk_next;
          mk_next = mask_word_next();
          if (nvalid < 32) mk |= (nvalid <= 0) ? 0xffffffffu : (0xffffffffu << nvalid);
          mbar_wait(&s_full[w], sf_ph, 27);
          sf_ph ^= 1;
          tc_fence_after();
          uint32_t rs[32], rp[32];
          tmem_ld_32x32(tS, rs);
          tmem_ld_32x32(tP, rp);
          tmem_ld_wait();
          tmem_ld_pin(rs);
          tmem_ld_pin(rp);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&s_empty[w]);
          float v[32];
          if (mk == 0u) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              v[e] = ex2_approx(fmaf(__uint_as_float(rs[e]), p.scale_log2, neg_lse)) *
                     fmaf(__uint_as_float(rp[e]), p.scale, nds);
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              v[e] = ((mk >> e) & 1u) ? 0.f
                                      : ex2_approx(fmaf(__uint_as_float(rs[e]), p.scale_log2, neg_lse)) *
                                            fmaf(__uint_as_float(rp[e]), p.scale, nds);
          }
          if (j > 0) {
            mbar_wait(&o_done[w], od_ph, 28);  // dA_w += dS_w B_{j-1} has read the staging rows
            od_ph ^= 1;
          }
#pragma unroll
          for (int jv = 0; jv < 4; ++jv) {
            uint4 qv;
            qv.x = pack2<T>(v[8 * jv + 0], v[8 * jv + 1]);
            qv.y = pack2<T>(v[8 * jv + 2], v[8 * jv + 3]);
            qv.z = pack2<T>(v[8 * jv + 4], v[8 * jv + 5]);
            qv.w = pack2<T>(v[8 * jv + 6], v[8 * jv + 7]);
            st_shared_v4(sRow + (((h * 4 + jv) ^ sw) * 16), qv);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_full[w]);
        }
      }
      // ---- row block finished: dA_w -> 16-bit -> global
      mbar_wait(&o_done[w], od_ph, 29);
      od_ph ^= 1;
      tc_fence_after();
      const uint32_t sbuf = sDSw + q * 4096;  // 32 rows of the dS tile double as the staging tile of (w, q)
      {
        uint32_t ra[32];
        tmem_ld_32x32(tA, ra);
        tmem_ld_wait();
        tmem_ld_pin(ra);
        float v[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(ra[e]);
        stage_row32<T>(sbuf, lane, v, h);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      named_bar_sync(pair_bar, 64);  // both halves of the staging tile are written
      const int rows_valid = max(0, min(32, p.M - row0));
      if (h == 0 && rows_valid > 0) {
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
      named_bar_sync(pair_bar, 64);  // the staging rows are dS rows of the next work item
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
// Column side: dV_j = sum_i P_ij dO_i and dB_j = sum_i dS_ij A_i over this rank's rows i, for the columns j of
// ALL ranks; partial sums for another rank's columns are stored straight into that rank's workspace and reduced
// there (GEMM -> reduce-scatter in one kernel, like OP_TN of the GEMM kernel).
// Work item = (batch, rank segment, pair of 128-column blocks, head); B_w and V_w (w = 0, 1) stay in shared
// memory, (A_i, dO_i, -L_i, -scale delta_i) stream by in 64-row steps.  The tiles are computed TRANSPOSED
// (S^T = B_w A_i^T: TMEM lanes = columns j, TMEM columns = rows i), so P^T and dS^T can be staged as K-major
// A operands and A_i / dO_i are read once as K-major (S^T, dP^T) and once as MN-major (dB, dV) operands.
//   TMEM: S^T_w [0,128)  dP^T_w [128,256)  dV_w [256,384)  dB_w [384,512)   (64 columns each)
// phase_mask bits: 1 = push the gathered operands, 2 = main loop, 4 = owner-side reduction (production: all).
// ============================================================================================================
constexpr int kAbColsStage = 2 * kAbHalfTile + 1024;  // A_i | dO_i | 64 x -L | 64 x -scale delta | pad

template <int DT, int STAGES>
__global__ void __launch_bounds__(kAtThreads, 1) attn_bwd_cols_kernel(const __grid_constant__ AttnBwdParams p) {
  using T = typename std::conditional<DT == DDP_BF16, __nv_bfloat16, __half>::type;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sB = smem;                     // B_0 B_1
  uint8_t* sV = sB + 2 * kAtTile;         // V_0 V_1
  uint8_t* sPt = sV + 2 * kAtTile;        // P^T_0 P^T_1  (128 columns j x 64 rows i, 16-bit)
  uint8_t* sDt = sPt + 2 * kAtTile;       // dS^T_0 dS^T_1
  uint8_t* sKV = sDt + 2 * kAtTile;       // STAGES x 17 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + STAGES * kAbColsStage);
  uint64_t* kv_full = bars;
  uint64_t* kv_empty = bars + STAGES;
  uint64_t* a_full = bars + 2 * STAGES;
  uint64_t* a_empty = a_full + 1;
  uint64_t* s_full = a_empty + 1;    // [2]
  uint64_t* s_empty = s_full + 2;    // [2]
  uint64_t* p_full = s_empty + 2;    // [2]
  uint64_t* o_done = p_full + 2;     // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler, too
  const int lane = threadIdx.x & 31;
  const int Bt = p.b0 * p.b1;
  const bool dist = p.world > 1;
  const bool do_push = (p.phase_mask & 1) != 0;
  const bool do_main = (p.phase_mask & 2) != 0;
  const bool do_reduce = (p.phase_mask & 4) != 0;
  const int n_mb = (p.Tn + kAtBM - 1) / kAtBM;  // 128-column blocks per segment
  const int per_seg = p.b1 * p.m_pairs * p.b0;
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
      mbar_init(&kv_empty[s], 2);  // one arrival per MMA issuer
    }
    mbar_init(a_full, 1);
    mbar_init(a_empty, 2);
    for (int w = 0; w < 2; ++w) {
      mbar_init(&s_full[w], 1);
      mbar_init(&s_empty[w], 8);
      mbar_init(&p_full[w], 8);
      mbar_init(&o_done[w], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // item -> (segment, outer batch, column-block pair, head); segments rank-rotated with the own one LAST, so
  // the partial sums for the other ranks leave early
  auto decode = [&](int it, int& seg, int& b1i, int& b0i, int& c0) {
    const int si = it / per_seg;
    int r = it - si * per_seg;
    seg = (p.rank + 1 + si) % p.world;
    b0i = r % p.b0;
    r /= p.b0;
    const int pair = r % p.m_pairs;
    b1i = r / p.m_pairs;
    c0 = pair * (2 * kAtBM);
  };

  if (warp < 4) {
    setmaxnreg_dec<kAtRegsLow>();
    if (warp == 0) {
      // ===================================================== TMA producer
      int stage = 0;
      uint32_t ph = 0, a_ph = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        int seg, b1i, b0i, c0;
        decode(it, seg, b1i, b0i, c0);
        const int b = b1i * p.b0 + b0i;
        const bool remote = dist && seg != p.rank;
        if (remote) attn_wait_gathered(p, Bt, seg, b, lane);
        if (lane == 0) {
          if (remote) fence_proxy_async();
          mbar_wait(a_empty, a_ph ^ 1, 31);
          mbar_arrive_expect_tx(a_full, 4 * kAtTile);
          if (remote) {
            tma_load_5d(sB, &p.tmBg, a_full, 0, 0, c0, b, seg);
            tma_load_5d(sB + kAtTile, &p.tmBg, a_full, 0, 0, c0 + kAtBM, b, seg);
            tma_load_5d(sV, &p.tmVg, a_full, 0, 0, c0, b, seg);
            tma_load_5d(sV + kAtTile, &p.tmVg, a_full, 0, 0, c0 + kAtBM, b, seg);
          } else {
            tma_load_5d(sB, &p.tmB, a_full, 0, 0, c0, b0i, b1i);
            tma_load_5d(sB + kAtTile, &p.tmB, a_full, 0, 0, c0 + kAtBM, b0i, b1i);
            tma_load_5d(sV, &p.tmV, a_full, 0, 0, c0, b0i, b1i);
            tma_load_5d(sV + kAtTile, &p.tmV, a_full, 0, 0, c0 + kAtBM, b0i, b1i);
          }
        }
        a_ph ^= 1;
        for (int nb = 0; nb < nj; ++nb) {
          if (lane == 0) {
            mbar_wait(&kv_empty[stage], ph ^ 1, 32);
            uint8_t* sb = sKV + stage * kAbColsStage;
            mbar_arrive_expect_tx(&kv_full[stage], 2 * kAbHalfTile + 512);
            const int i0 = nb * kAbStep;
            tma_load_5d(sb, &p.tmA, &kv_full[stage], 0, 0, i0, b0i, b1i);
            tma_load_5d(sb + kAbHalfTile, &p.tmG, &kv_full[stage], 0, 0, i0, b0i, b1i);
            const size_t so = static_cast<size_t>(b) * p.stat_ld + i0;
            bulk_load_1d(smem_u32(sb + 2 * kAbHalfTile), p.neg_lse + so, 256, &kv_full[stage]);
            bulk_load_1d(smem_u32(sb + 2 * kAbHalfTile + 256), p.nds + so, 256, &kv_full[stage]);
          }
          if (++stage == STAGES) {
            stage = 0;
            ph ^= 1;
          }
        }
      }
    } else if (warp == 1 || warp == 3) {
      // ===================================================== MMA issuers: warp 1 -> w = 0, warp 3 -> w = 1
      // Static schedule walked by the whole warp, one elected lane issues (see the row-side kernel): 16 MMAs per
      // step and column block here.
      {
        constexpr uint32_t kIdS = umma_idesc(DT, 0u, 0u, kAtBM, kAbStep);  // S^T, dP^T
        constexpr uint32_t kIdO = umma_idesc(DT, 0u, 1u, kAtBM, 64);       // dV, dB: B operand MN-major
        const bool leader = elect_one();
        const int w = (warp == 1) ? 0 : 1;
        uint32_t a_ph = 0, se_ph = 0, pf_ph = 0;
        const uint32_t sKV_u = smem_u32(sKV);
        const uint32_t tS_w = tmem_base + w * 64, tP_w = tmem_base + 128 + w * 64, tV_w = tmem_base + 256 + w * 64,
                       tB_w = tmem_base + 384 + w * 64;
        const uint64_t dB0 = umma_smem_desc(smem_u32(sB) + w * kAtTile, 16, 1024);  // the operands that do not move
        const uint64_t dV0 = umma_smem_desc(smem_u32(sV) + w * kAtTile, 16, 1024);
        const uint64_t dPt0 = umma_smem_desc(smem_u32(sPt) + w * kAtTile, 16, 1024);
        const uint64_t dDt0 = umma_smem_desc(smem_u32(sDt) + w * kAtTile, 16, 1024);
        uint32_t slot = 0, slot_ph = 0;
        auto issue_s = [&]() {
          mbar_wait(&kv_full[slot], slot_ph, 47);
          mbar_wait(&s_empty[w], se_ph ^ 1, 48);
          se_ph ^= 1;
          tc_fence_after();
          const uint64_t da = umma_smem_desc(sKV_u + slot * kAbColsStage, 16, 1024);
          const uint64_t dg = umma_smem_desc(sKV_u + slot * kAbColsStage + kAbHalfTile, 16, 1024);
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS_w, dB0 + 2 * kk, da + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tP_w, dV0 + 2 * kk, dg + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
            umma_commit(&s_full[w]);
          }
          __syncwarp();
          if (++slot == STAGES) {
            slot = 0;
            slot_ph ^= 1;
          }
        };
        uint32_t pslot = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x) {
          mbar_wait(a_full, a_ph, 34);
          a_ph ^= 1;
          tc_fence_after();
          issue_s();
          for (int j = 0; j < nj; ++j) {
            if (j + 1 < nj) issue_s();
            mbar_wait(&p_full[w], pf_ph, 49);
            pf_ph ^= 1;
            tc_fence_after();
            const uint64_t da = umma_smem_desc(sKV_u + pslot * kAbColsStage, kAbHalfTile, 1024, 2);
            const uint64_t dg = umma_smem_desc(sKV_u + pslot * kAbColsStage + kAbHalfTile, kAbHalfTile, 1024, 2);
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)  // dV_w += P^T_w dO_i
                umma_ss<false>(tV_w, dPt0 + 2 * kk, dg + 128 * kk, kIdO, (j | kk) != 0 ? 1u : 0u);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)  // dB_w += dS^T_w A_i
                umma_ss<false>(tB_w, dDt0 + 2 * kk, da + 128 * kk, kIdO, (j | kk) != 0 ? 1u : 0u);
              umma_commit(&o_done[w]);
              umma_commit(&kv_empty[pslot]);  // one of the two arrivals that free the ring slot
            }
            __syncwarp();
            if (++pslot == STAGES) pslot = 0;
          }
          if (leader) umma_commit(a_empty);
          __syncwarp();
        }
      }
    } else {
      if (dist && do_push) attn_push_gather(p, Bt, lane);  // warp 2
    }
  } else {
    setmaxnreg_inc<kAtRegsHigh>();
    // ======================================================= elementwise warps (thread = one column j x 32 rows i)
    const int idx = warp - 4;
    const int q = idx & 3;
    const int h = (idx >> 2) & 1;  // 32-row half of every 64-row step
    const int w = idx >> 3;        // column block of the pair
    const int row_in = q * 32 + lane;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t tS = lane_base + w * 64 + h * 32;
    const uint32_t tP = lane_base + 128 + w * 64 + h * 32;
    const uint32_t sPtw = smem_u32(sPt) + w * kAtTile;
    const uint32_t sDtw = smem_u32(sDt) + w * kAtTile;
    const uint32_t sw = lane & 7;
    const uint32_t pair_bar = 1 + w * 4 + q;
    uint32_t sf_ph = 0, od_ph = 0;
    uint32_t g = 0;         // running step count (ring slot of the step = g % STAGES)
    uint32_t done_ok = 0;   // peers whose done counter this warp has already checked
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int seg, b1i, b0i, c0;
      decode(it, seg, b1i, b0i, c0);
      const int b = b1i * p.b0 + b0i;
      const int col0 = c0 + w * kAtBM + q * 32;  // first column (row of B / V within the segment) of this warp
      const int col = col0 + lane;
      const uint32_t* mrow = nullptr;
      if (p.mask_bits != nullptr) {  // transposed + interleaved: the 32 columns of a warp read consecutive words
        const long long cg = static_cast<long long>(seg) * p.Tn + min(col, p.Tn - 1);
        mrow = p.mask_bits + b1i * p.mb_sb1 + (cg >> 5) * p.mb_ld * 32 + (cg & 31);
      }
      auto mask_word = [&](int jj) -> uint32_t {  // transposed mask: 32 rows of step jj, one step ahead
        if (mrow == nullptr || jj >= nj) return 0u;
        const int i0 = jj * kAbStep + h * 32;
        if (p.M - i0 <= 0) return 0u;
        return __ldg(mrow + (i0 >> 5) * 32);
      };
      uint32_t mk_next = mask_word(0);
      for (int j = 0; j < nj; ++j, ++g) {
        const int i0 = j * kAbStep + h * 32;
        const int nvalid = max(0, min(32, p.M - i0));
        uint32_t mk = mk_next;
        mk_next = mask_word(j + 1);
        if (nvalid < 32) mk |= (nvalid <= 0) ? 0xffffffffu : (0xffffffffu << nvalid);
        const uint32_t slot = g % STAGES;
        mbar_wait(&kv_full[slot], (g / STAGES) & 1, 37);  // the row statistics of the step (TMA -> our loads)
        mbar_wait(&s_full[w], sf_ph, 38);
        sf_ph ^= 1;
        tc_fence_after();
        uint32_t rs[32], rp[32];
        tmem_ld_32x32(tS, rs);
        tmem_ld_32x32(tP, rp);
        tmem_ld_wait();
        tmem_ld_pin(rs);
        tmem_ld_pin(rp);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[w]);
        const uint32_t sstat = smem_u32(sKV + slot * kAbColsStage + 2 * kAbHalfTile) + h * 128;
        if (j > 0) {
          mbar_wait(&o_done[w], od_ph, 39);  // the MMAs of step j-1 have read the staging rows
          od_ph ^= 1;
        }
        // P^T and dS^T of this thread's column for the 32 rows: 8 rows -> one 16-byte slot of each staging row
#pragma unroll
        for (int jv = 0; jv < 4; ++jv) {
          uint32_t pp[4], dd[4];
#pragma unroll
          for (int c2 = 0; c2 < 2; ++c2) {
            const int c = jv * 2 + c2;
            const uint4 nl = ld_shared_v4(sstat + c * 16);        // -L_i      (same address in every lane)
            const uint4 nd = ld_shared_v4(sstat + 256 + c * 16);  // -scale delta_i
            const float nlv[4] = {__uint_as_float(nl.x), __uint_as_float(nl.y), __uint_as_float(nl.z),
                                  __uint_as_float(nl.w)};
            const float ndv[4] = {__uint_as_float(nd.x), __uint_as_float(nd.y), __uint_as_float(nd.z),
                                  __uint_as_float(nd.w)};
            float pv[4], dv4[4];
            if (mk == 0u) {  // nothing masked in this thread's 32 rows: no per-element selects
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int ee = c * 4 + e;
                pv[e] = ex2_approx(fmaf(__uint_as_float(rs[ee]), p.scale_log2, nlv[e]));
                dv4[e] = pv[e] * fmaf(__uint_as_float(rp[ee]), p.scale, ndv[e]);
              }
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int ee = c * 4 + e;
                const bool dead = (mk >> ee) & 1u;
                const float pe = dead ? 0.f : ex2_approx(fmaf(__uint_as_float(rs[ee]), p.scale_log2, nlv[e]));
                pv[e] = pe;
                dv4[e] = dead ? 0.f : pe * fmaf(__uint_as_float(rp[ee]), p.scale, ndv[e]);
              }
            }
            pp[c2 * 2] = pack2<T>(pv[0], pv[1]);
            pp[c2 * 2 + 1] = pack2<T>(pv[2], pv[3]);
            dd[c2 * 2] = pack2<T>(dv4[0], dv4[1]);
            dd[c2 * 2 + 1] = pack2<T>(dv4[2], dv4[3]);
          }
          const uint32_t off = row_in * 128 + (((h * 4 + jv) ^ sw) * 16);
          st_shared_v4(sPtw + off, make_uint4(pp[0], pp[1], pp[2], pp[3]));
          st_shared_v4(sDtw + off, make_uint4(dd[0], dd[1], dd[2], dd[3]));
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[w]);
      }
      // ---- column block finished: dV_w, dB_w
      mbar_wait(&o_done[w], od_ph, 40);
      od_ph ^= 1;
      tc_fence_after();
      const int rows_valid = max(0, min(32, p.Tn - col0));
      if (!dist) {
        // final 16-bit results straight from the accumulators (TMA store through a staging tile)
#pragma unroll
        for (int kind = 0; kind < 2; ++kind) {
          const uint32_t sbuf = (kind ? sDtw : sPtw) + q * 4096;
          uint32_t ra[32];
          tmem_ld_32x32(lane_base + 256 + kind * 128 + w * 64 + h * 32, ra);
          tmem_ld_wait();
          tmem_ld_pin(ra);
          float v[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(ra[e]);
          stage_row32<T>(sbuf, lane, v, h);
        }
        tc_fence_before();
        fence_proxy_async_smem();
        named_bar_sync(pair_bar, 64);
        if (h == 0 && rows_valid > 0) {
#pragma unroll
          for (int kind = 0; kind < 2; ++kind) {
            const uint32_t sbuf = (kind ? sDtw : sPtw) + q * 4096;
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
        named_bar_sync(pair_bar, 64);  // the staging rows belong to the next work item again
      } else {
        // fp32 partial sums -> the owner's slot [src = this rank][b][Tn][cols]; one flag per (block, quadrant, half)
        if (seg != p.rank && !(done_ok & (1u << seg))) {
          if (lane == 0) wait_flag_ge(p.done[p.rank] + seg, p.seq - 2u, p.status, 4);
          __syncwarp();
          done_ok |= 1u << seg;
        }
#pragma unroll
        for (int kind = 0; kind < 2; ++kind) {
          const int ncols = kind ? p.d : p.dv;
          uint32_t ra[32];
          tmem_ld_32x32(lane_base + 256 + kind * 128 + w * 64 + h * 32, ra);
          tmem_ld_wait();
          tmem_ld_pin(ra);
          if (col < p.Tn) {
            float* dst = reinterpret_cast<float*>(p.ws[seg] + (kind ? p.ws_p1_off : p.ws_p0_off)) +
                         ((static_cast<size_t>(p.rank) * Bt + b) * p.Tn + col) * ncols + h * 32;
#pragma unroll
            for (int c = 0; c < 8; ++c)
              if (h * 32 + c * 4 < ncols) st_v4(dst + c * 4, make_uint4(ra[4 * c], ra[4 * c + 1], ra[4 * c + 2], ra[4 * c + 3]));
          }
        }
        tc_fence_before();
        __threadfence_system();
        __syncwarp();
        if (lane == 0) {
          const int mb = (c0 + w * kAtBM) / kAtBM;
          if (mb < n_mb) {
            const size_t fidx = static_cast<size_t>(p.rs_flag_off) +
                                (((static_cast<size_t>(p.rank) * Bt + b) * n_mb + mb) * 4 + q) * 2 + h;
            st_release_sys(p.flags[seg] + fidx, p.seq);
          }
        }
        __syncwarp();
      }
    }
    // ---------------------------------------------------- reduce-scatter: owner-side reduction
    if (dist && do_reduce) {
      const int r_items = Bt * n_mb * 8;
      const uint32_t* my_flags = p.flags[p.rank] + p.rs_flag_off;
      for (int item = blockIdx.x * 16 + idx; item < r_items; item += gridDim.x * 16) {
        const int hh = item & 1, qq = (item >> 1) & 3;
        int r = item >> 3;
        const int mb = r % n_mb;
        const int b = r / n_mb;
        if (lane < p.world) {
          const size_t fidx = (((static_cast<size_t>(lane) * Bt + b) * n_mb + mb) * 4 + qq) * 2 + hh;
          wait_flag_ge(my_flags + fidx, p.seq, p.status, 5);
        }
        __syncwarp();
        const int b1i = b / p.b0, b0i = b - b1i * p.b0;
        const int rrow = mb * kAtBM + qq * 32 + lane;
        if (rrow < p.Tn) {
#pragma unroll
          for (int kind = 0; kind < 2; ++kind) {
            const int ncols = kind ? p.d : p.dv;
            const float* part = reinterpret_cast<const float*>(p.ws[p.rank] + (kind ? p.ws_p1_off : p.ws_p0_off));
            const size_t src_stride = static_cast<size_t>(Bt) * p.Tn * ncols;
            const float* src0 = part + (static_cast<size_t>(b) * p.Tn + rrow) * ncols + hh * 32;
            T* out = reinterpret_cast<T*>(kind ? p.out1 : p.out0) + b1i * (kind ? p.o1_sb1 : p.o0_sb1) +
                     b0i * (kind ? p.o1_sb0 : p.o0_sb0) + static_cast<long long>(rrow) * (kind ? p.ld1 : p.ld0) +
                     hh * 32;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              if (hh * 32 + c * 4 < ncols) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int src = 0; src < p.world; ++src) {  // fixed order: deterministic
                  const float4 t = *reinterpret_cast<const float4*>(src0 + src * src_stride + c * 4);
                  acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
                }
                uint2 o;
                o.x = pack2<T>(acc.x, acc.y);
                o.y = pack2<T>(acc.z, acc.w);
                *reinterpret_cast<uint2*>(out + c * 4) = o;
              }
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
