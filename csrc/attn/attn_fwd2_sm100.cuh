// Slab-free attention forward, second generation: the probabilities never leave TENSOR MEMORY.
//
//     O[rows, :] = softmax(scale * A . gather(B)^T  masked) . gather(V)         (reference: module.py:61-69)
//
// Same decomposition as attn_fwd_sm100.cuh (persistent CTA per SM, two 128-row blocks per work item, every
// (row, 64-column half) its own online-softmax stream with its own O accumulator).  What changed, each step
// measured with the pipeline trace (`attn_harness_trace trace`) and the timing harness (`attn_harness perf`):
//   * the 16-bit P tile is written back over the fp32 score tile it came from (tcgen05.st) and enters
//     O += P.V as the TMEM A operand of tcgen05.mma: no st.shared of P, no fence.proxy.async, no 2 x 32 KB per
//     step read back from shared memory by the tensor core, and no wait for "O += P.V of the PREVIOUS step has
//     read the staging tile" in front of every exponential phase;
//   * the score tile of a stream is produced in two 32-column sub-tiles.  A softmax thread works on sub-tile 0
//     and sub-tile 1 of its 64 columns in turn, so the tensor-pipe round trip of one sub-tile
//     ([O += P_s V][S_s(j+1)], issued back to back - the tensor pipe executes one thread's MMAs in order, which
//     is what lets S_s(j+1) overwrite P_s) hides behind the exponentials of the other one;
//   * the MMA issuers are whole warps with one elected lane: issued from inside `if (lane == 0)` every
//     tcgen05.mma was wrapped by the compiler in an elect / R2UR / branch loop of ~14 dependent instructions
//     (~100 clk per MMA; 0.54 -> 0.44 ms for this kernel);
//   * the mask words of a step arrive in shared memory with the step's B / V tiles (TMA bulk copies of the
//     packed mask) instead of per-thread global loads one step ahead.
//
//   TMEM:  S_0 [0,128)  S_1 [128,256)  O_00 O_01 O_10 O_11 [256,512) (64 columns each)
//          sub-tile c = h * 2 + s of row block w: scores at w*128 + c*32 (32 columns), its P over the first 16
//   warp 0      TMA producer (A tiles, ring of (B_j, V_j, mask words) column blocks, peers' chunk flags)
//   warps 1, 3  MMA issuers, one per row block (four sub-tiles each), static schedule with blocking waits
//   warp 2      TMEM alloc; all-gather pusher
//   warps 4-19  softmax: warp = (row block w, column half h, lane quadrant q), thread = one row x 64 columns
#pragma once
#include "attn_fwd_sm100.cuh"

namespace ddp {

constexpr int kAt2StageO = 8 * 4096;  // O staging tiles of the epilogue: (w, q) -> 32 rows x 128 bytes
constexpr int kAt2MaskBytes = 4096;   // mask words of one step: 8 row groups x 4 column words x 32 rows
constexpr int kAt2StageBytes = kAtStageBytes + kAt2MaskBytes;  // B_j | V_j | mask tile

template <int DT, int STAGES, bool kRaw = false>
__global__ void __launch_bounds__(kAtThreads, 1) attn_fwd2_kernel(const __grid_constant__ AttnKParams p) {
  using T = typename std::conditional<DT == DDP_BF16, __nv_bfloat16, __half>::type;
  static_assert(DT == DDP_BF16 || DT == DDP_F16, "16-bit operands only");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                       // 2 x 16 KB
  uint8_t* sOst = sA + 2 * kAtTile;         // 8 x 4 KB
  uint8_t* sKV = sOst + kAt2StageO;         // STAGES x 36 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + STAGES * kAt2StageBytes);
  uint64_t* kv_full = bars;                 // [STAGES]
  uint64_t* kv_empty = bars + STAGES;       // [STAGES]
  uint64_t* a_full = bars + 2 * STAGES;     // [1]
  uint64_t* a_empty = a_full + 1;           // [1]
  uint64_t* s_full = a_empty + 1;           // [8]  sub-tile w*4 + c of S landed in TMEM
  uint64_t* p_full = s_full + 8;            // [8]  its P written over it (and O_wh rescaled)
  uint64_t* o_done = p_full + 8;            // [4]  an O_wh += P V product finished (two per step and stream)
  uint64_t* o_final = o_done + 4;           // [4]  the LAST product of the work item into O_wh finished
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_final + 4);
  float2* sStat = reinterpret_cast<float2*>(tmem_slot + 4);  // [2][2][128] (max, sum) of every stream, for the merge

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler, too
  const int lane = threadIdx.x & 31;
  const int Bt = p.b0 * p.b1;
  const bool dist = p.world > 1;
  const bool do_produce = (p.phase_mask & 1) != 0;
  const bool do_consume = (p.phase_mask & 2) != 0;
  const int items = do_consume ? p.b1 * p.m_pairs * p.b0 : 0;
  const int nj = p.world * p.n_blocks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmV);
    if (dist) {
      tma_prefetch_desc(&p.tmBg);
      tma_prefetch_desc(&p.tmVg);
    }
    if (p.o_tma_ok) tma_prefetch_desc(&p.tmO);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 2);  // one arrival per MMA issuer
    }
    mbar_init(a_full, 1);
    mbar_init(a_empty, 2);
    for (int i = 0; i < 8; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
    }
    for (int st = 0; st < 4; ++st) {
      mbar_init(&o_done[st], 1);
      mbar_init(&o_final[st], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kAtTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work item -> (outer batch, row-block pair, head); heads fastest so the CTAs that run at the same time share
  // their mask rows in L2
  auto decode = [&](int it, int& b1i, int& b0i, int& m0) {
    if (p.dbg & 1024) {  // row-block pairs fastest: concurrent CTAs stream the SAME (B_j, V_j) tiles of one head
      const int pair = it % p.m_pairs;
      const int r = it / p.m_pairs;
      b0i = r % p.b0;
      b1i = r / p.b0;
      m0 = pair * (2 * kAtBM);
      return;
    }
    b0i = it % p.b0;
    const int r = it / p.b0;
    const int pair = r % p.m_pairs;
    b1i = r / p.m_pairs;
    m0 = pair * (2 * kAtBM);
  };

  if (warp < 4) {
  setmaxnreg_dec<kAtRegsLow>();
  if (warp == 0) {
    // ======================================================= TMA producer (whole warp walks the loop)
    // (one elected lane issues the copies with warp-uniform operands, see the MMA issuers below)
    const bool leader = elect_one();
    int stage = 0;
    uint32_t ph = 0, a_ph = 0;
    const uint32_t* my_flags = p.flags[p.rank];
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int b1i, b0i, m0;
      decode(it, b1i, b0i, m0);
      const int b = b1i * p.b0 + b0i;
      mbar_wait(a_empty, a_ph ^ 1, 51);
      if (leader) {
        mbar_arrive_expect_tx(a_full, 2 * kAtTile);
        tma_load_5d(sA, &p.tmA, a_full, 0, 0, m0, b0i, b1i);
        tma_load_5d(sA + kAtTile, &p.tmA, a_full, 0, 0, m0 + kAtBM, b0i, b1i);
      }
      __syncwarp();
      a_ph ^= 1;
      for (int si = 0; si < p.world; ++si) {
        const int seg = (p.rank + si) % p.world;
        const bool remote = dist && seg != p.rank;
        bool need_fence = false;
        if (remote) {
          // every chunk of both gathered operands of rank `seg` for this batch entry, lane-parallel
          for (int c = lane; c < 2 * p.n_chunks; c += 32) {
            const int which = c / p.n_chunks, cc = c - which * p.n_chunks;
            const uint32_t* f =
                my_flags + ((static_cast<size_t>(which) * p.world + seg) * Bt + b) * p.n_chunks + cc;
            wait_flag_ge(f, p.seq, p.status, 6);
          }
          __syncwarp();
          need_fence = true;
        }
        for (int nb = 0; nb < p.n_blocks; ++nb) {
          {
            if (need_fence) fence_proxy_async();  // peer (generic-proxy) writes -> our TMA (async-proxy) reads
            mbar_wait(&kv_empty[stage], ph ^ 1, 52);
            uint8_t* sb = sKV + stage * kAt2StageBytes;
            const int n0 = nb * kAtBN;
            // mask words of the step: per 32-row group of the item, the 4 column words x 32 rows are one
            // contiguous run of the packed mask ([row group][column word][32 rows]); row groups past the last
            // row and column words past the end of the mask row are not copied (the consumers never use them)
            int mask_bytes = 0, mask_groups = 0;
            const long long w0 = (static_cast<long long>(seg) * p.Tn + n0) >> 5;
            if (p.mask_bits != nullptr) {
              const long long nw = min(4ll, p.mb_ld - w0);
              mask_bytes = static_cast<int>(nw) * 128;
              mask_groups = min(8, (p.M - m0 + 31) / 32);
            }
            if (leader) {
              mbar_arrive_expect_tx(&kv_full[stage], kAtStageBytes + mask_bytes * mask_groups);
              for (int gq = 0; gq < mask_groups; ++gq) {
                const uint32_t* src =
                    p.mask_bits + b1i * p.mb_sb1 + ((static_cast<long long>(m0 >> 5) + gq) * p.mb_ld + w0) * 32;
                bulk_load_1d(smem_u32(sb + kAtStageBytes + gq * 512), src, mask_bytes, &kv_full[stage]);
              }
              if (remote) {
                tma_load_5d(sb, &p.tmBg, &kv_full[stage], 0, 0, n0, b, seg);
                tma_load_5d(sb + kAtTile, &p.tmVg, &kv_full[stage], 0, 0, n0, b, seg);
              } else {
                tma_load_5d(sb, &p.tmB, &kv_full[stage], 0, 0, n0, b0i, b1i);
                tma_load_5d(sb + kAtTile, &p.tmV, &kv_full[stage], 0, 0, n0, b0i, b1i);
              }
            }
            __syncwarp();
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
    // ======================================================= MMA issuers, one per row block
    // Static schedule per 32-column sub-tile c = h * 2 + s of row block w (four per issuer and step), in the
    // order the softmax warps finish them (h0 s0, h1 s0, h0 s1, h1 s1):
    //     [wait P_c -> O_wh += P_c V_j[rows 32 c ..]]   [S_c(j+1) = A_w B_{j+1}[rows 32 c ..]^T]
    // The WHOLE warp walks the schedule (waits included) and one elected lane issues, so that the MMA operands
    // are warp-uniform and live in uniform registers.
    {
      constexpr uint32_t kIdS = umma_idesc(DT, 0u, 0u, kAtBM, 32);  // A K-major, B K-major, N = 32 columns
      constexpr uint32_t kIdO = umma_idesc(DT, 0u, 1u, kAtBM, 64);  // A = P (TMEM), B = V MN-major
      const bool leader = elect_one();
      const int w = (warp == 1) ? 0 : 1;
      uint32_t a_ph = 0, pf_ph = 0;
      const uint32_t sKV_u = smem_u32(sKV);
      const uint32_t tS_w = tmem_base + w * kAtBN;
      const uint32_t tO_w = tmem_base + 2 * kAtBN + w * 128;
      const uint64_t dA0 = umma_smem_desc(smem_u32(sA) + w * kAtTile, 16, 1024);
      // scores of sub-tile c of the step in ring slot `slot`: B rows [32 c, 32 c + 32) = 4 KB per sub-tile
      auto issue_s = [&](uint32_t slot, int c) {
        const uint64_t db = umma_smem_desc(sKV_u + slot * kAt2StageBytes + c * (32 * 128), 16, 1024);
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_ss<false>(tS_w + c * 32, dA0 + 2 * kk, db + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
          umma_commit(&s_full[w * 4 + c]);
        }
        __syncwarp();
      };
      uint32_t slot = 0, slot_ph = 0;  // ring position of the step whose P tiles are awaited
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        mbar_wait(a_full, a_ph, 54);
        a_ph ^= 1;
        mbar_wait(&kv_full[slot], slot_ph, 53);
        tc_fence_after();
#pragma unroll
        for (int e = 0; e < 4; ++e) issue_s(slot, (e & 1) * 2 + (e >> 1));
        for (int j = 0; j < nj; ++j) {
          uint32_t nslot = slot + 1, nslot_ph = slot_ph;
          if (nslot == STAGES) {
            nslot = 0;
            nslot_ph ^= 1;
          }
          const bool more = j + 1 < nj;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int sub = e >> 1, h = e & 1, c = h * 2 + sub;
            mbar_wait(&p_full[w * 4 + c], pf_ph, 55);
            if (e == 0 && more) mbar_wait(&kv_full[nslot], nslot_ph, 53);
            tc_fence_after();
            // V rows [32 c, 32 c + 32) of the step as [K = 32][N = dv], MN-major: two K = 16 slices of 2 KB
            const uint64_t dv =
                umma_smem_desc(sKV_u + slot * kAt2StageBytes + kAtTile + c * (2 * 2048), kAtTile, 1024, 2);
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < 2; ++kk)
                umma_ts(tO_w + h * 64, tS_w + c * 32 + kk * 8, dv + 128 * kk, kIdO, (j | sub | kk) != 0 ? 1u : 0u);
              umma_commit(&o_done[w * 2 + h]);
              if (!more && sub == 1) umma_commit(&o_final[w * 2 + h]);
            }
            __syncwarp();
            if (more) issue_s(nslot, c);
          }
          pf_ph ^= 1;
          if (leader) umma_commit(&kv_empty[slot]);  // one of the two arrivals that hand the ring slot back
          __syncwarp();
          slot = nslot;
          slot_ph = nslot_ph;
        }
        if (leader) umma_commit(a_empty);
        __syncwarp();
      }
    }
  } else {
    // ======================================================= all-gather pushers (B and V)
    if (dist && do_produce) {  // warp 2 only (warp 3 is the second MMA issuer)
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
        const char* src = p.push_src[which] + b1i * p.push_sb1[which] + b0i * p.push_sb0[which] +
                          static_cast<long long>(r0) * ld;
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
            const int peer = (p.rank - i + p.world) % p.world;  // rank-1 consumes our rows first
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
  }
  } else {
    setmaxnreg_inc<kAtRegsHigh>();
    // ======================================================= softmax warps (thread = one row x 64 columns)
    const int idx = warp - 4;
    const int q = idx & 3;         // TMEM lane quadrant this warp may access (= warp id mod 4)
    const int h = (idx >> 2) & 1;  // column half of every step
    const int w = idx >> 3;        // row block of the pair
    const int st = w * 2 + h;      // stream
    const int row_in = q * 32 + lane;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t tS = lane_base + w * kAtBN + h * 64;  // sub-tile s at + 32 s; its P over its first 16 columns
    const uint32_t tO = lane_base + 2 * kAtBN + st * 64;
    const uint32_t pair_bar = 1 + w * 4 + q;  // named barrier shared with the warp of the other column half
    uint64_t* const sf_bar = &s_full[w * 4 + h * 2];
    uint64_t* const pf_bar = &p_full[w * 4 + h * 2];
    const bool has_mask = p.mask_bits != nullptr;
    // this thread's two mask words of a step inside the stage's mask tile
    const uint32_t mask_off = smem_u32(sKV) + kAtStageBytes + (w * 4 + q) * 512 + (h * 2) * 128 + lane * 4;
    uint32_t sf_ph = 0, of_ph = 0;
    uint32_t npv = 0;                // O_st += P V products issued for this stream so far (two per step)
    uint32_t slot = 0;  // ring position of the current step (for the mask tile)
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int b1i, b0i, m0;
      decode(it, b1i, b0i, m0);
      const int b = b1i * p.b0 + b0i;
      const int row0 = m0 + w * kAtBM + q * 32;  // first row of this warp
      const int row = row0 + lane;
      float m_used = -INFINITY, l = 0.f;
      int cur_nb = 0;
      for (int j = 0; j < nj; ++j) {
        const int n0 = cur_nb * kAtBN + h * 64;  // first column of this thread's half
        if (++cur_nb == p.n_blocks) cur_nb = 0;
        const int nvalid = p.Tn - n0;            // live columns of the half (may be <= 0 or >= 64)
        uint32_t mk[2] = {0u, 0u};
        if (has_mask) {
          // the mask tile landed with the step's B / V tiles, which the score MMAs behind s_full have read; wait
          // for the first sub-tile here (the wait in the sub-tile loop then passes at once)
          mbar_wait(sf_bar, sf_ph, 56);
          mk[0] = ld_shared_u32(mask_off + slot * kAt2StageBytes);
          mk[1] = ld_shared_u32(mask_off + slot * kAt2StageBytes + 128);
          if (++slot == STAGES) slot = 0;
        }
        if (nvalid < 64) {
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int lo = i * 32;
            if (nvalid <= lo) {
              mk[i] = 0xffffffffu;
            } else if (nvalid < lo + 32) {
              mk[i] |= 0xffffffffu << (nvalid - lo);
            }
          }
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint32_t mks = mk[s];
          mbar_wait(sf_bar + s, sf_ph, 57);
          tc_fence_after();
          uint32_t r[32];
          tmem_ld_32x32(tS + s * 32, r);
          tmem_ld_wait();
          tmem_ld_pin(r);
          // ---- maximum of the raw scores over the live columns
          float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
          if (mks == 0u) {
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              mx0 = fmaxf(mx0, __uint_as_float(r[e]));
              mx1 = fmaxf(mx1, __uint_as_float(r[e + 1]));
              mx2 = fmaxf(mx2, __uint_as_float(r[e + 2]));
              mx3 = fmaxf(mx3, __uint_as_float(r[e + 3]));
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (!((mks >> e) & 1u)) mx0 = fmaxf(mx0, __uint_as_float(r[e]));
          }
          const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;  // scale_log2 > 0
          // Lazy rescale: the reference point of the exponentials only moves when the maximum grew by more than
          // 2^kAtTau, so P stays <= 2^kAtTau and the O accumulator is rarely touched.
          const bool need = !kRaw && (m_blk > m_used + kAtTau);
          if ((j | s) != 0 && __any_sync(0xffffffffu, need)) {
            // every product into O_st issued so far must have finished before O_st is rescaled (the only wait for
            // them inside the step loop); the next one is not issued before this warp arrives on p_full below
            mbar_wait(&o_done[st], (npv - 1u) & 1u, 58);
            tc_fence_after();
            const float f = need ? ex2_approx(m_used - m_blk) : 1.f;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              uint32_t ro[32];
              tmem_ld_32x32(tO + hh * 32, ro);
              tmem_ld_wait();
              tmem_ld_pin(ro);
#pragma unroll
              for (int e = 0; e < 32; ++e) ro[e] = __float_as_uint(__uint_as_float(ro[e]) * f);
              tmem_st_32x32(tO + hh * 32, ro);
            }
            l *= f;
          }
          if (need) m_used = m_blk;
          const float neg_m = (kRaw || m_used == -INFINITY) ? 0.f : -m_used;
          // ---- P = exp2(scale_log2 * S - m) as packed 16-bit pairs, written over the first 16 columns of the sub-tile
          uint32_t pk[16];
          float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
          if (kRaw) {
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
              const float v0 = ((mks >> e) & 1u) ? 0.f : __uint_as_float(r[e]) * p.scale_log2;
              const float v1 = ((mks >> (e + 1)) & 1u) ? 0.f : __uint_as_float(r[e + 1]) * p.scale_log2;
              pk[e / 2] = pack2<T>(v0, v1);
            }
          } else if (mks == 0u) {
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              float x0, x1, x2, x3;
              ffma2(x0, x1, __uint_as_float(r[e]), __uint_as_float(r[e + 1]), p.scale_log2, neg_m);
              ffma2(x2, x3, __uint_as_float(r[e + 2]), __uint_as_float(r[e + 3]), p.scale_log2, neg_m);
              x0 = ex2_approx(x0);
              x1 = ex2_approx(x1);
              x2 = ex2_approx(x2);
              x3 = ex2_approx(x3);
              fadd2(l0, l1, l0, l1, x0, x1);
              fadd2(l2, l3, l2, l3, x2, x3);
              pk[e / 2] = pack2<T>(x0, x1);
              pk[e / 2 + 1] = pack2<T>(x2, x3);
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
              const float v0 = ((mks >> e) & 1u) ? 0.f : ex2_approx(fmaf(__uint_as_float(r[e]), p.scale_log2, neg_m));
              const float v1 =
                  ((mks >> (e + 1)) & 1u) ? 0.f : ex2_approx(fmaf(__uint_as_float(r[e + 1]), p.scale_log2, neg_m));
              l0 += v0;
              l1 += v1;
              pk[e / 2] = pack2<T>(v0, v1);
            }
          }
          l += (l0 + l1) + (l2 + l3);
          tmem_st_32x16(tS + s * 32, pk);
          tmem_st_wait();  // P (and a rescaled O_st) are in tensor memory
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(pf_bar + s);
          ++npv;
        }
        sf_ph ^= 1;
      }
      // ---- row block finished: merge the two column-half streams of every row, O = (Oa fa + Ob fb) / l
      // (o_done cannot be used here: two of its phases may still be outstanding, which a parity wait cannot tell
      //  from none; inside the step loop the s_full wait in front of the rescale bounds it to one)
      mbar_wait(&o_final[st], of_ph, 59);
      of_ph ^= 1;
      tc_fence_after();
      sStat[(w * 2 + h) * kAtBM + row_in] = make_float2(m_used, l);
      tc_fence_before();
      named_bar_sync(pair_bar, 64);  // partner's statistics written; both streams' last MMAs have finished
      tc_fence_after();
      const float2 other = sStat[(w * 2 + (h ^ 1)) * kAtBM + row_in];
      const float m_all = fmaxf(m_used, other.x);
      float f_own = 1.f, f_oth = 1.f;
      if (!kRaw) {
        f_own = (m_used == -INFINITY) ? 0.f : ex2_approx(m_used - m_all);
        f_oth = (other.x == -INFINITY) ? 0.f : ex2_approx(other.x - m_all);
      }
      const float l_all = l * f_own + other.y * f_oth;
      const float inv = kRaw ? 1.f : 1.f / l_all;  // l == 0 (fully masked row): inf -> NaN row, as in the reference
      const float fa = (h == 0 ? f_own : f_oth) * inv, fb = (h == 0 ? f_oth : f_own) * inv;
      // this thread produces output columns [h*32, h*32+32) of its row from both accumulators
      const uint32_t sbuf = smem_u32(sOst) + (w * 4 + q) * 4096;
      {
        uint32_t ra[32], rb[32];
        tmem_ld_32x32(lane_base + 2 * kAtBN + (w * 2) * 64 + h * 32, ra);
        tmem_ld_32x32(lane_base + 2 * kAtBN + (w * 2 + 1) * 64 + h * 32, rb);
        tmem_ld_wait();
        tmem_ld_pin(ra);
        tmem_ld_pin(rb);
        float v[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          // 0 * inf = NaN marks a fully masked row; a dead stream (factor 0, accumulator 0) contributes 0 * inv
          v[e] = __uint_as_float(ra[e]) * fa + __uint_as_float(rb[e]) * fb;
        }
        stage_row32<T>(sbuf, lane, v, h);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      named_bar_sync(pair_bar, 64);  // both halves of the staging tile are written
      const int rows_valid = max(0, min(32, p.M - row0));
      if (h == 0 && rows_valid > 0) {
        if (p.o_tma_ok) {
          if (lane == 0) {
            tma_store_5d(&p.tmO, sbuf, 0, 0, row0, b0i, b1i);
            bulk_commit_group();
            bulk_wait_group_read<0>();
          }
          __syncwarp();
        } else {
          T* dst = reinterpret_cast<T*>(p.O) + b1i * p.o_sb1 + b0i * p.o_sb0 + static_cast<long long>(row0) * p.ldo;
          store_rows_generic<T>(sbuf, dst, p.ldo, rows_valid, p.dv, p.o_vec_ok != 0, lane);
          __syncwarp();
        }
        if (p.lse != nullptr && row < p.M) p.lse[static_cast<size_t>(b) * p.M + row] = m_all + log2f(l_all);
      }
      // the staging tile is written again at the end of the next work item, after the store above has read it
      named_bar_sync(pair_bar, 64);
    }
  }

  // ---------------------------------------------------------- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kAtTmemCols);
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
