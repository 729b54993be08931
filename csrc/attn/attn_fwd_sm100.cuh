// Slab-free attention forward for sm_100a with the all-gather fused in:
//
//     O[rows, :] = softmax(scale * A . gather(B)^T  masked) . gather(V)         (reference: module.py:61-69)
//
// A = this rank's rows (the reference's keys), B / V = the column-side operands of ALL ranks (the reference's
// queries / values).  The [T/N x T] score slab never exists: scores live in TMEM, probabilities go through
// shared memory straight back into the tensor core (online softmax), only O (and the row log-sum-exp) is
// written.  One persistent CTA per SM, 384 threads:
//
//   warp 0      TMA producer: A tiles (two 128-row blocks per work item) and a ring of (B_j, V_j) column blocks;
//               polls the peers' per-chunk flags before touching gathered rows
//   warps 1, 3  MMA issuers (one per row block w):   S_w = A_w . B_j^T (128x128, K = d) for the two row blocks w = 0, 1 and, per row block,
//               TWO independent accumulators O_{w,h} += P_w[:, half h] . V_j[half h] (128 x dv, K = 64); event
//               driven, so S_w(j+1) is computed while the softmax warps are still busy with step j
//   warp 2      TMEM alloc; all-gather pusher (B and V chunks -> every peer's workspace, one flag per chunk)
//   warps 4-19  softmax: warp = (row block w, column half h, lane quadrant q), thread = one row x 64 columns held
//               in registers.  Every (row, half) is its own online-softmax stream (own running max and sum, own
//               O accumulator), so no per-step exchange between the two threads of a row is needed; the two
//               streams are merged once, when the row block is finished.  The running max only moves when it grew
//               by more than 2^8 (lazy rescale of the O accumulator in TMEM).  Four softmax warps per scheduler
//               keep the MUFU pipe (ex2, the bound of this kernel at head dim 64) busy.
#pragma once
#include "../gemm/gemm_sm100.cuh"

namespace ddp {

constexpr int kAtThreads = 640;
constexpr int kAtBM = 128;                 // rows per warpgroup tile
constexpr int kAtBN = 128;                 // columns per step
constexpr int kAtTile = 128 * 128;         // 128 rows x 128 bytes
constexpr int kAtPBytes = 2 * kAtTile;     // P tile of one warpgroup: two 64-column k-chunks
constexpr int kAtStageBytes = 2 * kAtTile; // B_j | V_j
constexpr uint32_t kAtTmemCols = 512;      // S0 S1 (128 each) | O00 O01 O10 O11 (64 each)
constexpr float kAtTau = 8.f;              // lazy-rescale threshold (log2 units)
constexpr int kAtRegsLow = 56;             // registers per thread of warps 0-3 after setmaxnreg
constexpr int kAtRegsHigh = 104;           // ... of the softmax warps (128 x 56 + 512 x 104 <= 640 x 96)

struct alignas(64) AttnKParams {
  CUtensorMap tmA;   // [b1][b0][M][d]     box 64 x 128 rows
  CUtensorMap tmB;   // [b1][b0][Tn][d]    box 64 x 128 rows (K-major B operand of S)
  CUtensorMap tmV;   // [b1][b0][Tn][dv]   box 64 x 128 rows (MN-major B operand of O)
  CUtensorMap tmBg;  // gathered [world][Bt][Tn][d]
  CUtensorMap tmVg;  // gathered [world][Bt][Tn][dv]
  CUtensorMap tmO;   // [b1][b0][M][dv]    box 64 x 32 rows (TMA store)
  int M, Tn, dv;
  int b0, b1;
  int world, rank;
  int m_pairs;   // pairs of 128-row blocks per batch entry
  int n_blocks;  // 128-column blocks per rank segment
  void* O;
  long long ldo, o_sb0, o_sb1;
  int o_vec_ok, o_tma_ok;
  float* lse;  // [Bt][M]: row max + log2(row sum), log2 domain; may be null
  const uint32_t* mask_bits;
  long long mb_ld, mb_sb1;
  float scale_log2;  // scale * log2(e) > 0
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
  long long ws_v_off;  // byte offset of the gathered V region inside a workspace
  const char* push_src[2];
  long long push_ld[2], push_sb0[2], push_sb1[2];
  int push_rows, push_row_bytes[2];
  int dbg;  // bit 0 (harness): P = scale_log2 * S without max / exp / normalisation (linear, isolates the MMAs)
  // Pipeline trace (ddp_attn_debug_trace; null in production): CTA 0 records clock64() of the events of the first
  // kAtTraceSteps steps of its first work item, [actor][step][event]; actors 0-2 = softmax warps (w0 h0 q0),
  // (w0 h1 q0), (w1 h0 q0), actors 3-4 = the MMA issuers of row blocks 0 / 1, actors 5-7 = softmax warps
  // (w0 h0, q = 1, 2, 3).
  long long* trace;
};
constexpr int kAtTraceSteps = 64;
constexpr int kAtTraceEvents = 8;
// Compiled in only with -DDDP_ATTN_TRACE (python build.py trace): the extra live values cost the softmax loop
// registers it does not have (measured: 0.585 -> 0.76 ms for the forward kernel with the stamps compiled in).
__device__ __forceinline__ void attn_trace(long long* buf, bool on, int actor, int step, int ev) {
#ifdef DDP_ATTN_TRACE
  if (buf != nullptr && on && step < kAtTraceSteps)
    buf[(actor * kAtTraceSteps + step) * kAtTraceEvents + ev] = clock64();
#endif
}

template <int DT, int STAGES, bool kRaw = false>
__global__ void __launch_bounds__(kAtThreads, 1) attn_fwd_kernel(const __grid_constant__ AttnKParams p) {
  using T = typename std::conditional<DT == DDP_BF16, __nv_bfloat16, __half>::type;
  static_assert(DT == DDP_BF16 || DT == DDP_F16, "16-bit operands only");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                       // 2 x 16 KB
  uint8_t* sP = sA + 2 * kAtTile;           // 2 x 32 KB
  uint8_t* sKV = sP + 2 * kAtPBytes;        // STAGES x 32 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + STAGES * kAtStageBytes);
  uint64_t* kv_full = bars;                 // [STAGES]
  uint64_t* kv_empty = bars + STAGES;       // [STAGES]
  uint64_t* a_full = bars + 2 * STAGES;     // [1]
  uint64_t* a_empty = a_full + 1;           // [1]
  uint64_t* s_full = a_empty + 1;           // [2]  S_w landed in TMEM
  uint64_t* s_empty = s_full + 2;           // [2]  warpgroup w has read S_w
  uint64_t* p_full = s_empty + 2;           // [4]  half h of P_w staged (and O_{w,h} rescaled); index w*2+h
  uint64_t* o_done = p_full + 4;            // [4]  O_{w,h} += P_{w,h} V_j finished
  uint64_t* pp = o_done + 4;                // [4][2] exp-phase token of lane quadrant q: [q][w] = row block w may run
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pp + 8);
  float2* sStat = reinterpret_cast<float2*>(tmem_slot + 4);  // [2][2][128] (max, sum) of every stream, for the merge

  // dbg bit 8: the four control warps (producer, issuers, pusher) are the LAST warpgroup of the CTA instead of the
  // first (role index = (physical warp + 4) % 20; the softmax warps keep physical warp % 4 = TMEM lane quadrant)
  const int warp = (p.dbg & 8) ? ((threadIdx.x >> 5) + 4) % (kAtThreads / 32) : (threadIdx.x >> 5);
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
    for (int w = 0; w < 2; ++w) {
      mbar_init(&s_full[w], 1);
      mbar_init(&s_empty[w], 8);
    }
    for (int st = 0; st < 4; ++st) {
      mbar_init(&p_full[st], 4);
      mbar_init(&o_done[st], 1);
    }
    for (int i = 0; i < 8; ++i) mbar_init(&pp[i], 2);  // the two warps (column halves) of the other row block
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
    b0i = it % p.b0;
    const int r = it / p.b0;
    const int pair = r % p.m_pairs;
    b1i = r / p.m_pairs;
    m0 = pair * (2 * kAtBM);
  };

  // The softmax threads keep a whole 128-column score row in registers: the producer / MMA / pusher warpgroup
  // hands most of its registers to the two softmax warpgroups (setmaxnreg; the two sides only meet again at the
  // teardown, so the compiler allocates each side with its own budget).
  if (warp < 4) {
  setmaxnreg_dec<kAtRegsLow>();
  if (warp == 0) {
    // ======================================================= TMA producer (whole warp walks the loop)
    int stage = 0;
    uint32_t ph = 0, a_ph = 0;
    const uint32_t* my_flags = p.flags[p.rank];
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int b1i, b0i, m0;
      decode(it, b1i, b0i, m0);
      const int b = b1i * p.b0 + b0i;
      if (lane == 0) {
        mbar_wait(a_empty, a_ph ^ 1, 11);
        mbar_arrive_expect_tx(a_full, 2 * kAtTile);
        tma_load_5d(sA, &p.tmA, a_full, 0, 0, m0, b0i, b1i);
        tma_load_5d(sA + kAtTile, &p.tmA, a_full, 0, 0, m0 + kAtBM, b0i, b1i);
      }
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
          if (lane == 0) {
            if (need_fence) fence_proxy_async();  // peer (generic-proxy) writes -> our TMA (async-proxy) reads
            mbar_wait(&kv_empty[stage], ph ^ 1, 12);
            uint8_t* sb = sKV + stage * kAtStageBytes;
            mbar_arrive_expect_tx(&kv_full[stage], kAtStageBytes);
            const int n0 = nb * kAtBN;
            if (remote) {
              tma_load_5d(sb, &p.tmBg, &kv_full[stage], 0, 0, n0, b, seg);
              tma_load_5d(sb + kAtTile, &p.tmVg, &kv_full[stage], 0, 0, n0, b, seg);
            } else {
              tma_load_5d(sb, &p.tmB, &kv_full[stage], 0, 0, n0, b0i, b1i);
              tma_load_5d(sb + kAtTile, &p.tmV, &kv_full[stage], 0, 0, n0, b0i, b1i);
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
    // ======================================================= MMA issuers (event driven), one per row block
    // A softmax warp releases S_w as soon as its values sit in registers, so S_w(j+1) is computed while the
    // exponentials of step j are still running; the issuing thread polls its three conditions (S_w buffer free +
    // operands landed; half h of P_w staged, h = 0, 1) and issues whatever has become possible.  One issuing
    // thread for BOTH row blocks was the bottleneck of this kernel (~30 dependent instructions per tcgen05.mma,
    // 24 MMAs per step): warp 1 serves row block 0, warp 3 row block 1.
    if (lane == 0) {
      constexpr uint32_t kIdS = umma_idesc(DT, 0u, 0u, kAtBM, kAtBN);  // A K-major, B K-major
      constexpr uint32_t kIdO = umma_idesc(DT, 0u, 1u, kAtBM, 64);     // A = P K-major, B = V MN-major
      const int w = (warp == 1) ? 0 : 1;
      uint32_t a_ph = 0, se_ph = 0, pf_ph[2] = {0u, 0u};
      const uint32_t sKV_u = smem_u32(sKV);
      const uint32_t tS_w = tmem_base + w * kAtBN;
      const uint32_t tO_w = tmem_base + 2 * kAtBN + w * 128;
      // descriptors of the operands that do not move: A_w (4 k-slices) and the two halves of P_w (4 k-slices each)
      uint64_t dA[4], dP[2][4];
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        dA[kk] = umma_smem_desc(smem_u32(sA) + w * kAtTile + kk * 32, 16, 1024);
        dP[0][kk] = umma_smem_desc(smem_u32(sP) + w * kAtPBytes + kk * 32, 16, 1024);
        dP[1][kk] = umma_smem_desc(smem_u32(sP) + w * kAtPBytes + kAtTile + kk * 32, 16, 1024);
      }
      uint32_t g0 = 0;  // steps issued by earlier work items: step j of this item uses ring slot (g0 + j) % STAGES
      // dbg bit 2: the polls below use mbarrier.test_wait (never suspends) instead of try_wait;
      // dbg bit 4: static schedule with blocking waits, in the order the events arrive in steady state
      //            (S(j+1) as soon as the softmax warps hold S(j), then the two P.V products of step j)
      const bool poll_test = (p.dbg & 2) != 0;
      auto poll = [&](uint64_t* bar, uint32_t parity) {
        return poll_test ? mbar_test_wait(bar, parity) : mbar_try_wait(bar, parity);
      };
      if (p.dbg & 4) {
        for (int it = blockIdx.x; it < items; it += gridDim.x) {
          mbar_wait(a_full, a_ph, 14);
          a_ph ^= 1;
          tc_fence_after();
          auto issue_s = [&](int j) {
            const uint32_t g = g0 + j;
            mbar_wait(&kv_full[g % STAGES], (g / STAGES) & 1, 41);
            mbar_wait(&s_empty[w], se_ph ^ 1, 42);
            se_ph ^= 1;
            tc_fence_after();
            const uint64_t db = umma_smem_desc(sKV_u + (g % STAGES) * kAtStageBytes, 16, 1024);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS_w, dA[kk], db + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
            umma_commit(&s_full[w]);
          };
          issue_s(0);
          for (int j = 0; j < nj; ++j) {
            if (j + 1 < nj) issue_s(j + 1);
            const uint32_t g = g0 + j;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              mbar_wait(&p_full[w * 2 + h], pf_ph[h], 43);
              pf_ph[h] ^= 1;
              tc_fence_after();
              const uint64_t dv = umma_smem_desc(
                  sKV_u + (g % STAGES) * kAtStageBytes + kAtTile + h * (4 * 2048), kAtTile, 1024, 2);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                umma_ss<false>(tO_w + h * 64, dP[h][kk], dv + 128 * kk, kIdO, (j | kk) != 0 ? 1u : 0u);
              umma_commit(&o_done[w * 2 + h]);
            }
            umma_commit(&kv_empty[g % STAGES]);
          }
          umma_commit(a_empty);
          g0 += static_cast<uint32_t>(nj);
        }
      } else
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        mbar_wait(a_full, a_ph, 14);
        a_ph ^= 1;
        tc_fence_after();
        int nS = 0, nPV[2] = {0, 0};
        int kvr = 0;       // steps of this item whose operands are known to have landed
        int released = 0;  // steps whose ring slot this issuer has handed back
        uint32_t spins = 0;
        uint64_t t0 = 0;
        while (released < nj) {
          bool progress = false;
          if (nS < nj) {
            const uint32_t g = g0 + nS;
            if (nS >= kvr && poll(&kv_full[g % STAGES], (g / STAGES) & 1)) kvr = nS + 1;
            if (nS < kvr && poll(&s_empty[w], se_ph ^ 1)) {
              se_ph ^= 1;
              tc_fence_after();
              const uint64_t db = umma_smem_desc(sKV_u + (g % STAGES) * kAtStageBytes, 16, 1024);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) umma_ss<false>(tS_w, dA[kk], db + 2 * kk, kIdS, kk != 0 ? 1u : 0u);
              umma_commit(&s_full[w]);
              attn_trace(p.trace, blockIdx.x == 0 && it == 0, 3 + w, nS, 0);
              ++nS;
              progress = true;
            }
          }
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (nPV[h] < nS && poll(&p_full[w * 2 + h], pf_ph[h])) {
              pf_ph[h] ^= 1;
              tc_fence_after();
              const int j = nPV[h];
              const uint32_t g = g0 + j;
              const uint64_t dv = umma_smem_desc(
                  sKV_u + (g % STAGES) * kAtStageBytes + kAtTile + h * (4 * 2048), kAtTile, 1024, 2);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                umma_ss<false>(tO_w + h * 64, dP[h][kk], dv + 128 * kk, kIdO, (j | kk) != 0 ? 1u : 0u);
              umma_commit(&o_done[w * 2 + h]);
              attn_trace(p.trace, blockIdx.x == 0 && it == 0, 3 + w, j, 1 + h);
              nPV[h] = j + 1;
              progress = true;
            }
          }
          // both streams of this row block are through with column block `released`: one of the two arrivals that
          // hand its ring slot back to the producer
          const int lo = min(nPV[0], nPV[1]);
          while (released < lo) {
            umma_commit(&kv_empty[(g0 + released) % STAGES]);
            ++released;
          }
          if (progress) {
            spins = 0;
            t0 = 0;
          } else if ((++spins & 0xFFFu) == 0) {
            // same watchdog as mbar_wait: a broken pipeline must not hang the GPU
            const uint64_t now = global_timer_ns();
            if (t0 == 0) {
              t0 = now;
            } else if (now - t0 > 4000000000ull) {
              atomicMax(&g_ddp_timeout_site, 16u);
              break;
            }
            if (*reinterpret_cast<volatile unsigned int*>(&g_ddp_timeout_site) != 0) break;
          }
        }
        umma_commit(a_empty);
        g0 += static_cast<uint32_t>(nj);
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
    const uint32_t tS = lane_base + w * kAtBN + h * 64;
    const uint32_t tO = lane_base + 2 * kAtBN + st * 64;
    const uint32_t sPw = smem_u32(sP) + w * kAtPBytes;
    const uint32_t sPh = sPw + h * kAtTile + row_in * 128;  // this thread's 128-byte row of P k-chunk h
    const uint32_t sw = lane & 7;
    const uint32_t pair_bar = 1 + w * 4 + q;  // named barrier shared with the warp of the other column half
    uint32_t sf_ph = 0, od_ph = 0;
    // Ping-pong of the exponential phase (dbg bit 32): the four softmax warps of a scheduler (lane quadrant q) are
    // the two column halves of row block 0 and of row block 1.  Left alone they run in lock step - all four in
    // the MUFU-bound exponential phase at once, then all four in the TMEM-load / max / wait phases with the MUFU
    // idle.  The token makes the two row blocks alternate: block w runs its exponentials of step n only after
    // block w^1 has finished those of step n-1 (w = 1) / n (w = 0), so the other block's loads, maxima and waits
    // hide behind them.
    const bool pingpong = (p.dbg & 32) != 0;
    uint32_t pp_n = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      int b1i, b0i, m0;
      decode(it, b1i, b0i, m0);
      const int b = b1i * p.b0 + b0i;
      const int row0 = m0 + w * kAtBM + q * 32;  // first row of this warp
      const int row = row0 + lane;
      const uint32_t* mrow = nullptr;
      if (p.mask_bits != nullptr)
        mrow = p.mask_bits + b1i * p.mb_sb1 + static_cast<long long>(min(row, p.M - 1) >> 5) * p.mb_ld * 32 +
               (min(row, p.M - 1) & 31);  // interleaved layout: the 32 rows of a warp read consecutive words
      float m_used = -INFINITY, l = 0.f;
      // mask words of this row for the 64 columns of step jj (bit set = masked out); fetched ONE STEP AHEAD so
      // that their L2 latency never sits between the arrival of S and the exponentials
      int pf_nb = 0, pf_si = 0;  // cursor of the prefetch (no integer divisions in the step loop)
      auto mask_words_next = [&](uint32_t (&o)[2]) {
        o[0] = 0u;
        o[1] = 0u;
        if (mrow == nullptr || pf_si >= p.world) return;
        int seg = p.rank + pf_si;
        if (seg >= p.world) seg -= p.world;
        const int n0 = pf_nb * kAtBN + h * 64;
        const int nvalid = p.Tn - n0;
        const uint32_t* mp = mrow + ((static_cast<long long>(seg) * p.Tn + n0) >> 5) * 32;
        if (nvalid > 0) o[0] = __ldg(mp);
        if (nvalid > 32) o[1] = __ldg(mp + 32);
        if (++pf_nb == p.n_blocks) {
          pf_nb = 0;
          ++pf_si;
        }
      };
      uint32_t mk_next[2];
      mask_words_next(mk_next);
      int cur_nb = 0;
      for (int j = 0; j < nj; ++j) {
        {
          const int n0 = cur_nb * kAtBN + h * 64;              // first column of this thread's half
          if (++cur_nb == p.n_blocks) cur_nb = 0;
          const int nvalid = max(0, min(64, p.Tn - n0));
          uint32_t mk[2] = {mk_next[0], mk_next[1]};
          mask_words_next(mk_next);
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
          const bool tr_on = p.trace != nullptr && blockIdx.x == 0 && it == 0 && lane == 0 &&
                             ((q == 0 && (w == 0 || h == 0)) || (w == 0 && h == 0));
          const int tr_actor = q == 0 ? w * 2 + h : 4 + q;  // 0, 1, 2 and 5, 6, 7 = (w0 h0, q = 1, 2, 3)
          attn_trace(p.trace, tr_on, tr_actor, j, 0);
          mbar_wait(&s_full[w], sf_ph, 17);
          sf_ph ^= 1;
          tc_fence_after();
          attn_trace(p.trace, tr_on, tr_actor, j, 1);
          // ---- this thread's half row: one TMEM load, then S_w goes straight back to the MMA warp
          uint32_t r[64];
          tmem_ld_32x64(tS, r);
          tmem_ld_wait();
          tmem_ld_pin64(r);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&s_empty[w]);
          attn_trace(p.trace, tr_on, tr_actor, j, 2);
          // ---- maximum of the raw scores over the live columns (four independent chains)
          float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            if (mk[c] == 0u) {
#pragma unroll
              for (int e = 0; e < 32; e += 4) {
                mx0 = fmaxf(mx0, __uint_as_float(r[c * 32 + e]));
                mx1 = fmaxf(mx1, __uint_as_float(r[c * 32 + e + 1]));
                mx2 = fmaxf(mx2, __uint_as_float(r[c * 32 + e + 2]));
                mx3 = fmaxf(mx3, __uint_as_float(r[c * 32 + e + 3]));
              }
            } else {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                if (!((mk[c] >> e) & 1u)) mx0 = fmaxf(mx0, __uint_as_float(r[c * 32 + e]));
            }
          }
          const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;  // scale_log2 > 0
          // Lazy rescale: the reference point of the exponentials only moves when the maximum grew by more than
          // 2^kAtTau, so P stays <= 2^kAtTau and the O accumulator is rarely touched.
          const bool need = !kRaw && (m_blk > m_used + kAtTau);
          attn_trace(p.trace, tr_on, tr_actor, j, 3);
          if (j > 0) {
            // O_st += P V_{j-1} has finished: our P rows may be overwritten, O_st may be rescaled
            mbar_wait(&o_done[st], od_ph, 18);
            od_ph ^= 1;
            tc_fence_after();
            attn_trace(p.trace, tr_on, tr_actor, j, 4);
            if (__any_sync(0xffffffffu, need)) {
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
              tmem_st_wait();
              l *= f;
            }
          }
          if (need) m_used = m_blk;
          const float neg_m = (kRaw || m_used == -INFINITY) ? 0.f : -m_used;
          if (pingpong) mbar_wait(&pp[q * 2 + w], w == 0 ? ((pp_n & 1u) ^ 1u) : (pp_n & 1u), 50);
          // ---- P = exp2(scale_log2 * S - m) -> bf16 row in shared memory (K-major, 128B swizzle)
          float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float v[32];
            if (kRaw) {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                v[e] = ((mk[c] >> e) & 1u) ? 0.f : __uint_as_float(r[c * 32 + e]) * p.scale_log2;
            } else if (mk[c] == 0u) {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                v[e] = ex2_approx(fmaf(__uint_as_float(r[c * 32 + e]), p.scale_log2, neg_m));
            } else {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                v[e] = ((mk[c] >> e) & 1u) ? 0.f
                                           : ex2_approx(fmaf(__uint_as_float(r[c * 32 + e]), p.scale_log2, neg_m));
            }
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              l0 += v[e];
              l1 += v[e + 1];
              l2 += v[e + 2];
              l3 += v[e + 3];
            }
#pragma unroll
            for (int jv = 0; jv < 4; ++jv) {
              uint4 qv;
              qv.x = pack2<T>(v[8 * jv + 0], v[8 * jv + 1]);
              qv.y = pack2<T>(v[8 * jv + 2], v[8 * jv + 3]);
              qv.z = pack2<T>(v[8 * jv + 4], v[8 * jv + 5]);
              qv.w = pack2<T>(v[8 * jv + 6], v[8 * jv + 7]);
              st_shared_v4(sPh + (((c * 4 + jv) ^ sw) * 16), qv);
            }
          }
          l += (l0 + l1) + (l2 + l3);
          if (pingpong) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&pp[q * 2 + (w ^ 1)]);
            ++pp_n;
          }
          attn_trace(p.trace, tr_on, tr_actor, j, 5);
          tc_fence_before();           // a rescale of O_st (TMEM stores) precedes the MMA released below
          fence_proxy_async_smem();    // P (generic-proxy stores) -> tensor core (async proxy)
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_full[st]);
          attn_trace(p.trace, tr_on, tr_actor, j, 6);
        }
      }
      // ---- row block finished: merge the two column-half streams of every row, O = (Oa fa + Ob fb) / l
      mbar_wait(&o_done[st], od_ph, 19);
      od_ph ^= 1;
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
      const uint32_t sbuf = sPw + q * 4096;  // 32 rows of P k-chunk 0 double as the staging tile of (w, q)
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
            bulk_wait_group_read<0>();  // the staging rows are P rows of the next work item
          }
          __syncwarp();
        } else {
          T* dst = reinterpret_cast<T*>(p.O) + b1i * p.o_sb1 + b0i * p.o_sb0 + static_cast<long long>(row0) * p.ldo;
          store_rows_generic<T>(sbuf, dst, p.ldo, rows_valid, p.dv, p.o_vec_ok != 0, lane);
          __syncwarp();
        }
        if (p.lse != nullptr && row < p.M) p.lse[static_cast<size_t>(b) * p.M + row] = m_all + log2f(l_all);
      }
      // the partner must not overwrite the staging rows (its own P rows live in k-chunk 1, but the NEXT staging
      // pass writes k-chunk 0 again) before the store above has read them
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
