// Standalone validation + timing harness for ddp_attn_fwd (no Python).  Runs on ONE GPU; multi-rank cases are
// emulated like in gemm_harness.cu: every "rank" owns its workspace / flags on the same GPU, the produce phase of
// all ranks runs before the consume phase.
// Usage: attn_harness [quick|full|perf|perfall|bwd|bwdfull|multi|prof|one dt W b1 b0 M Tn d dv il maskmode chunk ramp dbg]
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ddp_b200.h"

#define CK(x)                                                                                 \
  do {                                                                                        \
    cudaError_t e_ = (x);                                                                     \
    if (e_ != cudaSuccess) {                                                                  \
      printf("CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                                \
    }                                                                                         \
  } while (0)

static uint32_t g_seed = 777;
static int irand(int lo, int hi) {
  g_seed = g_seed * 1664525u + 1013904223u;
  return lo + (int)((g_seed >> 8) % (uint32_t)(hi - lo + 1));
}

static void to_dev(const std::vector<float>& h, int dt, void* d) {
  if (dt == DDP_BF16) {
    std::vector<__nv_bfloat16> t(h.size());
    for (size_t i = 0; i < h.size(); ++i) t[i] = __float2bfloat16(h[i]);
    CK(cudaMemcpy(d, t.data(), t.size() * 2, cudaMemcpyHostToDevice));
  } else {
    std::vector<__half> t(h.size());
    for (size_t i = 0; i < h.size(); ++i) t[i] = __float2half(h[i]);
    CK(cudaMemcpy(d, t.data(), t.size() * 2, cudaMemcpyHostToDevice));
  }
}
static std::vector<float> from_dev(const void* d, size_t n, int dt) {
  std::vector<float> h(n);
  if (dt == DDP_BF16) {
    std::vector<__nv_bfloat16> t(n);
    CK(cudaMemcpy(t.data(), d, n * 2, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; ++i) h[i] = __bfloat162float(t[i]);
  } else {
    std::vector<__half> t(n);
    CK(cudaMemcpy(t.data(), d, n * 2, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; ++i) h[i] = __half2float(t[i]);
  }
  return h;
}

// Host tensor [b1][b0][rows][cols]; "interleave" = heads inside the row, like a [B, T, H*d] projection output.
struct HT {
  int b1, b0, rows, cols;
  long long ld, sb0, sb1;
  std::vector<float> v;
  size_t size() const { return v.size(); }
  float& at(int i1, int i0, int r, int c) { return v[i1 * sb1 + i0 * sb0 + (long long)r * ld + c]; }
  float at(int i1, int i0, int r, int c) const { return v[i1 * sb1 + i0 * sb0 + (long long)r * ld + c]; }
};
static HT make_ht(int b1, int b0, int rows, int cols, bool interleave, bool fill, float mul = 0.25f) {
  HT t;
  t.b1 = b1; t.b0 = b0; t.rows = rows; t.cols = cols;
  if (interleave) {
    t.sb0 = cols; t.ld = (long long)b0 * cols; t.sb1 = t.ld * rows;
  } else {
    t.ld = cols; t.sb0 = t.ld * rows; t.sb1 = t.sb0 * b0;
  }
  t.v.assign((size_t)t.sb1 * b1, 0.f);
  if (fill)
    for (auto& x : t.v) x = irand(-8, 8) * mul;
  return t;
}
static ddp_tensor dview(const HT& t, void* d, int dt) {
  ddp_tensor x;
  memset(&x, 0, sizeof x);
  x.ptr = d; x.rows = t.rows; x.cols = t.cols; x.ld = t.ld; x.b0 = t.b0; x.b1 = t.b1; x.sb0 = t.sb0; x.sb1 = t.sb1;
  x.dtype = dt;
  return x;
}

struct Case {
  int dt, world, b1, b0, M, Tn, d, dv;
  bool interleave;
  int maskmode;  // 0 none, 1 random 30 %, 2 block pattern with fully masked rows / tiles
  int chunk;
  int ramp;      // 1: later columns get larger scores (forces the lazy rescale of the O accumulator)
};

static int g_fail = 0, g_pass = 0;

static bool run_case(const Case& c, int dbg, bool verbose) {
  const int W = c.world, T = W * c.Tn;
  const float scale = 1.0f / sqrtf((float)c.d);
  std::vector<HT> A(W), B(W), V(W), O(W);
  std::vector<std::vector<uint8_t>> mask(W);
  for (int r = 0; r < W; ++r) {
    A[r] = make_ht(c.b1, c.b0, c.M, c.d, c.interleave, true);
    B[r] = make_ht(c.b1, c.b0, c.Tn, c.d, c.interleave, true);
    V[r] = make_ht(c.b1, c.b0, c.Tn, c.dv, c.interleave, true);
    O[r] = make_ht(c.b1, c.b0, c.M, c.dv, c.interleave, false);
    if (c.ramp)
      for (int i1 = 0; i1 < c.b1; ++i1)
        for (int i0 = 0; i0 < c.b0; ++i0)
          for (int j = 0; j < c.Tn; ++j) {
            const int gcol = r * c.Tn + j;
            const float f = (gcol * 4 / T) >= 2 ? 4.f : ((gcol * 4 / T) == 1 ? 2.f : 1.f);
            for (int k = 0; k < c.d; ++k) B[r].at(i1, i0, j, k) *= f;
          }
    if (c.maskmode) {
      mask[r].assign((size_t)c.b1 * c.M * T, 0);
      for (int i1 = 0; i1 < c.b1; ++i1)
        for (int m = 0; m < c.M; ++m)
          for (int t = 0; t < T; ++t) {
            uint8_t v = 0;
            if (c.maskmode == 1) v = irand(0, 9) < 3;
            if (c.maskmode == 2) v = (m % 37 == 5) || ((t / 128) % 3 == 1 && (m / 32) % 2 == 0) || (t > m + r * c.M);
            mask[r][((size_t)i1 * c.M + m) * T + t] = v;
          }
    }
  }
  std::vector<void*> dA(W), dB(W), dV(W), dO(W), dMask(W, nullptr), dBits(W, nullptr), dLse(W);
  const long long ldw = (T + 31) / 32;
  for (int r = 0; r < W; ++r) {
    CK(cudaMalloc(&dA[r], A[r].size() * 2));
    CK(cudaMalloc(&dB[r], B[r].size() * 2));
    CK(cudaMalloc(&dV[r], V[r].size() * 2));
    CK(cudaMalloc(&dO[r], O[r].size() * 2));
    CK(cudaMalloc(&dLse[r], (size_t)c.b1 * c.b0 * c.M * 4));
    to_dev(A[r].v, c.dt, dA[r]);
    to_dev(B[r].v, c.dt, dB[r]);
    to_dev(V[r].v, c.dt, dV[r]);
    CK(cudaMemset(dO[r], 0xFF, O[r].size() * 2));
    CK(cudaMemset(dLse[r], 0xFF, (size_t)c.b1 * c.b0 * c.M * 4));
    if (c.maskmode) {
      CK(cudaMalloc(&dMask[r], mask[r].size()));
      CK(cudaMemcpy(dMask[r], mask[r].data(), mask[r].size(), cudaMemcpyHostToDevice));
      CK(cudaMalloc(&dBits[r], ddp_attn_mask_words(c.b1, c.M, T, 0) * 4));
      if (ddp_attn_pack_mask((const uint8_t*)dMask[r], c.b1, c.M, T, T, (long long)c.M * T, (uint32_t*)dBits[r], 0, 0)) {
        printf("pack_mask failed: %s\n", ddp_last_error());
        return false;
      }
    }
  }
  std::vector<void*> ws(W, nullptr);
  std::vector<uint32_t*> flags(W, nullptr), done(W, nullptr), exitc(W, nullptr);
  if (W > 1) {
    size_t wsb = 0, nf = 0;
    ddp_tensor b0 = dview(B[0], dB[0], c.dt), v0 = dview(V[0], dV[0], c.dt);
    ddp_attn_fwd_workspace(&b0, &v0, W, c.chunk, &wsb, &nf);
    for (int r = 0; r < W; ++r) {
      CK(cudaMalloc(&ws[r], wsb + 256));
      CK(cudaMemset(ws[r], 0xFF, wsb + 256));
      CK(cudaMalloc((void**)&flags[r], nf * 4 + 256));
      CK(cudaMemset(flags[r], 0, nf * 4 + 256));
      CK(cudaMalloc((void**)&done[r], 256));
      CK(cudaMemset(done[r], 0, 256));
      CK(cudaMalloc((void**)&exitc[r], 256));
      CK(cudaMemset(exitc[r], 0, 256));
    }
  }
  bool ok = true;
  const int nseq = (W > 1) ? 2 : 1;
  for (int s = 1; s <= nseq && ok; ++s) {
    for (int phase = (W > 1 ? 1 : 3); phase <= (W > 1 ? 2 : 3); ++phase) {
      for (int r = 0; r < W; ++r) {
        ddp_comm cm;
        memset(&cm, 0, sizeof cm);
        cm.world = W; cm.rank = r; cm.seq = (uint32_t)s; cm.phase_mask = phase; cm.chunk_rows = c.chunk;
        for (int q = 0; q < W; ++q) {
          cm.ws[q] = ws[q]; cm.flags[q] = flags[q]; cm.done[q] = done[q];
        }
        cm.exit_counter = exitc[r];
        ddp_tensor a = dview(A[r], dA[r], c.dt), b = dview(B[r], dB[r], c.dt), v = dview(V[r], dV[r], c.dt),
                   o = dview(O[r], dO[r], c.dt);
        int e = ddp_attn_fwd(&a, &b, &v, &o, (float*)dLse[r], (const uint32_t*)dBits[r], ldw,
                             (long long)(ddp_attn_mask_words(c.b1, c.M, T, 0) / c.b1), scale, W > 1 ? &cm : nullptr, dbg, 0);
        if (e) {
          printf("  ddp_attn_fwd error %d: %s\n", e, ddp_last_error());
          ok = false;
          break;
        }
      }
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("  kernel fault: %s\n", cudaGetErrorString(e));
        exit(3);
      }
      if (!ok) break;
    }
  }
  const int site = ddp_debug_timeout_site_attn();
  if (site) {
    printf("  pipeline watchdog fired at site %d\n", site);
    ok = false;
  }
  double max_err = 0, max_lse_err = 0;
  long long nbad = 0, nshown = 0, nbad_lse = 0;
  long long bad_wq[2][4] = {{0}}, bad_cg[2] = {0};
  const bool raw = dbg & 1;
  if (ok) {
    const double sl2 = (double)scale * 1.4426950408889634;
    std::vector<double> s(T), pj(T);
    for (int r = 0; r < W; ++r) {
      std::vector<float> got = from_dev(dO[r], O[r].size(), c.dt);
      std::vector<float> lse((size_t)c.b1 * c.b0 * c.M);
      CK(cudaMemcpy(lse.data(), dLse[r], lse.size() * 4, cudaMemcpyDeviceToHost));
      for (int i1 = 0; i1 < c.b1; ++i1)
        for (int i0 = 0; i0 < c.b0; ++i0)
          for (int m = 0; m < c.M; ++m) {
            double mx = -INFINITY;
            for (int t = 0; t < T; ++t) {
              const int src = t / c.Tn, j = t % c.Tn;
              double acc = 0;
              for (int k = 0; k < c.d; ++k) acc += (double)A[r].at(i1, i0, m, k) * B[src].at(i1, i0, j, k);
              const bool dead = c.maskmode && mask[r][((size_t)i1 * c.M + m) * T + t];
              s[t] = dead ? -INFINITY : acc * sl2;
              if (s[t] > mx) mx = s[t];
            }
            double l = 0;
            for (int t = 0; t < T; ++t) {
              pj[t] = raw ? (s[t] == -INFINITY ? 0.0 : s[t]) : (mx == -INFINITY ? 0.0 : exp2(s[t] - mx));
              l += pj[t];
            }
            const bool nan_row = !raw && (mx == -INFINITY);
            for (int n = 0; n < c.dv; ++n) {
              double ref = 0, mag = 0;
              for (int t = 0; t < T; ++t) {
                const double pv = pj[t] * V[t / c.Tn].at(i1, i0, t % c.Tn, n);
                ref += pv;
                mag += fabs(pv);
              }
              if (!raw) { ref /= l; mag /= l; }
              const float g = got[i1 * O[r].sb1 + i0 * O[r].sb0 + (long long)m * O[r].ld + n];
              bool bad;
              double err = 0;
              if (nan_row) {
                bad = (g == g);  // must be NaN
              } else {
                err = fabs((double)g - ref);
                bad = !(err <= 0.01 * mag + 0.01 * fabs(ref) + 2e-3);
                if (err == err && err > max_err) max_err = err;
              }
              if (bad) {
                ++nbad;
                bad_wq[(m / 128) & 1][(m % 128) / 32]++;
                bad_cg[(n / 32) & 1]++;
                if (verbose && nshown < 8) {
                  printf("    bad rank %d b(%d,%d) m %d n %d: got %g want %g%s\n", r, i1, i0, m, n, g, ref,
                         nan_row ? " (NaN row)" : "");
                  ++nshown;
                }
              }
            }
            if (!raw && !nan_row) {
              const double lref = mx + log2(l);
              const double e = fabs((double)lse[((size_t)i1 * c.b0 + i0) * c.M + m] - lref);
              if (!(e <= 2e-2)) {
                ++nbad_lse;
                if (verbose && nbad_lse <= 4)
                  printf("    bad lse rank %d b(%d,%d) m %d: got %g want %g\n", r, i1, i0, m,
                         lse[((size_t)i1 * c.b0 + i0) * c.M + m], lref);
              }
              if (e == e && e > max_lse_err) max_lse_err = e;
            }
          }
    }
    if (nbad || nbad_lse) ok = false;
  }
  printf("%s ATTN %s W=%d b=(%d,%d) M=%d Tn=%d d=%d dv=%d il=%d mask=%d chunk=%d ramp=%d dbg=%d  max_err=%.3g bad=%lld "
         "lse_err=%.3g bad_lse=%lld\n",
         ok ? "PASS" : "FAIL", c.dt == DDP_BF16 ? "bf16" : "fp16", W, c.b1, c.b0, c.M, c.Tn, c.d, c.dv, (int)c.interleave,
         c.maskmode, c.chunk, c.ramp, dbg, max_err, nbad, max_lse_err, nbad_lse);
  if (nbad)
    printf("    bad by (row block, quadrant): [%lld %lld %lld %lld] [%lld %lld %lld %lld]  by column half: %lld %lld\n",
           bad_wq[0][0], bad_wq[0][1], bad_wq[0][2], bad_wq[0][3], bad_wq[1][0], bad_wq[1][1], bad_wq[1][2], bad_wq[1][3],
           bad_cg[0], bad_cg[1]);
  fflush(stdout);
  for (int r = 0; r < W; ++r) {
    cudaFree(dA[r]); cudaFree(dB[r]); cudaFree(dV[r]); cudaFree(dO[r]); cudaFree(dLse[r]);
    if (dMask[r]) cudaFree(dMask[r]);
    if (dBits[r]) cudaFree(dBits[r]);
    if (W > 1) { cudaFree(ws[r]); cudaFree(flags[r]); cudaFree(done[r]); cudaFree(exitc[r]); }
  }
  if (ok) ++g_pass; else ++g_fail;
  return ok;
}


// ---------------------------------------------------------------------------------------------- backward
// Validates ddp_attn_bwd_rows / ddp_attn_bwd_cols against a double-precision reference.  The row statistics
// (-L, -scale*delta) come from the CPU reference, so the two kernels are tested independently of the forward.
static bool run_bwd_case(const Case& c, bool verbose, int dbg = 0) {
  const int W = c.world, T = W * c.Tn;
  const float scale = 1.0f / sqrtf((float)c.d);
  const double sl2 = (double)scale * 1.4426950408889634;
  const int Bt = c.b1 * c.b0;
  const long long stat_ld = (c.M + 63) / 64 * 64;
  std::vector<HT> A(W), B(W), V(W), G(W), dA(W), dB(W), dV(W);
  std::vector<std::vector<uint8_t>> mask(W);
  for (int r = 0; r < W; ++r) {
    A[r] = make_ht(c.b1, c.b0, c.M, c.d, c.interleave, true);
    B[r] = make_ht(c.b1, c.b0, c.Tn, c.d, c.interleave, true);
    V[r] = make_ht(c.b1, c.b0, c.Tn, c.dv, c.interleave, true);
    G[r] = make_ht(c.b1, c.b0, c.M, c.dv, c.interleave, true);
    dA[r] = make_ht(c.b1, c.b0, c.M, c.d, c.interleave, false);
    dB[r] = make_ht(c.b1, c.b0, c.Tn, c.d, c.interleave, false);
    dV[r] = make_ht(c.b1, c.b0, c.Tn, c.dv, c.interleave, false);
    if (c.maskmode) {
      mask[r].assign((size_t)c.b1 * c.M * T, 0);
      for (int i1 = 0; i1 < c.b1; ++i1)
        for (int m = 0; m < c.M; ++m)
          for (int t = 0; t < T; ++t) {
            uint8_t v = 0;
            if (c.maskmode == 1) v = irand(0, 9) < 3;
            if (c.maskmode == 2) v = (m % 37 == 5) || ((t / 128) % 3 == 1 && (m / 32) % 2 == 0) || (t > m + r * c.M);
            mask[r][((size_t)i1 * c.M + m) * T + t] = v;
          }
    }
  }
  // ---- reference
  std::vector<std::vector<double>> rA(W), rB(W), rV(W), gA(W), gB(W), gV(W);  // values and magnitudes
  std::vector<std::vector<float>> neg_lse(W), nds(W);
  for (int r = 0; r < W; ++r) {
    rA[r].assign((size_t)Bt * c.M * c.d, 0); gA[r] = rA[r];
    rB[r].assign((size_t)Bt * c.Tn * c.d, 0); gB[r] = rB[r];
    rV[r].assign((size_t)Bt * c.Tn * c.dv, 0); gV[r] = rV[r];
    neg_lse[r].assign((size_t)Bt * stat_ld, 0.f);
    nds[r].assign((size_t)Bt * stat_ld, 0.f);
  }
  std::vector<double> s(T), pj(T), dsv(T), orow(c.dv);
  for (int r = 0; r < W; ++r)
    for (int i1 = 0; i1 < c.b1; ++i1)
      for (int i0 = 0; i0 < c.b0; ++i0) {
        const int b = i1 * c.b0 + i0;
        for (int m = 0; m < c.M; ++m) {
          double mx = -INFINITY;
          for (int t = 0; t < T; ++t) {
            const int src = t / c.Tn, j = t % c.Tn;
            double acc = 0;
            for (int k = 0; k < c.d; ++k) acc += (double)A[r].at(i1, i0, m, k) * B[src].at(i1, i0, j, k);
            const bool dead = c.maskmode && mask[r][((size_t)i1 * c.M + m) * T + t];
            s[t] = dead ? -INFINITY : acc * sl2;
            if (s[t] > mx) mx = s[t];
          }
          if (mx == -INFINITY) continue;  // fully masked row: contributes nothing (statistics stay 0)
          double l = 0;
          for (int t = 0; t < T; ++t) {
            pj[t] = exp2(s[t] - mx);
            l += pj[t];
          }
          for (int n = 0; n < c.dv; ++n) orow[n] = 0;
          for (int t = 0; t < T; ++t) {
            pj[t] /= l;
            for (int n = 0; n < c.dv; ++n) orow[n] += pj[t] * V[t / c.Tn].at(i1, i0, t % c.Tn, n);
          }
          double delta = 0;
          for (int n = 0; n < c.dv; ++n) delta += (double)G[r].at(i1, i0, m, n) * orow[n];
          neg_lse[r][(size_t)b * stat_ld + m] = (float)-(mx + log2(l));
          nds[r][(size_t)b * stat_ld + m] = (float)(-(double)scale * delta);
          for (int t = 0; t < T; ++t) {
            const int src = t / c.Tn, j = t % c.Tn;
            double dp = 0;
            for (int n = 0; n < c.dv; ++n) dp += (double)G[r].at(i1, i0, m, n) * V[src].at(i1, i0, j, n);
            const double ds = scale * pj[t] * (dp - delta);
            dsv[t] = ds;
            for (int k = 0; k < c.d; ++k) {
              const double ta = ds * B[src].at(i1, i0, j, k), tb = ds * A[r].at(i1, i0, m, k);
              rA[r][((size_t)b * c.M + m) * c.d + k] += ta;
              gA[r][((size_t)b * c.M + m) * c.d + k] += fabs(ta);
              rB[src][((size_t)b * c.Tn + j) * c.d + k] += tb;
              gB[src][((size_t)b * c.Tn + j) * c.d + k] += fabs(tb);
            }
            for (int n = 0; n < c.dv; ++n) {
              const double tv = pj[t] * G[r].at(i1, i0, m, n);
              rV[src][((size_t)b * c.Tn + j) * c.dv + n] += tv;
              gV[src][((size_t)b * c.Tn + j) * c.dv + n] += fabs(tv);
            }
          }
        }
      }
  // ---- device
  std::vector<void*> dAp(W), dBp(W), dVp(W), dGp(W), ddA(W), ddB(W), ddV(W), dMask(W, nullptr), dBits(W, nullptr),
      dBitsT(W, nullptr), dNl(W), dNd(W);
  const long long ldw = (T + 31) / 32, ldwt = (c.M + 31) / 32;
  for (int r = 0; r < W; ++r) {
    CK(cudaMalloc(&dAp[r], A[r].size() * 2)); to_dev(A[r].v, c.dt, dAp[r]);
    CK(cudaMalloc(&dBp[r], B[r].size() * 2)); to_dev(B[r].v, c.dt, dBp[r]);
    CK(cudaMalloc(&dVp[r], V[r].size() * 2)); to_dev(V[r].v, c.dt, dVp[r]);
    CK(cudaMalloc(&dGp[r], G[r].size() * 2)); to_dev(G[r].v, c.dt, dGp[r]);
    CK(cudaMalloc(&ddA[r], dA[r].size() * 2)); CK(cudaMemset(ddA[r], 0xFF, dA[r].size() * 2));
    CK(cudaMalloc(&ddB[r], dB[r].size() * 2)); CK(cudaMemset(ddB[r], 0xFF, dB[r].size() * 2));
    CK(cudaMalloc(&ddV[r], dV[r].size() * 2)); CK(cudaMemset(ddV[r], 0xFF, dV[r].size() * 2));
    CK(cudaMalloc(&dNl[r], neg_lse[r].size() * 4 + 256));
    CK(cudaMemcpy(dNl[r], neg_lse[r].data(), neg_lse[r].size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dNd[r], nds[r].size() * 4 + 256));
    CK(cudaMemcpy(dNd[r], nds[r].data(), nds[r].size() * 4, cudaMemcpyHostToDevice));
    if (c.maskmode) {
      CK(cudaMalloc(&dMask[r], mask[r].size()));
      CK(cudaMemcpy(dMask[r], mask[r].data(), mask[r].size(), cudaMemcpyHostToDevice));
      CK(cudaMalloc(&dBits[r], ddp_attn_mask_words(c.b1, c.M, T, 0) * 4));
      CK(cudaMalloc(&dBitsT[r], ddp_attn_mask_words(c.b1, c.M, T, 1) * 4));
      if (ddp_attn_pack_mask((const uint8_t*)dMask[r], c.b1, c.M, T, T, (long long)c.M * T, (uint32_t*)dBits[r], 0, 0) ||
          ddp_attn_pack_mask((const uint8_t*)dMask[r], c.b1, c.M, T, T, (long long)c.M * T, (uint32_t*)dBitsT[r], 1, 0)) {
        printf("pack_mask failed: %s\n", ddp_last_error());
        return false;
      }
    }
  }
  std::vector<void*> ws(W, nullptr);
  std::vector<uint32_t*> flags(W, nullptr), done(W, nullptr), exitc(W, nullptr);
  if (W > 1) {
    size_t wsb = 0, nf = 0, wsb2 = 0, nf2 = 0;
    ddp_tensor b0 = dview(B[0], dBp[0], c.dt), v0 = dview(V[0], dVp[0], c.dt);
    ddp_attn_fwd_workspace(&b0, &v0, W, c.chunk, &wsb, &nf);
    ddp_attn_bwd_cols_workspace(&b0, &v0, W, c.chunk, &wsb2, &nf2);
    if (wsb2 > wsb) wsb = wsb2;
    if (nf2 > nf) nf = nf2;
    for (int r = 0; r < W; ++r) {
      CK(cudaMalloc(&ws[r], wsb + 256)); CK(cudaMemset(ws[r], 0xFF, wsb + 256));
      CK(cudaMalloc((void**)&flags[r], nf * 4 + 256)); CK(cudaMemset(flags[r], 0, nf * 4 + 256));
      CK(cudaMalloc((void**)&done[r], 256)); CK(cudaMemset(done[r], 0, 256));
      CK(cudaMalloc((void**)&exitc[r], 256)); CK(cudaMemset(exitc[r], 0, 256));
    }
  }
  bool ok = true;
  uint32_t seq = 0;
  for (int kernel = 0; kernel < 2 && ok; ++kernel) {
    ++seq;
    const int phases_rows[2] = {1, 2}, phases_cols[3] = {1, 2, 4};
    const int nph = (W > 1) ? (kernel ? 3 : 2) : 1;
    for (int pi = 0; pi < nph && ok; ++pi) {
      const int phase = (W > 1) ? (kernel ? phases_cols[pi] : phases_rows[pi]) : 3;
      for (int r = 0; r < W; ++r) {
        ddp_comm cm;
        memset(&cm, 0, sizeof cm);
        cm.world = W; cm.rank = r; cm.seq = seq; cm.phase_mask = phase; cm.chunk_rows = c.chunk;
        for (int q = 0; q < W; ++q) {
          cm.ws[q] = ws[q]; cm.flags[q] = flags[q]; cm.done[q] = done[q];
        }
        cm.exit_counter = exitc[r];
        ddp_tensor a = dview(A[r], dAp[r], c.dt), b = dview(B[r], dBp[r], c.dt), v = dview(V[r], dVp[r], c.dt),
                   g = dview(G[r], dGp[r], c.dt), oa = dview(dA[r], ddA[r], c.dt), ob = dview(dB[r], ddB[r], c.dt),
                   ov = dview(dV[r], ddV[r], c.dt);
        int e;
        if (kernel == 0)
          e = ddp_attn_bwd_rows(&a, &g, &b, &v, &oa, (const float*)dNl[r], (const float*)dNd[r], stat_ld,
                                (const uint32_t*)dBits[r], ldw, (long long)(ddp_attn_mask_words(c.b1, c.M, T, 0) / c.b1), scale,
                                W > 1 ? &cm : nullptr, 0, 0);
        else
          e = ddp_attn_bwd_cols(&a, &g, &b, &v, &ov, &ob, (const float*)dNl[r], (const float*)dNd[r], stat_ld,
                                (const uint32_t*)dBitsT[r], ldwt, (long long)(ddp_attn_mask_words(c.b1, c.M, T, 1) / c.b1), scale,
                                W > 1 ? &cm : nullptr, 0, 0);
        if (e) {
          printf("  backward kernel %d error %d: %s\n", kernel, e, ddp_last_error());
          ok = false;
          break;
        }
      }
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("  kernel fault (kernel %d phase %d): %s\n", kernel, phase, cudaGetErrorString(e));
        exit(3);
      }
    }
    const int site = ddp_debug_timeout_site_attn();
    if (site) {
      printf("  pipeline watchdog fired at site %d (kernel %d)\n", site, kernel);
      ok = false;
    }
  }
  // ---- compare
  double maxerr[3] = {0, 0, 0};
  long long nbad[3] = {0, 0, 0};
  if (ok) {
    for (int r = 0; r < W; ++r) {
      for (int which = 0; which < 3; ++which) {
        const HT& t = which == 0 ? dA[r] : which == 1 ? dB[r] : dV[r];
        const void* dp = which == 0 ? ddA[r] : which == 1 ? ddB[r] : ddV[r];
        const std::vector<double>& ref = which == 0 ? rA[r] : which == 1 ? rB[r] : rV[r];
        const std::vector<double>& mag = which == 0 ? gA[r] : which == 1 ? gB[r] : gV[r];
        std::vector<float> got = from_dev(dp, t.size(), c.dt);
        int shown = 0;
        for (int i1 = 0; i1 < c.b1; ++i1)
          for (int i0 = 0; i0 < c.b0; ++i0)
            for (int m = 0; m < t.rows; ++m)
              for (int n = 0; n < t.cols; ++n) {
                const size_t ri = ((size_t)(i1 * c.b0 + i0) * t.rows + m) * t.cols + n;
                const float g = got[i1 * t.sb1 + i0 * t.sb0 + (long long)m * t.ld + n];
                const double err = fabs((double)g - ref[ri]);
                const bool bad = !(err <= 0.01 * mag[ri] + 0.01 * fabs(ref[ri]) + 2e-3);
                if (err == err && err > maxerr[which]) maxerr[which] = err;
                if (bad) {
                  ++nbad[which];
                  if (verbose && shown < 5) {
                    printf("    bad %s rank %d b(%d,%d) row %d col %d: got %g want %g\n",
                           which == 0 ? "dA" : which == 1 ? "dB" : "dV", r, i1, i0, m, n, g, ref[ri]);
                    ++shown;
                  }
                }
              }
      }
    }
    if (nbad[0] || nbad[1] || nbad[2]) ok = false;
  }
  printf("%s ATTN-BWD %s W=%d b=(%d,%d) M=%d Tn=%d d=%d dv=%d il=%d mask=%d chunk=%d  dA err=%.3g bad=%lld  dB err=%.3g "
         "bad=%lld  dV err=%.3g bad=%lld\n",
         ok ? "PASS" : "FAIL", c.dt == DDP_BF16 ? "bf16" : "fp16", W, c.b1, c.b0, c.M, c.Tn, c.d, c.dv, (int)c.interleave,
         c.maskmode, c.chunk, maxerr[0], nbad[0], maxerr[1], nbad[1], maxerr[2], nbad[2]);
  fflush(stdout);
  for (int r = 0; r < W; ++r) {
    cudaFree(dAp[r]); cudaFree(dBp[r]); cudaFree(dVp[r]); cudaFree(dGp[r]); cudaFree(ddA[r]); cudaFree(ddB[r]);
    cudaFree(ddV[r]); cudaFree(dNl[r]); cudaFree(dNd[r]);
    if (dMask[r]) cudaFree(dMask[r]);
    if (dBits[r]) cudaFree(dBits[r]);
    if (dBitsT[r]) cudaFree(dBitsT[r]);
    if (W > 1) { cudaFree(ws[r]); cudaFree(flags[r]); cudaFree(done[r]); cudaFree(exitc[r]); }
  }
  if (ok) ++g_pass; else ++g_fail;
  return ok;
}

static void perf_bwd(int dt, int H, int M, int Tn, int d, int iters, bool with_mask = false) {
  HT A = make_ht(1, H, M, d, true, false), B = make_ht(1, H, Tn, d, true, false);
  void *dA_, *dB_, *dV_, *dG_, *oA, *oB, *oV;
  float *nl, *nd;
  CK(cudaMalloc(&dA_, A.size() * 2)); CK(cudaMalloc(&dG_, A.size() * 2)); CK(cudaMalloc(&oA, A.size() * 2));
  CK(cudaMalloc(&dB_, B.size() * 2)); CK(cudaMalloc(&dV_, B.size() * 2)); CK(cudaMalloc(&oB, B.size() * 2));
  CK(cudaMalloc(&oV, B.size() * 2));
  CK(cudaMemset(dA_, 0x3c, A.size() * 2)); CK(cudaMemset(dG_, 0x3c, A.size() * 2));
  CK(cudaMemset(dB_, 0x3c, B.size() * 2)); CK(cudaMemset(dV_, 0x3c, B.size() * 2));
  CK(cudaMalloc((void**)&nl, (size_t)H * M * 4)); CK(cudaMemset(nl, 0, (size_t)H * M * 4));
  CK(cudaMalloc((void**)&nd, (size_t)H * M * 4)); CK(cudaMemset(nd, 0, (size_t)H * M * 4));
  ddp_tensor a = dview(A, dA_, dt), g = dview(A, dG_, dt), b = dview(B, dB_, dt), v = dview(B, dV_, dt),
             oa = dview(A, oA, dt), ob = dview(B, oB, dt), ov = dview(B, oV, dt);
  const float scale = 1.0f / sqrtf((float)d);
  uint32_t *mb = nullptr, *mbt = nullptr;
  const long long ldw = (Tn + 31) / 32, ldwt = (M + 31) / 32;
  if (with_mask) {
    CK(cudaMalloc((void**)&mb, (size_t)M * ldw * 4)); CK(cudaMemset(mb, 0, (size_t)M * ldw * 4));
    CK(cudaMalloc((void**)&mbt, (size_t)Tn * ldwt * 4)); CK(cudaMemset(mbt, 0, (size_t)Tn * ldwt * 4));
  }
  cudaEvent_t e0, e1, e2;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));
  for (int i = 0; i < 2; ++i) {
    if (ddp_attn_bwd_rows(&a, &g, &b, &v, &oa, nl, nd, M, mb, ldw, (long long)M * ldw, scale, nullptr, 0, 0) ||
        ddp_attn_bwd_cols(&a, &g, &b, &v, &ov, &ob, nl, nd, M, mbt, ldwt, (long long)Tn * ldwt, scale, nullptr, 0, 0)) {
      printf("perf launch failed: %s\n", ddp_last_error());
      return;
    }
  }
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < iters; ++i) ddp_attn_bwd_rows(&a, &g, &b, &v, &oa, nl, nd, M, mb, ldw, (long long)M * ldw, scale, nullptr, 0, 0);
  CK(cudaEventRecord(e1));
  for (int i = 0; i < iters; ++i) ddp_attn_bwd_cols(&a, &g, &b, &v, &ov, &ob, nl, nd, M, mbt, ldwt, (long long)Tn * ldwt, scale, nullptr, 0, 0);
  CK(cudaEventRecord(e2));
  CK(cudaEventSynchronize(e2));
  float ms_r, ms_c;
  CK(cudaEventElapsedTime(&ms_r, e0, e1));
  CK(cudaEventElapsedTime(&ms_c, e1, e2));
  ms_r /= iters; ms_c /= iters;
  const double tile = 2.0 * H * (double)M * Tn * d;
  const int site = ddp_debug_timeout_site_attn();
  printf("PERF ATTN-BWD %s H=%d M=%d T=%d d=%d mask=%d: rows (dA) %.3f ms %.1f TFLOP/s   cols (dB, dV) %.3f ms %.1f TFLOP/s%s\n",
         dt == DDP_BF16 ? "bf16" : "fp16", H, M, Tn, d, (int)with_mask, ms_r, 3 * tile / ms_r * 1e-9, ms_c, 4 * tile / ms_c * 1e-9,
         site ? "  WATCHDOG FIRED" : "");
  fflush(stdout);
}

static void perf_case(int dt, int H, int M, int Tn, int d, bool with_mask, int iters, int dbg = 0) {
  HT A = make_ht(1, H, M, d, true, false), B = make_ht(1, H, Tn, d, true, false), V = make_ht(1, H, Tn, d, true, false),
     O = make_ht(1, H, M, d, true, false);
  void *dA, *dB, *dV, *dO, *dBits = nullptr;
  float* dLse;
  CK(cudaMalloc(&dA, A.size() * 2));
  CK(cudaMalloc(&dB, B.size() * 2));
  CK(cudaMalloc(&dV, V.size() * 2));
  CK(cudaMalloc(&dO, O.size() * 2));
  CK(cudaMalloc((void**)&dLse, (size_t)H * M * 4));
  // small deterministic values: 0x3c3c as bf16 is ~0.0115
  CK(cudaMemset(dA, 0x3c, A.size() * 2));
  CK(cudaMemset(dB, 0x3c, B.size() * 2));
  CK(cudaMemset(dV, 0x3c, V.size() * 2));
  const long long ldw = (Tn + 31) / 32;
  if (with_mask) {
    CK(cudaMalloc(&dBits, (size_t)M * ldw * 4));
    CK(cudaMemset(dBits, 0, (size_t)M * ldw * 4));
  }
  ddp_tensor a = dview(A, dA, dt), b = dview(B, dB, dt), v = dview(V, dV, dt), o = dview(O, dO, dt);
  const float scale = 1.0f / sqrtf((float)d);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 2; ++i)
    if (ddp_attn_fwd(&a, &b, &v, &o, dLse, (const uint32_t*)dBits, ldw, (long long)M * ldw, scale, nullptr, dbg, 0)) {
      printf("perf launch failed: %s\n", ddp_last_error());
      return;
    }
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < iters; ++i)
    ddp_attn_fwd(&a, &b, &v, &o, dLse, (const uint32_t*)dBits, ldw, (long long)M * ldw, scale, nullptr, dbg, 0);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  ms /= iters;
  const double flops = 4.0 * H * (double)M * Tn * d;
  const int site = ddp_debug_timeout_site_attn();
  printf("PERF ATTN %s H=%d M=%d T=%d d=%d mask=%d dbg=%d: %.3f ms  %.1f TFLOP/s  (%.0f clk per 128x128 tile and SM)%s\n",
         dt == DDP_BF16 ? "bf16" : "fp16", H, M, Tn, d, (int)with_mask, dbg, ms, flops / ms * 1e-9,
         ms * 1e-3 * 1.965e9 * 148.0 / ((double)H * (M / 128.0) * (Tn / 128.0)), site ? "  WATCHDOG FIRED" : "");
  fflush(stdout);
  cudaFree(dA); cudaFree(dB); cudaFree(dV); cudaFree(dO); cudaFree(dLse);
  if (dBits) cudaFree(dBits);
}

int main(int argc, char** argv) {
  std::string mode = argc > 1 ? argv[1] : "quick";
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("device: %s sm_%d%d, %d SMs\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  if (mode == "one" && argc >= 14) {
    Case c;
    c.dt = atoi(argv[2]); c.world = atoi(argv[3]); c.b1 = atoi(argv[4]); c.b0 = atoi(argv[5]); c.M = atoi(argv[6]);
    c.Tn = atoi(argv[7]); c.d = atoi(argv[8]); c.dv = atoi(argv[9]); c.interleave = atoi(argv[10]) != 0;
    c.maskmode = atoi(argv[11]); c.chunk = atoi(argv[12]); c.ramp = atoi(argv[13]);
    run_case(c, argc > 14 ? atoi(argv[14]) : 0, true);
    return g_fail ? 1 : 0;
  }
  if (mode == "trace") {  // pipeline time stamps of CTA 0 (ddp_attn_debug_trace), flagship shape; dbg = argv[2]
    int na, ns, ne;
    ddp_attn_debug_trace_dims(&na, &ns, &ne);
    const size_t n = (size_t)na * ns * ne;
    long long* dbuf;
    CK(cudaMalloc((void**)&dbuf, n * 8));
    CK(cudaMemset(dbuf, 0, n * 8));
    const int dbg = argc > 2 ? atoi(argv[2]) : 0;
    const bool with_mask = argc > 3 ? atoi(argv[3]) != 0 : false;
    perf_case(DDP_BF16, 16, 8192, 8192, 64, with_mask, 2, dbg);  // warm
    ddp_attn_debug_trace(dbuf);
    perf_case(DDP_BF16, 16, 8192, 8192, 64, with_mask, 1, dbg);
    ddp_attn_debug_trace(nullptr);
    std::vector<long long> h(n);
    CK(cudaMemcpy(h.data(), dbuf, n * 8, cudaMemcpyDeviceToHost));
    auto at = [&](int a, int s_, int e) { return h[((size_t)a * ns + s_) * ne + e]; };
    const long long t0 = at(3, 0, 0);
    printf("TRACE dbg=%d mask=%d: clocks relative to the first S issue of issuer 0\n", dbg, (int)with_mask);
    printf("step | iss0: S PV0 PV1 | iss1: S PV0 PV1 | warp(w0h0): wait_s got_s ld_done max_done got_odone exp_done pfull | warp(w0h1) ... | warp(w1h0) ...\n");
    for (int s_ = 8; s_ < 14; ++s_) {
      printf("%3d |", s_);
      for (int a = 3; a < 5; ++a) printf(" %7lld %7lld %7lld |", at(a, s_, 0) - t0, at(a, s_, 1) - t0, at(a, s_, 2) - t0);
      for (int a = 0; a < 3; ++a) {
        for (int e = 0; e < 7; ++e) printf(" %7lld", at(a, s_, e) - t0);
        printf(" |");
      }
      printf("\n");
    }
    // average phase lengths over steps 8..55 for the first warp
    for (int a = 0; a < 8; ++a) {
      if (a == 3 || a == 4) continue;
      double d[7] = {0};
      int cnt = 0;
      for (int s_ = 8; s_ < 56; ++s_, ++cnt) {
        d[0] += at(a, s_, 1) - at(a, s_, 0);   // waiting for S
        d[1] += at(a, s_, 2) - at(a, s_, 1);   // TMEM load + s_empty arrive
        d[2] += at(a, s_, 3) - at(a, s_, 2);   // max
        d[3] += at(a, s_, 4) - at(a, s_, 3);   // waiting for o_done
        d[4] += at(a, s_, 5) - at(a, s_, 4);   // (rescale) + exp + pack + st.shared
        d[5] += at(a, s_, 6) - at(a, s_, 5);   // fences + p_full arrive
        d[6] += at(a, s_ + 1, 0) - at(a, s_, 0);  // whole step
      }
      printf("TRACE-AVG actor %d: wait_S %.0f  ld %.0f  max %.0f  wait_odone %.0f  exp+st %.0f  fence+arrive %.0f  | step %.0f clk\n",
             a, d[0] / cnt, d[1] / cnt, d[2] / cnt, d[3] / cnt, d[4] / cnt, d[5] / cnt, d[6] / cnt);
    }
    {  // skew between the four lane quadrants of stream (w0, h0): ld_done and p_full arrive relative to q = 0
      const int qa[4] = {0, 5, 6, 7};
      for (int e : {2, 6}) {
        double sk[4] = {0};
        int cnt = 0;
        for (int s_ = 8; s_ < 56; ++s_, ++cnt)
          for (int q = 0; q < 4; ++q) sk[q] += at(qa[q], s_, e) - at(0, s_, e);
        printf("TRACE-SKEW event %d (2 = S loaded, 6 = P staged) of q1 q2 q3 vs q0: %.0f %.0f %.0f clk\n", e, sk[1] / cnt,
               sk[2] / cnt, sk[3] / cnt);
      }
    }
    {
      double sS = 0, sPV0 = 0, sPV1 = 0, lat_s = 0, lat_pv = 0;
      int cnt = 0;
      for (int s_ = 8; s_ < 56; ++s_, ++cnt) {
        sS += at(3, s_ + 1, 0) - at(3, s_, 0);
        // S(j) issue -> softmax warp 0 sees s_full(j)
        lat_s += at(0, s_, 1) - at(3, s_, 0);
        // p_full arrive of warp 0 (step j) -> PV0(j) issued ; PV0(j) issued -> o_done seen by warp 0 in step j+1
        sPV0 += at(3, s_, 1) - at(0, s_, 6);
        lat_pv += at(0, s_ + 1, 4) - at(3, s_, 1);
        sPV1 += at(3, s_, 2) - at(3, s_, 1);
      }
      printf("TRACE-AVG issuer 0: S period %.0f  | S issue -> warp sees S %.0f | warp p_full arrive -> PV0 issue %.0f | PV0 issue -> warp sees o_done %.0f | PV1 - PV0 issue %.0f\n",
             sS / cnt, lat_s / cnt, sPV0 / cnt, lat_pv / cnt, sPV1 / cnt);
    }
    return 0;
  }
  if (mode == "perflong") {  // the per-GPU work of the 8-GPU flagship (8192 rows x 65536 columns) on one GPU
    perf_case(DDP_BF16, 16, 8192, 65536, 64, true, 2);
    perf_bwd(DDP_BF16, 16, 8192, 65536, 64, 2, true);
    return 0;
  }
  if (mode == "perfall") {  // timing only, all three kernels, with and without a (zero) bit mask
    perf_case(DDP_BF16, 16, 8192, 8192, 64, false, 5);
    perf_case(DDP_BF16, 16, 8192, 8192, 64, true, 5);
    perf_bwd(DDP_BF16, 16, 8192, 8192, 64, 5);
    perf_bwd(DDP_BF16, 16, 8192, 8192, 64, 5, true);
    return 0;
  }
  if (mode != "perf" && mode != "bwd" && mode != "bwdfull" && mode != "prof" && mode != "multi") {
    const int bf = DDP_BF16;
    //        dt  W b1 b0 M    Tn   d   dv  il     mask chunk ramp
    // stage 1: linear mode (no max / exp): isolates S MMA -> P staging -> PV MMA -> O store
    if (!run_case({bf, 1, 1, 1, 256, 128, 64, 64, false, 0, 64, 0}, 1, true)) {
      run_case({bf, 1, 1, 1, 256, 128, 64, 64, false, 0, 64, 0}, 1 | 16, true);  // generic O stores
      run_case({bf, 1, 1, 1, 256, 128, 64, 64, false, 0, 64, 0}, 0, true);
      printf("SUMMARY pass=%d fail=%d (stopped early)\n", g_pass, g_fail);
      return 1;
    }
    if (!run_case({bf, 1, 1, 1, 256, 384, 64, 64, false, 0, 64, 0}, 1, true)) {
      printf("SUMMARY pass=%d fail=%d (stopped early)\n", g_pass, g_fail);
      return 1;
    }
    // stage 2: the real thing
    run_case({bf, 1, 1, 1, 256, 128, 64, 64, false, 0, 64, 0}, 0, true);
    run_case({bf, 1, 1, 1, 256, 512, 64, 64, false, 0, 64, 0}, 0, true);
    run_case({bf, 1, 1, 1, 256, 512, 64, 64, false, 0, 64, 1}, 0, true);   // lazy rescale
    run_case({bf, 1, 1, 2, 512, 640, 64, 64, true, 1, 64, 1}, 0, true);    // heads interleaved, random mask
    run_case({bf, 1, 2, 3, 200, 300, 32, 32, true, 2, 64, 1}, 0, true);    // ragged, fully masked rows / tiles
    run_case({bf, 1, 1, 2, 100, 72, 40, 24, false, 1, 64, 0}, 0, true);    // d != dv, everything ragged
    run_case({DDP_F16, 1, 1, 2, 384, 256, 64, 64, true, 1, 64, 0}, 0, true);
    run_case({bf, 1, 1, 4, 1024, 1024, 64, 64, true, 0, 64, 0}, 0, true);  // several items per CTA? (32 items)
    // stage 3: emulated ranks
    if (mode == "full" || mode == "quick") {
      run_case({bf, 2, 1, 2, 256, 256, 64, 64, true, 0, 64, 0}, 0, true);
      run_case({bf, 2, 1, 2, 256, 256, 64, 64, true, 1, 128, 1}, 0, true);
      run_case({bf, 4, 2, 2, 160, 160, 32, 32, true, 2, 64, 1}, 0, true);
      run_case({bf, 4, 1, 3, 100, 96, 64, 64, false, 1, 32, 0}, 0, true);
    }
    printf("SUMMARY pass=%d fail=%d\n", g_pass, g_fail);
  }
  if (mode == "multi") {  // more work items than CTAs: every barrier phase / ring slot is carried across items
    const int bf = DDP_BF16;
    //        dt  W b1 b0  M     Tn   d   dv  il    mask chunk ramp
    run_case({bf, 1, 1, 16, 2560, 128, 64, 64, true, 1, 64, 1}, 0, true);       // forward: 160 items
    run_bwd_case({bf, 1, 1, 16, 2560, 128, 64, 64, true, 1, 64, 0}, true);      // row side: 160 items
    run_bwd_case({bf, 1, 1, 16, 128, 2560, 64, 64, true, 1, 64, 0}, true);      // column side: 160 items
    run_bwd_case({bf, 2, 1, 16, 128, 1280, 64, 64, true, 0, 256, 0}, true);     // column side, 2 ranks: 2 x 80 items
    printf("SUMMARY-MULTI pass=%d fail=%d\n", g_pass, g_fail);
    return g_fail ? 1 : 0;
  }
  if (mode == "prof") {  // exactly one launch of every attention kernel on the flagship per-GPU shape (for ncu)
    perf_case(DDP_BF16, 16, 8192, 8192, 64, true, 1);
    perf_bwd(DDP_BF16, 16, 8192, 8192, 64, 1, true);
    return 0;
  }
  if (mode == "bwd" || mode == "bwdfull") {
    const int bf = DDP_BF16;
    //            dt  W b1 b0 M    Tn   d   dv  il     mask chunk ramp
    run_bwd_case({bf, 1, 1, 1, 256, 128, 64, 64, false, 0, 64, 0}, true);
    run_bwd_case({bf, 1, 1, 2, 256, 512, 64, 64, true, 1, 64, 0}, true);
    run_bwd_case({bf, 1, 2, 3, 200, 300, 32, 32, true, 2, 64, 0}, true);
    run_bwd_case({bf, 1, 1, 2, 100, 72, 40, 24, false, 1, 64, 0}, true);
    run_bwd_case({DDP_F16, 1, 1, 2, 384, 256, 64, 64, true, 1, 64, 0}, true);
    run_bwd_case({bf, 1, 1, 4, 1024, 1024, 64, 64, true, 0, 64, 0}, true);
    run_bwd_case({bf, 2, 1, 2, 256, 256, 64, 64, true, 0, 64, 0}, true);
    run_bwd_case({bf, 2, 1, 2, 256, 256, 64, 64, true, 1, 128, 0}, true);
    run_bwd_case({bf, 4, 2, 2, 160, 160, 32, 32, true, 2, 64, 0}, true);
    run_bwd_case({bf, 4, 1, 3, 100, 96, 64, 64, false, 1, 32, 0}, true);
    printf("SUMMARY-BWD pass=%d fail=%d\n", g_pass, g_fail);
    if (mode == "bwdfull") {
      perf_bwd(DDP_BF16, 16, 8192, 8192, 64, 5);
      perf_bwd(DDP_BF16, 16, 8192, 8192, 64, 5, true);
    }
    return g_fail ? 1 : 0;
  }
  if (mode == "perf" || mode == "full") {
    perf_case(DDP_BF16, 16, 8192, 8192, 64, false, 5);
    perf_case(DDP_BF16, 16, 8192, 8192, 64, true, 5);
    perf_case(DDP_BF16, 16, 8192, 8192, 64, false, 5, 1);  // linear mode: MMA + staging without the exponentials
  }
  return g_fail ? 1 : 0;
}
