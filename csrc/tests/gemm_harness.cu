// Standalone validation + timing harness for ddp_gemm (no Python).  Runs on ONE GPU:
//   * world == 1 cases for NT / NN / TN in bf16, fp16 and fp32(tf32), ragged and batched/strided shapes
//   * emulated multi-rank cases: every "rank" owns its own workspace/flag buffers on the same GPU and the
//     produce phase of all ranks is launched before the consume phase (kernels that wait on each other must
//     not be co-scheduled on one GPU)
//   * a few large shapes timed with CUDA events.
// Usage: gemm_harness [quick|full|perf]
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

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess) {                                                               \
      printf("CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                             \
    }                                                                                      \
  } while (0)

static uint32_t g_seed = 12345;
static int irand(int lo, int hi) {
  g_seed = g_seed * 1664525u + 1013904223u;
  return lo + (int)((g_seed >> 8) % (uint32_t)(hi - lo + 1));
}

static int esize(int dt) { return dt == DDP_F32 ? 4 : 2; }
static const char* dtname(int dt) { return dt == DDP_F32 ? "tf32" : dt == DDP_BF16 ? "bf16" : "fp16"; }
static const char* opname(int op) { return op == DDP_OP_NT ? "NT" : op == DDP_OP_NN ? "NN" : "TN"; }

static void to_dev(const std::vector<float>& h, int dt, void* d) {
  if (dt == DDP_F32) {
    CK(cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  } else if (dt == DDP_BF16) {
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
  if (dt == DDP_F32) {
    CK(cudaMemcpy(h.data(), d, n * 4, cudaMemcpyDeviceToHost));
  } else if (dt == DDP_BF16) {
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

// Host tensor [b1][b0][rows][cols] stored with padding: ld >= cols; optional "head interleaved" layout where
// b0 is interleaved inside the row (sb0 = cols, ld = b0*cols) like a [B, T, H, dh] projection output.
struct HT {
  int b1, b0, rows, cols;
  long long ld, sb0, sb1;
  std::vector<float> v;  // physical storage
  size_t size() const { return v.size(); }
  float& at(int i1, int i0, int r, int c) { return v[i1 * sb1 + i0 * sb0 + (long long)r * ld + c]; }
  float at(int i1, int i0, int r, int c) const { return v[i1 * sb1 + i0 * sb0 + (long long)r * ld + c]; }
};
static HT make_ht(int b1, int b0, int rows, int cols, bool interleave, int pad, int lo, int hi, bool fill) {
  HT t;
  t.b1 = b1; t.b0 = b0; t.rows = rows; t.cols = cols;
  if (interleave) {
    t.sb0 = cols; t.ld = (long long)b0 * cols + pad; t.sb1 = t.ld * rows;
  } else {
    t.ld = cols + pad; t.sb0 = t.ld * rows; t.sb1 = t.sb0 * b0;
  }
  t.v.assign((size_t)t.sb1 * b1, 0.f);
  if (fill)
    for (int i1 = 0; i1 < b1; ++i1)
      for (int i0 = 0; i0 < b0; ++i0)
        for (int r = 0; r < rows; ++r)
          for (int c = 0; c < cols; ++c) t.at(i1, i0, r, c) = (float)irand(lo, hi);
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
  int op, dt, world;
  int b1, b0;
  int M;      // rows of left per rank
  int Tn;     // rows of right per rank
  int D;
  bool interleave;
  int pad;
  int chunk;
  int val;    // input magnitude
};

static int g_fail = 0, g_pass = 0;

static bool run_case(const Case& c, int dbg, bool verbose) {
  const int W = c.world, es = esize(c.dt);
  const int T = W * c.Tn;
  // per-rank operands
  std::vector<HT> A(W), B(W), C(W);
  for (int r = 0; r < W; ++r) {
    if (c.op == DDP_OP_NT) {
      A[r] = make_ht(c.b1, c.b0, c.M, c.D, c.interleave, c.pad, -c.val, c.val, true);
      B[r] = make_ht(c.b1, c.b0, c.Tn, c.D, c.interleave, c.pad, -c.val, c.val, true);
      C[r] = make_ht(c.b1, c.b0, c.M, T, false, c.pad, 0, 0, false);
    } else if (c.op == DDP_OP_NN) {
      A[r] = make_ht(c.b1, c.b0, c.M, T, false, c.pad, -c.val, c.val, true);
      B[r] = make_ht(c.b1, c.b0, c.Tn, c.D, c.interleave, c.pad, -c.val, c.val, true);
      C[r] = make_ht(c.b1, c.b0, c.M, c.D, c.interleave, c.pad, 0, 0, false);
    } else {
      A[r] = make_ht(c.b1, c.b0, c.Tn, T, false, c.pad, -c.val, c.val, true);
      B[r] = make_ht(c.b1, c.b0, c.Tn, c.D, c.interleave, c.pad, -c.val, c.val, true);
      C[r] = make_ht(c.b1, c.b0, c.Tn, c.D, c.interleave, c.pad, 0, 0, false);
    }
  }
  std::vector<void*> dA(W), dB(W), dC(W);
  for (int r = 0; r < W; ++r) {
    CK(cudaMalloc(&dA[r], A[r].size() * es));
    CK(cudaMalloc(&dB[r], B[r].size() * es));
    CK(cudaMalloc(&dC[r], C[r].size() * es));
    to_dev(A[r].v, c.dt, dA[r]);
    to_dev(B[r].v, c.dt, dB[r]);
    CK(cudaMemset(dC[r], 0xFF, C[r].size() * es));  // NaN pattern: unwritten outputs show up
  }
  // workspaces
  std::vector<void*> ws(W, nullptr);
  std::vector<uint32_t*> flags(W, nullptr), done(W, nullptr), exitc(W, nullptr);
  if (W > 1) {
    size_t wsb = 0, nf = 0;
    ddp_tensor a0 = dview(A[0], dA[0], c.dt), b0 = dview(B[0], dB[0], c.dt);
    if (ddp_gemm_workspace(c.op, &a0, &b0, W, c.chunk, &wsb, &nf)) {
      printf("workspace query failed: %s\n", ddp_last_error());
      return false;
    }
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
  const int nseq = (W > 1) ? 2 : 1;  // run twice in emulated mode to exercise seq / done handling
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
        ddp_tensor a = dview(A[r], dA[r], c.dt), b = dview(B[r], dB[r], c.dt), cc = dview(C[r], dC[r], c.dt);
        int e = ddp_gemm(c.op, &a, &b, &cc, 1.0f, nullptr, W > 1 ? &cm : nullptr, dbg, 0);
        if (e) {
          printf("  ddp_gemm error %d: %s\n", e, ddp_last_error());
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
  // reference + compare
  double max_err = 0;
  long long nbad = 0, nshown = 0;
  if (ok) {
    for (int r = 0; r < W; ++r) {
      std::vector<float> got = from_dev(dC[r], C[r].size(), c.dt);
      for (int i1 = 0; i1 < c.b1; ++i1)
        for (int i0 = 0; i0 < c.b0; ++i0)
          for (int m = 0; m < C[r].rows; ++m)
            for (int n = 0; n < C[r].cols; ++n) {
              double ref = 0;
              if (c.op == DDP_OP_NT) {
                const int src = n / c.Tn, j = n % c.Tn;
                for (int k = 0; k < c.D; ++k) ref += (double)A[r].at(i1, i0, m, k) * B[src].at(i1, i0, j, k);
              } else if (c.op == DDP_OP_NN) {
                for (int k = 0; k < T; ++k) ref += (double)A[r].at(i1, i0, m, k) * B[k / c.Tn].at(i1, i0, k % c.Tn, n);
              } else {
                for (int p = 0; p < W; ++p)
                  for (int k = 0; k < c.Tn; ++k) ref += (double)A[p].at(i1, i0, k, r * c.Tn + m) * B[p].at(i1, i0, k, n);
              }
              const float g = got[i1 * C[r].sb1 + i0 * C[r].sb0 + (long long)m * C[r].ld + n];
              const double tol = (c.dt == DDP_F32 ? 1e-5 : 1.0 / 100) * (fabs(ref) + 1.0);
              const double err = fabs((double)g - ref);
              if (!(err <= tol)) {
                ++nbad;
                if (verbose && nshown < 6) {
                  printf("    bad rank %d b(%d,%d) m %d n %d: got %g want %g\n", r, i1, i0, m, n, g, ref);
                  ++nshown;
                }
              }
              if (err == err && err > max_err) max_err = err;
            }
    }
    if (nbad) ok = false;
  }
  printf("%s %s %s W=%d b=(%d,%d) M=%d Tn=%d D=%d il=%d pad=%d chunk=%d dbg=%d  max_err=%.3g bad=%lld\n",
         ok ? "PASS" : "FAIL", opname(c.op), dtname(c.dt), W, c.b1, c.b0, c.M, c.Tn, c.D, (int)c.interleave, c.pad,
         c.chunk, dbg, max_err, nbad);
  fflush(stdout);
  for (int r = 0; r < W; ++r) {
    cudaFree(dA[r]); cudaFree(dB[r]); cudaFree(dC[r]);
    if (W > 1) { cudaFree(ws[r]); cudaFree(flags[r]); cudaFree(done[r]); cudaFree(exitc[r]); }
  }
  if (ok) ++g_pass; else ++g_fail;
  return ok;
}

static void perf_case(int op, int dt, int b0, int M, int Tn, int D, int iters, int dbg = 0) {
  const int es = esize(dt);
  HT A, B, C;
  if (op == DDP_OP_NT) {
    A = make_ht(1, b0, M, D, false, 0, 0, 0, false);
    B = make_ht(1, b0, Tn, D, false, 0, 0, 0, false);
    C = make_ht(1, b0, M, Tn, false, 0, 0, 0, false);
  } else if (op == DDP_OP_NN) {
    A = make_ht(1, b0, M, Tn, false, 0, 0, 0, false);
    B = make_ht(1, b0, Tn, D, false, 0, 0, 0, false);
    C = make_ht(1, b0, M, D, false, 0, 0, 0, false);
  } else {
    A = make_ht(1, b0, Tn, Tn, false, 0, 0, 0, false);
    B = make_ht(1, b0, Tn, D, false, 0, 0, 0, false);
    C = make_ht(1, b0, Tn, D, false, 0, 0, 0, false);
  }
  void *dA, *dB, *dC;
  CK(cudaMalloc(&dA, A.size() * es));
  CK(cudaMalloc(&dB, B.size() * es));
  CK(cudaMalloc(&dC, C.size() * es));
  CK(cudaMemset(dA, 0x3c, A.size() * es));
  CK(cudaMemset(dB, 0x3c, B.size() * es));
  ddp_tensor a = dview(A, dA, dt), b = dview(B, dB, dt), c = dview(C, dC, dt);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i)
    if (ddp_gemm(op, &a, &b, &c, 1.f, nullptr, nullptr, dbg, 0)) {
      printf("perf launch failed: %s\n", ddp_last_error());
      return;
    }
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < iters; ++i) ddp_gemm(op, &a, &b, &c, 1.f, nullptr, nullptr, dbg, 0);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  ms /= iters;
  const double rowsA = (op == DDP_OP_TN) ? Tn : M;
  double flops, bytes;
  if (op == DDP_OP_NT) {
    flops = 2.0 * b0 * M * (double)Tn * D;
  } else if (op == DDP_OP_NN) {
    flops = 2.0 * b0 * M * (double)Tn * D;
  } else {
    flops = 2.0 * b0 * Tn * (double)Tn * D;
  }
  (void)rowsA;
  bytes = (double)(A.size() + B.size() + C.size()) * es;
  printf("PERF dbg=%d %s %s b0=%d M=%d Tn=%d D=%d: %.3f ms  %.1f TFLOP/s  %.1f GB/s (algorithmic)\n", dbg, opname(op), dtname(dt),
         b0, M, Tn, D, ms, flops / ms * 1e-9, bytes / ms * 1e-6);
  fflush(stdout);
  cudaFree(dA); cudaFree(dB); cudaFree(dC);
}

int main(int argc, char** argv) {
  std::string mode = argc > 1 ? argv[1] : "quick";
  int dev_count = 0;
  CK(cudaGetDeviceCount(&dev_count));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("device: %s sm_%d%d, %d SMs\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);

  if (mode == "one" && argc >= 13) {
    Case c;
    c.op = atoi(argv[2]); c.dt = atoi(argv[3]); c.world = atoi(argv[4]); c.b1 = atoi(argv[5]); c.b0 = atoi(argv[6]);
    c.M = atoi(argv[7]); c.Tn = atoi(argv[8]); c.D = atoi(argv[9]); c.interleave = atoi(argv[10]) != 0;
    c.pad = atoi(argv[11]); c.chunk = atoi(argv[12]); c.val = 2;
    run_case(c, argc > 13 ? atoi(argv[13]) : 0, true);
    return g_fail ? 1 : 0;
  }
  if (mode != "perf") {
    // ---- stage 1: the smallest K-major case; if this fails nothing else is meaningful
    const int dts[3] = {DDP_BF16, DDP_F32, DDP_F16};
    for (int di = 0; di < 3; ++di) {
      const int dt = dts[di];
      //            op        dt  W  b1 b0 M    Tn   D    il     pad chunk val
      run_case({DDP_OP_NT, dt, 1, 1, 1, 128, 64, 64, false, 0, 64, 2}, 0, true);
      run_case({DDP_OP_NT, dt, 1, 1, 1, 128, 256, 128, false, 0, 64, 2}, 0, true);
      run_case({DDP_OP_NT, dt, 1, 2, 3, 200, 300, 96, false, 8, 64, 2}, 0, true);
      run_case({DDP_OP_NT, dt, 1, 1, 4, 150, 150, 32, true, 0, 64, 2}, 0, true);
      // NN (B MN-major)
      if (!run_case({DDP_OP_NN, dt, 1, 1, 1, 128, 128, 64, false, 0, 64, 2}, 0, true))
        for (int dbg = 1; dbg < 4; ++dbg) run_case({DDP_OP_NN, dt, 1, 1, 1, 128, 128, 64, false, 0, 64, 2}, dbg, true);
      run_case({DDP_OP_NN, dt, 1, 1, 1, 128, 256, 128, false, 0, 64, 2}, 0, true);
      run_case({DDP_OP_NN, dt, 1, 2, 2, 200, 328, 264, false, 8, 64, 2}, 0, true);
      run_case({DDP_OP_NN, dt, 1, 1, 4, 100, 160, 32, true, 0, 64, 2}, 0, true);
      // TN (both MN-major)
      if (!run_case({DDP_OP_TN, dt, 1, 1, 1, 0, 128, 64, false, 0, 64, 2}, 0, true))
        for (int dbg = 1; dbg < 4; ++dbg) run_case({DDP_OP_TN, dt, 1, 1, 1, 0, 128, 64, false, 0, 64, 2}, dbg, true);
      run_case({DDP_OP_TN, dt, 1, 1, 1, 0, 256, 128, false, 0, 64, 2}, 0, true);
      run_case({DDP_OP_TN, dt, 1, 2, 2, 0, 200, 264, false, 8, 64, 2}, 0, true);
      run_case({DDP_OP_TN, dt, 1, 1, 4, 0, 136, 32, true, 0, 64, 2}, 0, true);
    }
    // ---- stage 2: emulated ranks
    for (int di = 0; di < 2; ++di) {
      const int dt = dts[di];
      for (int W = 2; W <= 4; W += 2) {
        run_case({DDP_OP_NT, dt, W, 1, 2, 128, 128, 64, false, 0, 64, 2}, 0, true);
        run_case({DDP_OP_NT, dt, W, 2, 2, 96, 96, 64, true, 0, 32, 2}, 0, true);
        run_case({DDP_OP_NT, dt, W, 1, 1, 328, 328, 96, false, 0, 100, 2}, 0, true);
        run_case({DDP_OP_NN, dt, W, 1, 2, 128, 128, 64, false, 0, 64, 2}, 0, true);
        run_case({DDP_OP_NN, dt, W, 2, 2, 96, 96, 64, true, 0, 32, 2}, 0, true);
        run_case({DDP_OP_NN, dt, W, 1, 1, 328, 328, 96, false, 0, 100, 2}, 0, true);
        run_case({DDP_OP_NT, dt, W, 1, 2, 100, 100, 72, false, 0, 64, 2}, 0, true);
        run_case({DDP_OP_TN, dt, W, 1, 2, 0, 128, 64, false, 0, 64, 2}, 0, true);
        run_case({DDP_OP_TN, dt, W, 2, 2, 0, 96, 64, true, 0, 32, 2}, 0, true);
        run_case({DDP_OP_TN, dt, W, 1, 1, 0, 328, 96, false, 0, 100, 2}, 0, true);
      }
    }
    printf("SUMMARY pass=%d fail=%d\n", g_pass, g_fail);
  }
  if (mode == "epi") {
    perf_case(DDP_OP_NT, DDP_BF16, 16, 8192, 8192, 64, 10, 0);
    perf_case(DDP_OP_NT, DDP_BF16, 16, 8192, 8192, 64, 10, 4);
    perf_case(DDP_OP_NT, DDP_BF16, 16, 8192, 8192, 64, 10, 8);
    perf_case(DDP_OP_NT, DDP_F32, 16, 8192, 8192, 64, 10, 0);
    perf_case(DDP_OP_NT, DDP_F32, 16, 8192, 8192, 64, 10, 4);
    perf_case(DDP_OP_NT, DDP_F32, 16, 8192, 8192, 64, 10, 8);
  }
  if (mode == "perf" || mode == "full") {
    // flagship per-rank shapes at world 1 (T = 8192) and compute-bound shapes
    perf_case(DDP_OP_NT, DDP_BF16, 16, 8192, 8192, 64, 10);
    perf_case(DDP_OP_NN, DDP_BF16, 16, 8192, 8192, 64, 10);
    perf_case(DDP_OP_TN, DDP_BF16, 16, 8192, 8192, 64, 10);
    perf_case(DDP_OP_NT, DDP_BF16, 1, 8192, 8192, 1024, 10);
    perf_case(DDP_OP_NT, DDP_BF16, 1, 8192, 8192, 8192, 5);
    perf_case(DDP_OP_NN, DDP_BF16, 1, 8192, 8192, 1024, 10);
    perf_case(DDP_OP_TN, DDP_BF16, 1, 8192, 8192, 1024, 10);
    perf_case(DDP_OP_NT, DDP_F32, 1, 8192, 8192, 768, 10);
    perf_case(DDP_OP_NN, DDP_F32, 1, 8192, 8192, 768, 10);
    perf_case(DDP_OP_TN, DDP_F32, 1, 8192, 8192, 768, 10);
  }
  return g_fail ? 1 : 0;
}
