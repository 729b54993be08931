// Host side of the fused collective GEMM: shape checks, TMA tensor-map construction, tile-shape
// selection and launch.  C ABI declared in include/ddp_b200.h.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <type_traits>

#include "../common/host_util.h"
#include "gemm_sm100.cuh"
#include "tma_map.h"

namespace ddp {

using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                              const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                              CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn get_encode_fn() {
  static EncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &p, 12000, cudaEnableDefault, &st) ==
            cudaSuccess &&
        st == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeFn>(p);
  });
  return fn;
}

static int elem_size(int dt) { return dt == DDP_F32 ? 4 : 2; }

// dims/strides innermost first; strides in bytes for dims 1..4.
int make_map(CUtensorMap* m, int dt, void* ptr, const uint64_t dims[5], const uint64_t strides[4],
             const uint32_t box[5], bool mn_major) {
  EncodeFn enc = get_encode_fn();
  if (!enc) return set_error(-10, "cuTensorMapEncodeTiled is not available from the driver");
  // The encoder is a driver-API call and needs a current context on THIS thread.  A fresh thread (e.g. the
  // autograd engine's worker when our node is the first CUDA work it sees) has none until a runtime call
  // binds the primary context: do that once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(nullptr);
    ctx_bound = true;
  }
  if (reinterpret_cast<uintptr_t>(ptr) & 15) return set_error(-2, "tensor base address is not 16-byte aligned");
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < 5; ++i) {
    gd[i] = dims[i] ? dims[i] : 1;
    bx[i] = box[i];
    if (box[i] == 0 || box[i] > 256) return set_error(-3, "TMA box extent out of range");
  }
  for (int i = 0; i < 4; ++i) {
    gs[i] = strides[i];
    if (gs[i] & 15) return set_error(-2, "tensor stride is not a multiple of 16 bytes");
    if (gs[i] == 0) gs[i] = 16;
  }
  const CUtensorMapDataType t = dt == DDP_F32    ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                : dt == DDP_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  // MN-major tf32 operands only exist in the 32-byte-atom flavour of the 128-byte swizzle.
  const CUtensorMapSwizzle sw =
      (mn_major && dt == DDP_F32) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = enc(m, t, 5, ptr, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof buf,
             "cuTensorMapEncodeTiled failed (%d): dims %llu %llu %llu %llu %llu strides %llu %llu %llu %llu box %u %u "
             "%u %u %u",
             (int)r, (unsigned long long)gd[0], (unsigned long long)gd[1], (unsigned long long)gd[2],
             (unsigned long long)gd[3], (unsigned long long)gd[4], (unsigned long long)gs[0],
             (unsigned long long)gs[1], (unsigned long long)gs[2], (unsigned long long)gs[3], bx[0], bx[1], bx[2],
             bx[3], bx[4]);
    return set_error(-11, buf);
  }
  return 0;
}

template <int OP, int DT, int BN, int EPI, int EWX = 0>
static int launch_one(const GemmKParams& p, int grid, cudaStream_t stream) {
  // The exp-heavy softmax statistics epilogue runs on 8 epilogue warps (two per TMEM lane quadrant).
  // (The store path of the 8-warp configuration is not enabled: on B200 the second warp of a quadrant
  //  intermittently drained zeros for its first 32 columns after a TMA store; see DESIGN.md.  dbg bit 128
  //  selects it for experiments.)
  // (two warps per quadrant split a tile's columns, so the tile must hold at least two 128-byte chunks per row)
  constexpr int EW = ((EPI == 1 || EPI == 2 || EWX == 8) && BN >= 128) ? 8 : 4;
  // smem ring depth: ~190 KB of operand stages, less when the epilogue needs more staging / landing buffers
  constexpr int STAGES = EPI == 1   ? ((BN == 256) ? 2 : (BN == 128 ? 3 : 4))
                         : EPI >= 2 ? ((BN == 256) ? 3 : (BN == 128 ? 5 : 6))
                                    : ((BN == 256) ? (EW == 8 ? 3 : 4) : (BN == 128 ? 6 : 8));
  constexpr int kStageBytes = kStageABytes + BN * 128;
  constexpr int smem = STAGES * kStageBytes + (EPI == 1 ? 4 * EW : 2 * EW) * 4096 + (2 * STAGES + 36) * 8 + 16 + 1024;
  static_assert(smem <= 232448, "shared memory budget exceeded");
  auto kern = gemm_kernel<OP, DT, BN, STAGES, EPI, EW>;
  static bool configured[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return set_error(-12, cudaGetErrorString(e));
    configured[dev] = true;
  }
  kern<<<grid, kThreadsBase + 32 * EW, smem, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(-13, cudaGetErrorString(e));
  return 0;
}

template <int OP, int DT, int EPI>
static int launch_bn(const GemmKParams& p, int bn, int grid, cudaStream_t stream) {
  switch (bn) {
    case 64: return launch_one<OP, DT, 64, EPI>(p, grid, stream);
    case 128: return launch_one<OP, DT, 128, EPI>(p, grid, stream);
    default:
      if constexpr (EPI == 3 && DT == DDP_BF16)
        if (p.dbg & 128) return launch_one<OP, DT, 256, EPI, 8>(p, grid, stream);
      // Shallow-K score slabs (K = head dim) are bound by the epilogue, not the MMA: give them 8 epilogue
      // warps (at the price of one smem stage, irrelevant when a tile has one or two k-blocks).
      // (8 epilogue warps were tried for the plain shallow-K score slab as well: the TMEM drain gets faster,
      //  0.144 vs 0.167 ms, but the slab write gets slower, 0.48 vs 0.42 ms; dbg bit 256 selects it.)
      if constexpr (OP == DDP_OP_NT && EPI == 0)
        if (p.k_blocks <= 2 && (p.dbg & 256)) return launch_one<OP, DT, 256, EPI, 8>(p, grid, stream);
      return launch_one<OP, DT, 256, EPI>(p, grid, stream);
  }
}

template <int OP, int EPI = 0>
static int launch_dt(const GemmKParams& p, int dt, int bn, int grid, cudaStream_t stream) {
  switch (dt) {
    case DDP_F16: return launch_bn<OP, DDP_F16, EPI>(p, bn, grid, stream);
    case DDP_BF16: return launch_bn<OP, DDP_BF16, EPI>(p, bn, grid, stream);
    default: return launch_bn<OP, DDP_F32, EPI>(p, bn, grid, stream);
  }
}

// Widest N tile that still gives every SM a tile; narrow problems fall back to smaller tiles.
static int pick_bn(long long n, long long tiles_per_n_block) {
  const int cands[3] = {256, 128, 64};
  const int sms = ddp_num_sms();
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (i < 2 && n <= cands[i + 1]) continue;  // a narrower tile already covers N
    const long long tiles = tiles_per_n_block * ((n + bn - 1) / bn);
    if (tiles >= sms || i == 2) return bn;
  }
  return 64;
}

struct Shapes {
  long long M, N, K;  // per segment
  long long push_rows;
  int bn, m_blocks, n_blocks, k_blocks, n_chunks;
};

static int derive_shapes(int op, const ddp_tensor* A, const ddp_tensor* B, int world, int chunk_rows, Shapes* s) {
  const int es = elem_size(A->dtype);
  const int bk = 128 / es;
  if (world < 1 || world > DDP_MAX_WORLD) return set_error(-4, "world size out of range");
  if (A->dtype != B->dtype) return set_error(-4, "A and B must have the same dtype");
  if (A->b0 != B->b0 || A->b1 != B->b1) return set_error(-4, "A and B batch extents differ");
  if (op == DDP_OP_NT) {
    if (A->cols != B->cols) return set_error(-4, "nt: inner dimensions differ");
    s->M = A->rows; s->N = B->rows; s->K = A->cols; s->push_rows = B->rows;
  } else if (op == DDP_OP_NN) {
    if (A->cols != world * B->rows) return set_error(-4, "all: left.size(-1) must equal world * right.size(-2)");
    s->M = A->rows; s->N = B->cols; s->K = B->rows; s->push_rows = B->rows;
  } else if (op == DDP_OP_TN) {
    if (A->rows != B->rows) return set_error(-4, "tn: row counts differ");
    if (A->cols % world) return set_error(-4, "tn: left.size(-1) must be divisible by world");
    s->M = A->cols / world; s->N = B->cols; s->K = A->rows; s->push_rows = 0;
  } else {
    return set_error(-4, "unknown op");
  }
  if (s->M < 1 || s->N < 1 || s->K < 1) return set_error(-4, "empty operand");
  s->m_blocks = (int)((s->M + kBlockM - 1) / kBlockM);
  s->bn = pick_bn(s->N, (long long)A->b0 * A->b1 * s->m_blocks * (op == DDP_OP_NN ? 1 : world));
  s->n_blocks = (int)((s->N + s->bn - 1) / s->bn);
  s->k_blocks = (int)((s->K + bk - 1) / bk);
  if (chunk_rows < 1) chunk_rows = 1;
  s->n_chunks = s->push_rows ? (int)((s->push_rows + chunk_rows - 1) / chunk_rows) : 0;
  return 0;
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_gemm_workspace(int op, const ddp_tensor* A, const ddp_tensor* B, int world, int chunk_rows,
                                  size_t* ws_bytes, size_t* n_flags) {
  Shapes s;
  if (int e = derive_shapes(op, A, B, world, chunk_rows, &s)) return e;
  const size_t Bt = (size_t)A->b0 * A->b1;
  if (world == 1) {
    *ws_bytes = 0;
    *n_flags = 0;
    return 0;
  }
  if (op == DDP_OP_TN) {
    *ws_bytes = (size_t)world * Bt * s.M * s.N * 4;
    *n_flags = (size_t)world * Bt * s.m_blocks * s.n_blocks * 4;
  } else {
    *ws_bytes = (size_t)world * Bt * B->rows * B->cols * elem_size(B->dtype);
    *n_flags = (size_t)world * Bt * s.n_chunks;
  }
  return 0;
}

extern "C" int ddp_debug_timeout_site(void) {
  unsigned int v = 0, zero = 0;
  if (cudaMemcpyFromSymbol(&v, ddp::g_ddp_timeout_site, sizeof v) != cudaSuccess) return -1;
  if (v) cudaMemcpyToSymbol(ddp::g_ddp_timeout_site, &zero, sizeof zero);
  return static_cast<int>(v);
}

extern "C" int ddp_gemm_stats_tiles(const ddp_tensor* A, const ddp_tensor* B, int world) {
  Shapes s;
  if (derive_shapes(DDP_OP_NT, A, B, world, 1 << 30, &s)) return -1;
  return world * s.n_blocks * (s.bn >= 128 ? 2 : 1);  // tiles of >= 128 columns are split into two halves
}

extern "C" int ddp_gemm(int op, const ddp_tensor* A, const ddp_tensor* B, const ddp_tensor* C, float alpha,
                        const void* bias, const ddp_comm* comm, int dbg, void* stream_) {
  return ddp_gemm_ex(op, A, B, C, alpha, bias, comm, nullptr, dbg, stream_);
}

extern "C" int ddp_gemm_ex(int op, const ddp_tensor* A, const ddp_tensor* B, const ddp_tensor* C, float alpha,
                           const void* bias, const ddp_comm* comm, const ddp_epilogue* epi, int dbg, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int world = comm ? comm->world : 1;
  const int rank = comm ? comm->rank : 0;
  int chunk_rows = comm ? comm->chunk_rows : 1 << 30;
  if (chunk_rows < 1) chunk_rows = 1;
  Shapes s;
  if (int e = derive_shapes(op, A, B, world, chunk_rows, &s)) return e;
  const int dt = A->dtype;
  const int es = elem_size(dt);
  const uint32_t bk = 128 / es, ch = 128 / es;
  if (!C) return set_error(-4, "C is required (pass the result tensor even for the statistics pass)");
  if (C->dtype != dt) return set_error(-4, "C must have the dtype of A and B");
  if (C->b0 != A->b0 || C->b1 != A->b1) return set_error(-4, "C batch extents differ");
  const long long c_cols = (op == DDP_OP_NT) ? world * s.N : s.N;
  if (C->rows != s.M || C->cols != c_cols) return set_error(-4, "C has the wrong shape");
  if (world > 1 && bias) return set_error(-4, "bias is only supported for world == 1");

  GemmKParams p;
  memset(&p, 0, sizeof p);
  p.M = (int)s.M; p.N = (int)s.N; p.K = (int)s.K;
  p.b0 = (int)A->b0; p.b1 = (int)A->b1;
  p.world = world; p.rank = rank; p.nseg = world;
  p.m_blocks = s.m_blocks; p.n_blocks = s.n_blocks; p.k_blocks = s.k_blocks;
  p.C = C->ptr; p.ldc = C->ld; p.c_sb0 = C->sb0; p.c_sb1 = C->sb1;
  p.alpha = alpha; p.bias = bias; p.dbg = dbg;
  p.c_vec_ok = ((reinterpret_cast<uintptr_t>(C->ptr) & 15) == 0) && ((C->ld * es) % 16 == 0) &&
               ((C->sb0 * es) % 16 == 0) && ((C->sb1 * es) % 16 == 0) &&
               (op != DDP_OP_NT || world == 1 || (s.N * es) % 16 == 0);
  const uint64_t Bt = (uint64_t)A->b0 * A->b1;

  // ---- A
  {
    uint64_t dims[5], strides[4];
    uint32_t box[5] = {1, 1, 1, 1, 1};
    if (op == DDP_OP_TN) {
      if (world > 1 && (s.M * es) % 16)
        return set_error(-2, "tn: T/N * element size must be a multiple of 16 bytes (pad the segments)");
      dims[0] = A->cols; dims[1] = 1; dims[2] = A->rows; dims[3] = A->b0; dims[4] = A->b1;
      strides[0] = A->ld * es; strides[1] = A->ld * es; strides[2] = A->sb0 * es; strides[3] = A->sb1 * es;
      box[0] = ch; box[2] = bk;
    } else {
      // TMA needs a 16-byte aligned innermost start coordinate, so a rank segment must start on one.
      if (op == DDP_OP_NN && world > 1 && (s.K * es) % 16)
        return set_error(-2, "all: T/N * element size must be a multiple of 16 bytes (pad the segments)");
      const bool clipped = (op == DDP_OP_NN) && world > 1;
      if (clipped) {
        dims[0] = s.K; dims[1] = world; strides[0] = s.K * es;
        p.a_seg_c0 = 0; p.a_seg_c1 = 1;
      } else {
        dims[0] = A->cols; dims[1] = 1; strides[0] = A->ld * es;
        p.a_seg_c0 = (op == DDP_OP_NN) ? (int)s.K : 0; p.a_seg_c1 = 0;
      }
      dims[2] = A->rows; dims[3] = A->b0; dims[4] = A->b1;
      strides[1] = A->ld * es; strides[2] = A->sb0 * es; strides[3] = A->sb1 * es;
      box[0] = bk; box[2] = kBlockM;
    }
    if (A->b0 == 1) strides[2] = strides[1];
    if (A->b1 == 1) strides[3] = strides[2];
    if (int e = make_map(&p.tmA, dt, A->ptr, dims, strides, box, op == DDP_OP_TN)) return e;
  }
  // ---- B (local) and B (gathered)
  {
    uint64_t dims[5] = {(uint64_t)B->cols, 1, (uint64_t)B->rows, (uint64_t)B->b0, (uint64_t)B->b1};
    uint64_t strides[4] = {(uint64_t)B->ld * es, (uint64_t)B->ld * es, (uint64_t)B->sb0 * es, (uint64_t)B->sb1 * es};
    uint32_t box[5] = {1, 1, 1, 1, 1};
    if (op == DDP_OP_NT) {
      box[0] = bk; box[2] = s.bn;
    } else {
      box[0] = ch; box[2] = bk;
    }
    if (B->b0 == 1) strides[2] = strides[1];
    if (B->b1 == 1) strides[3] = strides[2];
    if (int e = make_map(&p.tmB, dt, B->ptr, dims, strides, box, op != DDP_OP_NT)) return e;
    if (world > 1 && op != DDP_OP_TN) {
      const uint64_t row_bytes = (uint64_t)B->cols * es;
      if (row_bytes % 16) return set_error(-2, "gathered operand rows must be a multiple of 16 bytes");
      if (reinterpret_cast<uintptr_t>(B->ptr) & 15 || (B->ld * es) % 16 || (B->sb0 * es) % 16 || (B->sb1 * es) % 16)
        return set_error(-2, "gathered operand is not 16-byte aligned");
      uint64_t gd[5] = {(uint64_t)B->cols, 1, (uint64_t)B->rows, Bt, (uint64_t)world};
      uint64_t gs[4] = {row_bytes, row_bytes, row_bytes * B->rows, row_bytes * B->rows * Bt};
      if (int e = make_map(&p.tmBg, dt, comm->ws[rank], gd, gs, box, op != DDP_OP_NT)) return e;
      p.push_src = static_cast<const char*>(B->ptr);
      p.push_ld = B->ld * es; p.push_sb0 = B->sb0 * es; p.push_sb1 = B->sb1 * es;
      p.push_rows = (int)B->rows; p.push_row_bytes = (int)row_bytes;
    } else {
      p.tmBg = p.tmB;
    }
  }
  // ---- C (TMA store path): 32 rows x 128 bytes per box; generic stores cover everything it cannot
  p.c_tma_ok = 0;
  if (p.c_vec_ok && !(op == DDP_OP_TN && world > 1) && !(dbg & 16)) {  // dbg 16: force generic stores
    uint64_t dims[5] = {(uint64_t)c_cols, 1, (uint64_t)C->rows, (uint64_t)C->b0, (uint64_t)C->b1};
    uint64_t strides[4] = {(uint64_t)C->ld * es, (uint64_t)C->ld * es, (uint64_t)C->sb0 * es, (uint64_t)C->sb1 * es};
    uint32_t box[5] = {(uint32_t)(128 / es), 1, 32, 1, 1};
    if (C->b0 == 1) strides[2] = strides[1];
    if (C->b1 == 1) strides[3] = strides[2];
    if (make_map(&p.tmC, dt, C->ptr, dims, strides, box, false) == 0) p.c_tma_ok = 1;
  }
  if (!p.c_tma_ok) p.tmC = p.tmA;
  // ---- fused softmax-backward epilogue (nt only): needs P addressable exactly like C
  const bool fused = epi && epi->mode == DDP_EPI_SOFTMAX_BWD;
  p.tmP = p.tmA;
  if (fused) {
    const ddp_tensor* P = epi->P;
    if (op != DDP_OP_NT) return set_error(-4, "the softmax-backward epilogue belongs to the nt product");
    if (!P || !epi->delta) return set_error(-4, "softmax-backward epilogue: P and delta are required");
    if (P->dtype != dt || P->rows != C->rows || P->cols != C->cols || P->b0 != C->b0 || P->b1 != C->b1)
      return set_error(-4, "softmax-backward epilogue: P must have the shape and dtype of the result");
    if (world > 1 && (s.N * es) % 16) return set_error(-2, "softmax-backward epilogue: T/N rows must be 16-byte aligned");
    uint64_t dims[5] = {(uint64_t)c_cols, 1, (uint64_t)P->rows, (uint64_t)P->b0, (uint64_t)P->b1};
    uint64_t strides[4] = {(uint64_t)P->ld * es, (uint64_t)P->ld * es, (uint64_t)P->sb0 * es, (uint64_t)P->sb1 * es};
    uint32_t box[5] = {(uint32_t)(128 / es), 1, 32, 1, 1};
    if (P->b0 == 1) strides[2] = strides[1];
    if (P->b1 == 1) strides[3] = strides[2];
    if (int e = make_map(&p.tmP, dt, P->ptr, dims, strides, box, false)) return e;
    p.delta = epi->delta;
    p.epi_scale = epi->scale;
  }
  const bool sm_stats = epi && epi->mode == DDP_EPI_SOFTMAX_STATS;
  const bool sm_apply = epi && epi->mode == DDP_EPI_SOFTMAX_APPLY;
  if (sm_stats || sm_apply) {
    if (op != DDP_OP_NT) return set_error(-4, "the softmax-forward epilogues belong to the nt product");
    if (world > 1 && s.N % 32) return set_error(-2, "softmax-forward epilogue: T/N must be a multiple of 32");
    if (sm_stats && !epi->stats) return set_error(-4, "softmax-stats epilogue: stats buffer required");
    if (sm_apply && !epi->rowstat) return set_error(-4, "softmax-apply epilogue: rowstat required");
    p.mask_bits = epi->mask_bits;
    p.mb_ld = epi->mask_ld_words;
    p.mb_sb1 = epi->mask_sb1_words;
    p.stats = reinterpret_cast<float2*>(epi->stats);
    p.rowstat = reinterpret_cast<const float2*>(epi->rowstat);
    p.stats_nt = world * s.n_blocks * (s.bn >= 128 ? 2 : 1);  // two column halves per tile with 8 epilogue warps
    p.epi_scale = epi->scale * 1.4426950408889634f;  // scores in the log2 domain
  }
  if (world > 1) {
    p.seq = comm->seq; p.phase_mask = comm->phase_mask;
    p.chunk_rows = chunk_rows; p.n_chunks = s.n_chunks;
    for (int i = 0; i < world; ++i) {
      p.ws[i] = static_cast<char*>(comm->ws[i]);
      p.mc_ws = static_cast<char*>(comm->mc_ws);
      p.flags[i] = comm->flags[i];
      p.done[i] = comm->done[i];
    }
    p.exit_counter = comm->exit_counter;
    p.status = comm->status;
  } else {
    p.phase_mask = 3; p.chunk_rows = 1 << 30; p.n_chunks = 1;
  }
  const long long tiles = (long long)Bt * s.m_blocks * s.n_blocks * (op == DDP_OP_NN ? 1 : world);
  const int sms = ddp_num_sms();
  int grid = world > 1 ? sms : (int)(tiles < sms ? tiles : sms);
  if (grid < 1) grid = 1;
  switch (op) {
    case DDP_OP_NT:
      if (fused) return launch_dt<DDP_OP_NT, 1>(p, dt, s.bn, grid, stream);
      if (sm_stats) return launch_dt<DDP_OP_NT, 2>(p, dt, s.bn, grid, stream);
      if (sm_apply) return launch_dt<DDP_OP_NT, 3>(p, dt, s.bn, grid, stream);
      return launch_dt<DDP_OP_NT>(p, dt, s.bn, grid, stream);
    case DDP_OP_NN: return launch_dt<DDP_OP_NN>(p, dt, s.bn, grid, stream);
    default: return launch_dt<DDP_OP_TN>(p, dt, s.bn, grid, stream);
  }
}
