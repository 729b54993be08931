// Host side of the slab-free attention forward: shape checks, TMA tensor maps, launch.
// C ABI declared in include/ddp_b200.h.
#include <cstdlib>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstring>

#include "../common/host_util.h"
#include "../gemm/tma_map.h"
#include "attn_bwd_sm100.cuh"
#include "attn_fwd2_sm100.cuh"
#include "attn_bwd2_sm100.cuh"

namespace ddp {

constexpr int kAtStages = 3;
constexpr int kAtSmem = 2 * kAtTile + 2 * kAtPBytes + kAtStages * kAtStageBytes + (2 * kAtStages + 22) * 8 + 16 + 4 * 128 * 8 + 1024;
static_assert(kAtSmem <= 232448, "shared memory budget exceeded");

static int attn_check(const ddp_tensor* A, const ddp_tensor* B, const ddp_tensor* V, int world) {
  if (!A || !B || !V) return set_error(-4, "attn: A, B and V are required");
  if (world < 1 || world > DDP_MAX_WORLD) return set_error(-4, "world size out of range");
  const int dt = A->dtype;
  if (dt != DDP_BF16 && dt != DDP_F16) return set_error(-6, "attn: bf16 / fp16 operands only");
  if (B->dtype != dt || V->dtype != dt) return set_error(-4, "attn: operand dtypes differ");
  if (A->cols != B->cols) return set_error(-4, "attn: A and B inner dimensions differ");
  if (A->cols < 8 || A->cols > 64 || A->cols % 8) return set_error(-6, "attn: head dim must be a multiple of 8, <= 64");
  if (V->cols < 8 || V->cols > 64 || V->cols % 8) return set_error(-6, "attn: value dim must be a multiple of 8, <= 64");
  if (B->rows != V->rows) return set_error(-4, "attn: B and V row counts differ");
  if (A->b0 != B->b0 || A->b1 != B->b1 || A->b0 != V->b0 || A->b1 != V->b1)
    return set_error(-4, "attn: batch extents differ");
  if (A->rows < 1 || B->rows < 1) return set_error(-4, "attn: empty operand");
  return 0;
}

static int tile_map(CUtensorMap* m, const ddp_tensor* t, uint32_t box_rows, bool mn_major) {
  const uint64_t es = 2;
  uint64_t dims[5] = {(uint64_t)t->cols, 1, (uint64_t)t->rows, (uint64_t)t->b0, (uint64_t)t->b1};
  uint64_t strides[4] = {(uint64_t)t->ld * es, (uint64_t)t->ld * es, (uint64_t)t->sb0 * es, (uint64_t)t->sb1 * es};
  uint32_t box[5] = {64, 1, box_rows, 1, 1};
  if (t->b0 == 1) strides[2] = strides[1];
  if (t->b1 == 1) strides[3] = strides[2];
  return make_map(m, t->dtype, t->ptr, dims, strides, box, mn_major);
}

// DDP_ATTN_DBG=<bits> ORs experiment bits into the dbg argument of the three attention entry points (read once).
static int attn_env_dbg() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DDP_ATTN_DBG");
    v = e ? atoi(e) : 0;
  }
  return v;
}

static long long* g_attn_trace = nullptr;  // ddp_attn_debug_trace

constexpr int kAt2Stages = 4;
constexpr int kAt2Smem = 2 * kAtTile + kAt2StageO + kAt2Stages * kAt2StageBytes + (2 * kAt2Stages + 26) * 8 + 16 + 4 * 128 * 8 + 1024;
static_assert(kAt2Smem <= 232448, "shared memory budget exceeded");

// Second-generation forward (P stays in tensor memory), see attn_fwd2_sm100.cuh.
template <int DT>
static int attn_launch2(const AttnKParams& p, int grid, cudaStream_t stream) {
  auto kern = (p.dbg & 1) ? attn_fwd2_kernel<DT, kAt2Stages, true> : attn_fwd2_kernel<DT, kAt2Stages, false>;
  static bool configured[2][64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 64 && !configured[p.dbg & 1][dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kAt2Smem);
    if (e != cudaSuccess) return set_error(-12, cudaGetErrorString(e));
    configured[p.dbg & 1][dev] = true;
  }
  kern<<<grid, kAtThreads, kAt2Smem, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(-13, cudaGetErrorString(e));
  return 0;
}

template <int DT>
static int attn_launch(const AttnKParams& p, int grid, cudaStream_t stream) {
  auto kern = (p.dbg & 1) ? attn_fwd_kernel<DT, kAtStages, true> : attn_fwd_kernel<DT, kAtStages, false>;
  static bool configured[2][64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 64 && !configured[p.dbg & 1][dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kAtSmem);
    if (e != cudaSuccess) return set_error(-12, cudaGetErrorString(e));
    configured[p.dbg & 1][dev] = true;
  }
  kern<<<grid, kAtThreads, kAtSmem, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(-13, cudaGetErrorString(e));
  return 0;
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_attn_fwd_workspace(const ddp_tensor* B, const ddp_tensor* V, int world, int chunk_rows,
                                      size_t* ws_bytes, size_t* n_flags) {
  if (!B || !V) return set_error(-4, "attn: B and V are required");
  if (world <= 1) {
    *ws_bytes = 0;
    *n_flags = 0;
    return 0;
  }
  if (chunk_rows < 1) chunk_rows = 1;
  const size_t Bt = (size_t)B->b0 * B->b1;
  const size_t n_chunks = ((size_t)B->rows + chunk_rows - 1) / chunk_rows;
  const size_t reg_b = ((size_t)world * Bt * B->rows * B->cols * 2 + 1023) & ~(size_t)1023;
  const size_t reg_v = (size_t)world * Bt * V->rows * V->cols * 2;
  *ws_bytes = reg_b + reg_v;
  *n_flags = 2 * (size_t)world * Bt * n_chunks;
  return 0;
}

extern "C" void ddp_attn_debug_trace(long long* device_buf) { g_attn_trace = device_buf; }
extern "C" int ddp_attn_debug_trace_dims(int* actors, int* steps, int* events) {
  *actors = 8; *steps = kAtTraceSteps; *events = kAtTraceEvents;
  return 0;
}

extern "C" int ddp_debug_timeout_site_attn(void) {
  unsigned int v = 0, zero = 0;
  if (cudaMemcpyFromSymbol(&v, ddp::g_ddp_timeout_site, sizeof v) != cudaSuccess) return -1;
  if (v) cudaMemcpyToSymbol(ddp::g_ddp_timeout_site, &zero, sizeof zero);
  return static_cast<int>(v);
}

extern "C" int ddp_attn_fwd(const ddp_tensor* A, const ddp_tensor* B, const ddp_tensor* V, const ddp_tensor* O,
                            float* lse, const uint32_t* mask_bits, int64_t mask_ld_words, int64_t mask_sb1_words,
                            float scale, const ddp_comm* comm, int dbg, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int world = comm ? comm->world : 1;
  const int rank = comm ? comm->rank : 0;
  if (int e = attn_check(A, B, V, world)) return e;
  if (!O) return set_error(-4, "attn: O is required");
  const int dt = A->dtype;
  if (O->dtype != dt || O->rows != A->rows || O->cols != V->cols || O->b0 != A->b0 || O->b1 != A->b1)
    return set_error(-4, "attn: O has the wrong shape or dtype");
  if (!(scale > 0.f)) return set_error(-4, "attn: scale must be positive");
  if (mask_bits && world > 1 && (B->rows % 32)) return set_error(-6, "attn: masked, world > 1: T/N must be a multiple of 32");
  int chunk_rows = comm ? comm->chunk_rows : 1 << 30;
  if (chunk_rows < 1) chunk_rows = 1;

  AttnKParams p;
  memset(&p, 0, sizeof p);
  p.M = (int)A->rows; p.Tn = (int)B->rows; p.dv = (int)V->cols;
  p.b0 = (int)A->b0; p.b1 = (int)A->b1;
  p.world = world; p.rank = rank;
  const int m_blocks = (p.M + kAtBM - 1) / kAtBM;
  p.m_pairs = (m_blocks + 1) / 2;
  p.n_blocks = (p.Tn + kAtBN - 1) / kAtBN;
  p.O = O->ptr; p.ldo = O->ld; p.o_sb0 = O->sb0; p.o_sb1 = O->sb1;
  p.lse = lse;
  p.mask_bits = mask_bits; p.mb_ld = mask_ld_words; p.mb_sb1 = mask_sb1_words;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.dbg = dbg | attn_env_dbg();
  p.trace = g_attn_trace;
  const uint64_t Bt = (uint64_t)A->b0 * A->b1;

  if (int e = tile_map(&p.tmA, A, kAtBM, false)) return e;
  if (int e = tile_map(&p.tmB, B, kAtBN, false)) return e;
  if (int e = tile_map(&p.tmV, V, kAtBN, true)) return e;
  p.o_vec_ok = ((reinterpret_cast<uintptr_t>(O->ptr) & 15) == 0) && ((O->ld * 2) % 16 == 0) &&
               ((O->sb0 * 2) % 16 == 0) && ((O->sb1 * 2) % 16 == 0);
  p.o_tma_ok = 0;
  if (p.o_vec_ok && !(dbg & 16)) {
    if (tile_map(&p.tmO, O, 32, false) == 0) p.o_tma_ok = 1;
  }
  if (!p.o_tma_ok) p.tmO = p.tmA;
  p.tmBg = p.tmB;
  p.tmVg = p.tmV;
  if (world > 1) {
    const ddp_tensor* src[2] = {B, V};
    size_t wsb = 0, nf = 0;
    ddp_attn_fwd_workspace(B, V, world, chunk_rows, &wsb, &nf);
    p.ws_v_off = (long long)(((size_t)world * Bt * B->rows * B->cols * 2 + 1023) & ~(size_t)1023);
    for (int i = 0; i < 2; ++i) {
      const ddp_tensor* t = src[i];
      const uint64_t row_bytes = (uint64_t)t->cols * 2;
      if (reinterpret_cast<uintptr_t>(t->ptr) & 15 || (t->ld * 2) % 16 || (t->sb0 * 2) % 16 || (t->sb1 * 2) % 16)
        return set_error(-2, "attn: gathered operand is not 16-byte aligned");
      uint64_t gd[5] = {(uint64_t)t->cols, 1, (uint64_t)t->rows, Bt, (uint64_t)world};
      uint64_t gs[4] = {row_bytes, row_bytes, row_bytes * t->rows, row_bytes * t->rows * Bt};
      uint32_t box[5] = {64, 1, (uint32_t)kAtBN, 1, 1};
      char* base = static_cast<char*>(comm->ws[rank]) + (i ? p.ws_v_off : 0);
      if (int e = make_map(i ? &p.tmVg : &p.tmBg, dt, base, gd, gs, box, i == 1)) return e;
      p.push_src[i] = static_cast<const char*>(t->ptr);
      p.push_ld[i] = t->ld * 2; p.push_sb0[i] = t->sb0 * 2; p.push_sb1[i] = t->sb1 * 2;
      p.push_row_bytes[i] = (int)row_bytes;
    }
    p.push_rows = (int)B->rows;
    p.seq = comm->seq; p.phase_mask = comm->phase_mask;
    p.chunk_rows = chunk_rows;
    p.n_chunks = (int)((B->rows + chunk_rows - 1) / chunk_rows);
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
  const long long items = (long long)p.b1 * p.m_pairs * p.b0;
  const int sms = ddp_num_sms();
  int grid = world > 1 ? sms : (int)(items < sms ? items : sms);
  if (grid < 1) grid = 1;
  if (!(p.dbg & 64))  // dbg bit 64: the first-generation kernel (P staged through shared memory)
    return dt == DDP_BF16 ? attn_launch2<DDP_BF16>(p, grid, stream) : attn_launch2<DDP_F16>(p, grid, stream);
  return dt == DDP_BF16 ? attn_launch<DDP_BF16>(p, grid, stream) : attn_launch<DDP_F16>(p, grid, stream);
}

// ---------------------------------------------------------------------------------------------- backward
namespace ddp {

constexpr int kAbRowsStages = 6;
constexpr int kAbRowsSmem = 6 * kAtTile + kAbRowsStages * 2 * kAbHalfTile + (2 * kAbRowsStages + 10) * 8 + 16 + 1024;
static_assert(kAbRowsSmem <= 232448, "shared memory budget exceeded");

template <typename Kern>
static int attn_bwd_launch(Kern kern, bool* configured, int smem, const AttnBwdParams& p, int grid, cudaStream_t stream) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return set_error(-12, cudaGetErrorString(e));
    configured[dev] = true;
  }
  kern<<<grid, kAtThreads, smem, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(-13, cudaGetErrorString(e));
  return 0;
}

// Fills the collective part of the parameters (gather of B and V) - shared by both backward kernels.
static int attn_bwd_comm(AttnBwdParams& p, const ddp_tensor* B, const ddp_tensor* V, const ddp_comm* comm,
                         uint32_t box_rows) {
  const int world = p.world, rank = p.rank;
  const uint64_t Bt = (uint64_t)B->b0 * B->b1;
  int chunk_rows = comm->chunk_rows < 1 ? 1 : comm->chunk_rows;
  const ddp_tensor* src[2] = {B, V};
  p.ws_v_off = (long long)(((size_t)world * Bt * B->rows * B->cols * 2 + 1023) & ~(size_t)1023);
  for (int i = 0; i < 2; ++i) {
    const ddp_tensor* t = src[i];
    const uint64_t row_bytes = (uint64_t)t->cols * 2;
    if (reinterpret_cast<uintptr_t>(t->ptr) & 15 || (t->ld * 2) % 16 || (t->sb0 * 2) % 16 || (t->sb1 * 2) % 16)
      return set_error(-2, "attn: gathered operand is not 16-byte aligned");
    uint64_t gd[5] = {(uint64_t)t->cols, 1, (uint64_t)t->rows, Bt, (uint64_t)world};
    uint64_t gs[4] = {row_bytes, row_bytes, row_bytes * t->rows, row_bytes * t->rows * Bt};
    uint32_t box[5] = {64, 1, box_rows, 1, 1};
    // second-generation kernels only: a still valid gather of the same operands (ddp_comm::gather_ws) is read in place
    const bool reuse = comm->gather_ws != nullptr && !(p.dbg & 64);
    p.gather_ready = reuse ? 1 : 0;
    char* base = (reuse ? static_cast<char*>(const_cast<void*>(comm->gather_ws)) : static_cast<char*>(comm->ws[rank])) +
                 (i ? p.ws_v_off : 0);
    if (int e = make_map(i ? &p.tmVg : &p.tmBg, t->dtype, base, gd, gs, box, false)) return e;
    p.push_src[i] = static_cast<const char*>(t->ptr);
    p.push_ld[i] = t->ld * 2; p.push_sb0[i] = t->sb0 * 2; p.push_sb1[i] = t->sb1 * 2;
    p.push_row_bytes[i] = (int)row_bytes;
  }
  p.push_rows = (int)B->rows;
  p.seq = comm->seq; p.phase_mask = comm->phase_mask;
  p.chunk_rows = chunk_rows;
  p.n_chunks = (int)((B->rows + chunk_rows - 1) / chunk_rows);
  for (int i = 0; i < world; ++i) {
    p.ws[i] = static_cast<char*>(comm->ws[i]);
    p.mc_ws = static_cast<char*>(comm->mc_ws);
    p.flags[i] = comm->flags[i];
    p.done[i] = comm->done[i];
  }
  p.exit_counter = comm->exit_counter;
  p.status = comm->status;
  return 0;
}

static bool out_vec_ok(const ddp_tensor* O) {
  return ((reinterpret_cast<uintptr_t>(O->ptr) & 15) == 0) && ((O->ld * 2) % 16 == 0) && ((O->sb0 * 2) % 16 == 0) &&
         ((O->sb1 * 2) % 16 == 0);
}

}  // namespace ddp

extern "C" int ddp_attn_bwd_rows(const ddp_tensor* A, const ddp_tensor* G, const ddp_tensor* B, const ddp_tensor* V,
                                 const ddp_tensor* dA, const float* neg_lse, const float* nds, int64_t stat_ld,
                                 const uint32_t* mask_bits, int64_t mask_ld_words, int64_t mask_sb1_words, float scale,
                                 const ddp_comm* comm, int dbg, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int world = comm ? comm->world : 1;
  const int rank = comm ? comm->rank : 0;
  if (int e = attn_check(A, B, V, world)) return e;
  if (!G || !dA || !neg_lse || !nds) return set_error(-4, "attn bwd: dO, dA, neg_lse and nds are required");
  const int dt = A->dtype;
  if (G->dtype != dt || G->rows != A->rows || G->cols != V->cols || G->b0 != A->b0 || G->b1 != A->b1)
    return set_error(-4, "attn bwd: dO has the wrong shape or dtype");
  if (dA->dtype != dt || dA->rows != A->rows || dA->cols != A->cols || dA->b0 != A->b0 || dA->b1 != A->b1)
    return set_error(-4, "attn bwd: dA has the wrong shape or dtype");
  if (!(scale > 0.f)) return set_error(-4, "attn: scale must be positive");
  if (stat_ld < A->rows) return set_error(-4, "attn bwd: stat_ld < rows");
  if (mask_bits && world > 1 && (B->rows % 32)) return set_error(-6, "attn: masked, world > 1: T/N must be a multiple of 32");

  AttnBwdParams p;
  memset(&p, 0, sizeof p);
  p.M = (int)A->rows; p.Tn = (int)B->rows; p.d = (int)A->cols; p.dv = (int)V->cols;
  p.b0 = (int)A->b0; p.b1 = (int)A->b1;
  p.world = world; p.rank = rank;
  const int m_blocks = (p.M + kAtBM - 1) / kAtBM;
  p.m_pairs = (m_blocks + 1) / 2;
  p.n_steps = (p.Tn + kAbStep - 1) / kAbStep;
  p.out0 = dA->ptr; p.ld0 = dA->ld; p.o0_sb0 = dA->sb0; p.o0_sb1 = dA->sb1;
  p.neg_lse = neg_lse; p.nds = nds; p.stat_ld = stat_ld;
  p.mask_bits = mask_bits; p.mb_ld = mask_ld_words; p.mb_sb1 = mask_sb1_words;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  p.dbg = dbg | attn_env_dbg();
  if (int e = tile_map(&p.tmA, A, kAtBM, false)) return e;
  if (int e = tile_map(&p.tmG, G, kAtBM, false)) return e;
  if (int e = tile_map(&p.tmB, B, kAbStep, false)) return e;
  if (int e = tile_map(&p.tmV, V, kAbStep, false)) return e;
  p.o0_vec_ok = out_vec_ok(dA);
  if (p.o0_vec_ok && !(dbg & 16) && tile_map(&p.tmOut0, dA, 32, false) == 0) p.o0_tma_ok = 1;
  if (!p.o0_tma_ok) p.tmOut0 = p.tmA;
  p.tmOut1 = p.tmA;
  p.tmBg = p.tmB;
  p.tmVg = p.tmV;
  if (world > 1) {
    if (int e = attn_bwd_comm(p, B, V, comm, kAbStep)) return e;
  } else {
    p.phase_mask = 3; p.chunk_rows = 1 << 30; p.n_chunks = 1;
  }
  const int sms = ddp_num_sms();
  if (!(p.dbg & 64)) {  // second generation: one 128-row block per item, dS stays in tensor memory
    constexpr int kSmem2 = 3 * kAtTile + kAb2Stages * kAb2RowsStage + (2 * kAb2Stages + 12) * 8 + 16 + 1024;
    static_assert(kSmem2 <= 232448, "shared memory budget exceeded");
    const long long items2 = (long long)p.b1 * m_blocks * p.b0;
    int grid2 = world > 1 ? sms : (int)(items2 < sms ? items2 : sms);
    if (grid2 < 1) grid2 = 1;
    static bool cfg2_bf[64] = {}, cfg2_h[64] = {};
    return dt == DDP_BF16 ? attn_bwd_launch(attn_bwd2_rows_kernel<DDP_BF16>, cfg2_bf, kSmem2, p, grid2, stream)
                          : attn_bwd_launch(attn_bwd2_rows_kernel<DDP_F16>, cfg2_h, kSmem2, p, grid2, stream);
  }
  const long long items = (long long)p.b1 * p.m_pairs * p.b0;
  int grid = world > 1 ? sms : (int)(items < sms ? items : sms);
  if (grid < 1) grid = 1;
  static bool cfg_bf[64] = {}, cfg_h[64] = {};
  return dt == DDP_BF16
             ? attn_bwd_launch(attn_bwd_rows_kernel<DDP_BF16, kAbRowsStages>, cfg_bf, kAbRowsSmem, p, grid, stream)
             : attn_bwd_launch(attn_bwd_rows_kernel<DDP_F16, kAbRowsStages>, cfg_h, kAbRowsSmem, p, grid, stream);
}

// ---------------------------------------------------------------------------------------------- column side
namespace ddp {

constexpr int kAbColsStages = 5;
constexpr int kAbColsSmem = 8 * kAtTile + kAbColsStages * kAbColsStage + (2 * kAbColsStages + 10) * 8 + 16 + 1024;
static_assert(kAbColsSmem <= 232448, "shared memory budget exceeded");

struct ColsLayout {
  size_t v_off, p0_off, p1_off, bytes, gather_flags, flags;
};
static ColsLayout cols_layout(const ddp_tensor* B, const ddp_tensor* V, int world, int chunk_rows) {
  ColsLayout L;
  const size_t Bt = (size_t)B->b0 * B->b1;
  const size_t al = 1023;
  if (chunk_rows < 1) chunk_rows = 1;
  const size_t n_chunks = ((size_t)B->rows + chunk_rows - 1) / chunk_rows;
  const size_t n_mb = ((size_t)B->rows + kAtBM - 1) / kAtBM;
  L.v_off = ((size_t)world * Bt * B->rows * B->cols * 2 + al) & ~al;
  L.p0_off = (L.v_off + (size_t)world * Bt * V->rows * V->cols * 2 + al) & ~al;
  L.p1_off = (L.p0_off + (size_t)world * Bt * V->rows * V->cols * 4 + al) & ~al;
  L.bytes = L.p1_off + (size_t)world * Bt * B->rows * B->cols * 4;
  L.gather_flags = 2 * (size_t)world * Bt * n_chunks;
  L.flags = L.gather_flags + (size_t)world * Bt * n_mb * 16;  // one per (source, batch, block, quadrant, part)
  return L;
}

// Bit-packed masks of the attention kernels, in a layout whose loads coalesce: the kernels' threads own one row
// (row-side kernels) or one column (column-side kernel) each and read one 32-bit word per 32 elements of the
// other axis, so the words of 32 neighbouring rows (columns) are interleaved:
//   plain:      bits[b1][ceil(rows/32)][ld_words][32]   word w of row r   at ((r >> 5) * ld_words + w) * 32 + (r & 31)
//               bit k of that word <- mask[b1][r][32 w + k] != 0
//   transposed: bits[b1][ceil(cols/32)][ld_words][32]   word w of column c at ((c >> 5) * ld_words + w) * 32 + (c & 31)
//               bit k of that word <- mask[b1][32 w + k][c] != 0
// Elements outside the mask are packed as "masked".
__global__ void pack_mask_il_kernel(const uint8_t* __restrict__ mask, long long b1, long long rows, long long cols,
                                    long long m_ld, long long m_sb1, uint32_t* __restrict__ bits, long long ld_words,
                                    int transposed) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (!transposed) {
    const long long rows_p = (rows + 31) / 32 * 32;
    if (idx >= b1 * rows_p * ld_words) return;
    const long long w = idx % ld_words;  // adjacent threads -> adjacent words of one row (coalesced mask reads)
    long long r = idx / ld_words;
    const long long b1i = r / rows_p;
    r -= b1i * rows_p;
    uint32_t word = 0xffffffffu;
    if (r < rows) {
      const uint8_t* src = mask + b1i * m_sb1 + r * m_ld + w * 32;
      word = 0;
      if (w * 32 + 32 <= cols && (reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        const uint4 x = *reinterpret_cast<const uint4*>(src);
        const uint4 y = *reinterpret_cast<const uint4*>(src + 16);
        const uint32_t v[8] = {x.x, x.y, x.z, x.w, y.x, y.y, y.z, y.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if ((v[i] >> (8 * k)) & 0xffu) word |= 1u << (4 * i + k);
      } else {
        for (int k = 0; k < 32; ++k) {
          const long long c = w * 32 + k;
          if (c >= cols || src[k] != 0) word |= 1u << k;
        }
      }
    }
    bits[(b1i * (rows_p / 32) + (r >> 5)) * ld_words * 32 + w * 32 + (r & 31)] = word;
  } else {
    const long long cols_p = (cols + 31) / 32 * 32;
    if (idx >= b1 * cols_p * ld_words) return;
    const long long c = idx % cols_p;  // adjacent threads -> adjacent columns (coalesced mask reads)
    long long r = idx / cols_p;
    const long long w = r % ld_words;
    const long long b1i = r / ld_words;
    uint32_t word = 0xffffffffu;
    if (c < cols) {
      const uint8_t* src = mask + b1i * m_sb1 + c;
      word = 0;
      for (int k = 0; k < 32; ++k) {
        const long long row = w * 32 + k;
        if (row >= rows || src[row * m_ld] != 0) word |= 1u << k;
      }
    }
    bits[(b1i * (cols_p / 32) + (c >> 5)) * ld_words * 32 + w * 32 + (c & 31)] = word;
  }
}

}  // namespace ddp

// Row statistics of the slab-free backward in one launch: delta = rowsum(dO * O) per (row, head), then
//   neg_lse[b][h][r] = -L            nds[b][h][r] = -scale * delta          (non-finite values -> 0: fully masked rows)
// for r < rows and 0 for the padding rows r in [rows, stat_ld) - the two arrays the backward kernels stream.
namespace ddp {
template <typename T>
__global__ void attn_bwd_stats_kernel(const T* __restrict__ g, const T* __restrict__ o, const float* __restrict__ lse,
                                      float* __restrict__ neg_lse, float* __restrict__ nds, long long rows, int cols,
                                      long long g_ld, long long g_sb0, long long g_sb1, long long o_ld, long long o_sb0,
                                      long long o_sb1, int b0, long long n_bh, long long stat_ld, float scale) {
  const long long wid = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (wid >= n_bh * stat_ld) return;
  const int lane = threadIdx.x & 31;
  const long long bh = wid / stat_ld, r = wid - bh * stat_ld;
  float nl = 0.f, nd = 0.f;
  if (r < rows) {
    const long long b1i = bh / b0, b0i = bh - b1i * b0;
    const T* gp = g + b1i * g_sb1 + b0i * g_sb0 + r * g_ld;
    const T* op = o + b1i * o_sb1 + b0i * o_sb0 + r * o_ld;
    float acc = 0.f;
    for (int c = lane; c < cols; c += 32) acc += static_cast<float>(gp[c]) * static_cast<float>(op[c]);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    nd = -scale * acc;
    nl = -lse[bh * rows + r];
    if (!(fabsf(nd) <= 3.0e38f)) nd = 0.f;  // NaN / inf
    if (!(fabsf(nl) <= 3.0e38f)) nl = 0.f;
  }
  if (lane == 0) {
    neg_lse[wid] = nl;
    nds[wid] = nd;
  }
}
// The same with 16-byte loads: 8 lanes per row (8 elements each, cols <= 64), four rows per warp.
template <typename T>
__global__ void attn_bwd_stats_vec_kernel(const T* __restrict__ g, const T* __restrict__ o, const float* __restrict__ lse,
                                          float* __restrict__ neg_lse, float* __restrict__ nds, long long rows, int cols,
                                          long long g_ld, long long g_sb0, long long g_sb1, long long o_ld,
                                          long long o_sb0, long long o_sb1, int b0, long long n_bh, long long stat_ld,
                                          float scale) {
  const long long idx = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 3;  // (bh, row)
  const int sub = threadIdx.x & 7;
  const bool in_range = idx < n_bh * stat_ld;
  const long long bh = in_range ? idx / stat_ld : 0, r = in_range ? idx - bh * stat_ld : rows;
  float acc = 0.f;
  if (r < rows && sub * 8 < cols) {
    const long long b1i = bh / b0, b0i = bh - b1i * b0;
    const uint4 gv = *reinterpret_cast<const uint4*>(g + b1i * g_sb1 + b0i * g_sb0 + r * g_ld + sub * 8);
    const uint4 ov = *reinterpret_cast<const uint4*>(o + b1i * o_sb1 + b0i * o_sb0 + r * o_ld + sub * 8);
    const T* ge = reinterpret_cast<const T*>(&gv);
    const T* oe = reinterpret_cast<const T*>(&ov);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += static_cast<float>(ge[i]) * static_cast<float>(oe[i]);
  }
#pragma unroll
  for (int off = 4; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (in_range && sub == 0) {
    float nl = 0.f, nd = 0.f;
    if (r < rows) {
      nd = -scale * acc;
      nl = -lse[bh * rows + r];
      if (!(fabsf(nd) <= 3.0e38f)) nd = 0.f;  // NaN / inf
      if (!(fabsf(nl) <= 3.0e38f)) nl = 0.f;
    }
    neg_lse[idx] = nl;
    nds[idx] = nd;
  }
}
}  // namespace ddp

extern "C" int ddp_attn_bwd_stats(const ddp_tensor* G, const ddp_tensor* O, const float* lse, float* neg_lse, float* nds,
                                  int64_t stat_ld, float scale, void* stream_) {
  if (!G || !O || !lse || !neg_lse || !nds) return set_error(-4, "attn bwd stats: null argument");
  if (G->dtype != O->dtype || G->rows != O->rows || G->cols != O->cols || G->b0 != O->b0 || G->b1 != O->b1)
    return set_error(-4, "attn bwd stats: dO and O differ in shape or dtype");
  if (stat_ld < G->rows) return set_error(-4, "attn bwd stats: stat_ld < rows");
  const long long n_bh = (long long)G->b0 * G->b1;
  const long long warps = n_bh * stat_ld;
  const int wpb = 8;
  const unsigned grid = (unsigned)((warps + wpb - 1) / wpb);
  cudaStream_t st = static_cast<cudaStream_t>(stream_);
  auto vec_ok = [](const ddp_tensor* t) {
    return !(reinterpret_cast<uintptr_t>(t->ptr) & 15) && t->ld % 8 == 0 && t->sb0 % 8 == 0 && t->sb1 % 8 == 0;
  };
  if (G->cols % 8 == 0 && G->cols <= 64 && vec_ok(G) && vec_ok(O) && (G->dtype == DDP_BF16 || G->dtype == DDP_F16)) {
    const long long threads = warps * 8;
    const unsigned vgrid = (unsigned)((threads + 255) / 256);
    if (G->dtype == DDP_BF16)
      ddp::attn_bwd_stats_vec_kernel<__nv_bfloat16><<<vgrid, 256, 0, st>>>(
          (const __nv_bfloat16*)G->ptr, (const __nv_bfloat16*)O->ptr, lse, neg_lse, nds, G->rows, (int)G->cols, G->ld,
          G->sb0, G->sb1, O->ld, O->sb0, O->sb1, (int)G->b0, n_bh, stat_ld, scale);
    else
      ddp::attn_bwd_stats_vec_kernel<__half><<<vgrid, 256, 0, st>>>(
          (const __half*)G->ptr, (const __half*)O->ptr, lse, neg_lse, nds, G->rows, (int)G->cols, G->ld, G->sb0, G->sb1,
          O->ld, O->sb0, O->sb1, (int)G->b0, n_bh, stat_ld, scale);
    return ddp::cuda_check(cudaGetLastError(), "attn_bwd_stats launch");
  }
  if (G->dtype == DDP_BF16)
    ddp::attn_bwd_stats_kernel<__nv_bfloat16><<<grid, wpb * 32, 0, st>>>(
        (const __nv_bfloat16*)G->ptr, (const __nv_bfloat16*)O->ptr, lse, neg_lse, nds, G->rows, (int)G->cols, G->ld, G->sb0,
        G->sb1, O->ld, O->sb0, O->sb1, (int)G->b0, n_bh, stat_ld, scale);
  else if (G->dtype == DDP_F16)
    ddp::attn_bwd_stats_kernel<__half><<<grid, wpb * 32, 0, st>>>((const __half*)G->ptr, (const __half*)O->ptr, lse, neg_lse,
                                                                  nds, G->rows, (int)G->cols, G->ld, G->sb0, G->sb1, O->ld,
                                                                  O->sb0, O->sb1, (int)G->b0, n_bh, stat_ld, scale);
  else
    return set_error(-4, "attn bwd stats: 16-bit tensors only");
  return ddp::cuda_check(cudaGetLastError(), "attn_bwd_stats launch");
}

extern "C" size_t ddp_attn_mask_words(int64_t b1, int64_t rows, int64_t cols, int transposed) {
  const int64_t own = transposed ? cols : rows, other = transposed ? rows : cols;
  return (size_t)b1 * (size_t)((own + 31) / 32) * (size_t)((other + 31) / 32) * 32;
}

extern "C" int ddp_attn_pack_mask(const uint8_t* mask, int64_t b1, int64_t rows, int64_t cols, int64_t mask_ld,
                                  int64_t mask_sb1, uint32_t* bits, int transposed, void* stream) {
  const long long own = transposed ? cols : rows, other = transposed ? rows : cols;
  const long long ld_words = (other + 31) / 32;
  const long long total = b1 * ((own + 31) / 32 * 32) * ld_words;
  if (total <= 0) return 0;
  const int threads = 256;
  const long long blocks = (total + threads - 1) / threads;
  ddp::pack_mask_il_kernel<<<static_cast<unsigned>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      mask, b1, rows, cols, mask_ld, mask_sb1, bits, ld_words, transposed);
  return ddp::cuda_check(cudaGetLastError(), "attn_pack_mask launch");
}

extern "C" int ddp_attn_bwd_cols_workspace(const ddp_tensor* B, const ddp_tensor* V, int world, int chunk_rows,
                                           size_t* ws_bytes, size_t* n_flags) {
  if (!B || !V) return set_error(-4, "attn: B and V are required");
  if (world <= 1) {
    *ws_bytes = 0;
    *n_flags = 0;
    return 0;
  }
  const ColsLayout L = cols_layout(B, V, world, chunk_rows);
  *ws_bytes = L.bytes;
  *n_flags = L.flags;
  return 0;
}

extern "C" int ddp_attn_bwd_cols(const ddp_tensor* A, const ddp_tensor* G, const ddp_tensor* B, const ddp_tensor* V,
                                 const ddp_tensor* dV, const ddp_tensor* dB, const float* neg_lse, const float* nds,
                                 int64_t stat_ld, const uint32_t* mask_bits_t, int64_t mask_ld_words,
                                 int64_t mask_sb1_words, float scale, const ddp_comm* comm, int dbg, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int world = comm ? comm->world : 1;
  const int rank = comm ? comm->rank : 0;
  if (int e = attn_check(A, B, V, world)) return e;
  if (!G || !dV || !dB || !neg_lse || !nds) return set_error(-4, "attn bwd: dO, dV, dB, neg_lse and nds are required");
  const int dt = A->dtype;
  if (G->dtype != dt || G->rows != A->rows || G->cols != V->cols || G->b0 != A->b0 || G->b1 != A->b1)
    return set_error(-4, "attn bwd: dO has the wrong shape or dtype");
  if (dV->dtype != dt || dV->rows != V->rows || dV->cols != V->cols || dV->b0 != A->b0 || dV->b1 != A->b1)
    return set_error(-4, "attn bwd: dV has the wrong shape or dtype");
  if (dB->dtype != dt || dB->rows != B->rows || dB->cols != B->cols || dB->b0 != A->b0 || dB->b1 != A->b1)
    return set_error(-4, "attn bwd: dB has the wrong shape or dtype");
  if (!(scale > 0.f)) return set_error(-4, "attn: scale must be positive");
  if (stat_ld < A->rows || stat_ld % 64) return set_error(-4, "attn bwd: stat_ld must be a multiple of 64 >= rows");
  if (reinterpret_cast<uintptr_t>(neg_lse) & 15 || reinterpret_cast<uintptr_t>(nds) & 15)
    return set_error(-2, "attn bwd: the row statistics must be 16-byte aligned");

  AttnBwdParams p;
  memset(&p, 0, sizeof p);
  p.M = (int)A->rows; p.Tn = (int)B->rows; p.d = (int)A->cols; p.dv = (int)V->cols;
  p.b0 = (int)A->b0; p.b1 = (int)A->b1;
  p.world = world; p.rank = rank;
  const int n_mb = (p.Tn + kAtBM - 1) / kAtBM;
  p.m_pairs = (n_mb + 1) / 2;
  p.n_steps = (p.M + kAbStep - 1) / kAbStep;
  p.out0 = dV->ptr; p.ld0 = dV->ld; p.o0_sb0 = dV->sb0; p.o0_sb1 = dV->sb1;
  p.out1 = dB->ptr; p.ld1 = dB->ld; p.o1_sb0 = dB->sb0; p.o1_sb1 = dB->sb1;
  p.neg_lse = neg_lse; p.nds = nds; p.stat_ld = stat_ld;
  p.mask_bits = mask_bits_t; p.mb_ld = mask_ld_words; p.mb_sb1 = mask_sb1_words;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  p.dbg = dbg | attn_env_dbg();
  if (int e = tile_map(&p.tmA, A, kAbStep, false)) return e;
  if (int e = tile_map(&p.tmG, G, kAbStep, false)) return e;
  if (int e = tile_map(&p.tmB, B, kAtBM, false)) return e;
  if (int e = tile_map(&p.tmV, V, kAtBM, false)) return e;
  p.o0_vec_ok = out_vec_ok(dV);
  p.o1_vec_ok = out_vec_ok(dB);
  if (world > 1 && !(p.o0_vec_ok && p.o1_vec_ok)) return set_error(-6, "attn bwd: world > 1 needs 16-byte aligned outputs");
  if (p.o0_vec_ok && !(dbg & 16) && tile_map(&p.tmOut0, dV, 32, false) == 0) p.o0_tma_ok = 1;
  if (p.o1_vec_ok && !(dbg & 16) && tile_map(&p.tmOut1, dB, 32, false) == 0) p.o1_tma_ok = 1;
  if (!p.o0_tma_ok) p.tmOut0 = p.tmA;
  if (!p.o1_tma_ok) p.tmOut1 = p.tmA;
  p.tmBg = p.tmB;
  p.tmVg = p.tmV;
  if (world > 1) {
    if (int e = attn_bwd_comm(p, B, V, comm, kAtBM)) return e;
    const ColsLayout L = cols_layout(B, V, world, p.chunk_rows);
    if ((size_t)p.ws_v_off != L.v_off) return set_error(-20, "attn bwd: workspace layout mismatch");
    p.ws_p0_off = (long long)L.p0_off;
    p.ws_p1_off = (long long)L.p1_off;
    p.rs_flag_off = (long long)L.gather_flags;
    if (p.phase_mask == 3) p.phase_mask = 7;  // production: push, main loop and reduction in one launch
  } else {
    p.phase_mask = 7; p.chunk_rows = 1 << 30; p.n_chunks = 1;
  }
  const int sms = ddp_num_sms();
  if (!(p.dbg & 64)) {  // second generation: one 128-column block per item, P^T / dS^T stay in tensor memory
    constexpr int kSmem2 = 2 * kAtTile + kAb2ColsOut + kAb2ColsStages * kAb2ColsStage + (2 * kAb2ColsStages + 20) * 8 + 16 + 1024;
    static_assert(kSmem2 <= 232448, "shared memory budget exceeded");
    const long long items2 = (long long)world * p.b1 * n_mb * p.b0;
    int grid2 = world > 1 ? sms : (int)(items2 < sms ? items2 : sms);
    if (grid2 < 1) grid2 = 1;
    static bool cfg2_bf[64] = {}, cfg2_h[64] = {};
    return dt == DDP_BF16 ? attn_bwd_launch(attn_bwd2_cols_kernel<DDP_BF16>, cfg2_bf, kSmem2, p, grid2, stream)
                          : attn_bwd_launch(attn_bwd2_cols_kernel<DDP_F16>, cfg2_h, kSmem2, p, grid2, stream);
  }
  const long long items = (long long)world * p.b1 * p.m_pairs * p.b0;
  int grid = world > 1 ? sms : (int)(items < sms ? items : sms);
  if (grid < 1) grid = 1;
  static bool cfg_bf[64] = {}, cfg_h[64] = {};
  return dt == DDP_BF16
             ? attn_bwd_launch(attn_bwd_cols_kernel<DDP_BF16, kAbColsStages>, cfg_bf, kAbColsSmem, p, grid, stream)
             : attn_bwd_launch(attn_bwd_cols_kernel<DDP_F16, kAbColsStages>, cfg_h, kAbColsSmem, p, grid, stream);
}
