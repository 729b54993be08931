// Host side of the slab-free attention forward: shape checks, TMA tensor maps, launch.
// C ABI declared in include/ddp_b200.h.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstring>

#include "../common/host_util.h"
#include "../gemm/tma_map.h"
#include "attn_fwd_sm100.cuh"

namespace ddp {

constexpr int kAtStages = 3;
constexpr int kAtSmem = 2 * kAtTile + 2 * kAtPBytes + kAtStages * kAtStageBytes + (2 * kAtStages + 14) * 8 + 16 + 4 * 128 * 8 + 1024;
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
  p.dbg = dbg;
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
  return dt == DDP_BF16 ? attn_launch<DDP_BF16>(p, grid, stream) : attn_launch<DDP_F16>(p, grid, stream);
}
