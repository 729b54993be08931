/* C ABI of libddp_b200.so — the native layer of the B200 distributed-dot-product framework.
 *
 * Every entry point takes raw device pointers and a cudaStream_t (passed as void*), returns 0 on
 * success or a negative error code (ddp_last_error() gives the text), and never synchronises the
 * host with the device unless stated.  Python binds these through ctypes
 * (distributed_dot_product/_native.py); C/C++ callers include this header and link the .so.
 */
#ifndef DDP_B200_H_
#define DDP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDP_MAX_WORLD 16

/* element types */
enum { DDP_F16 = 0, DDP_BF16 = 1, DDP_F32 = 2 /* fp32 storage, TF32 tensor-core math */ };

/* distributed products (reference: distributed_dot_product/multiplication/functions.py) */
enum {
  DDP_OP_NT = 0, /* C[*,M,T]  = A[*,M,D]  * gather(B)[*,T,D]^T   functions.py:44-99   */
  DDP_OP_NN = 1, /* C[*,M,D]  = A[*,M,T]  * gather(B)[*,T,D]     functions.py:160-212 */
  DDP_OP_TN = 2  /* C[*,Tn,D] = sum_p A_p[*,:,rank-block]^T B_p  functions.py:102-148 */
};

/* A strided view [b1][b0][rows][cols]; strides in ELEMENTS; the column stride is 1. */
typedef struct ddp_tensor {
  void* ptr;
  int64_t rows, cols;
  int64_t ld;       /* row stride */
  int64_t b0, b1;   /* batch extents (b0 = inner, e.g. heads; b1 = outer) */
  int64_t sb0, sb1; /* batch strides */
  int32_t dtype;    /* DDP_F16 / DDP_BF16 / DDP_F32 */
  int32_t _pad;
} ddp_tensor;

/* Peer-memory context of ONE collective call.  world == 1 -> everything else is ignored.
 * ws[r] / flags[r] / done[r] are device pointers valid on THIS GPU that alias rank r's symmetric
 * workspace, flag pad and done counters (own rank included). */
typedef struct ddp_comm {
  int32_t world, rank;
  uint32_t seq;        /* call sequence number, identical on all ranks, strictly increasing */
  int32_t phase_mask;  /* 1 = produce (push / partial GEMM), 2 = consume; 3 = fused (production) */
  int32_t chunk_rows;  /* staging granularity in rows of the gathered operand (from `offset`) */
  int32_t _pad;
  void* ws[DDP_MAX_WORLD];
  uint32_t* flags[DDP_MAX_WORLD];
  uint32_t* done[DDP_MAX_WORLD]; /* done[r][q] = last seq rank q finished, stored in rank r's memory */
  uint32_t* exit_counter;        /* one zero-initialised uint32 in local memory */
  uint32_t* status;              /* sticky error word in local memory (0 = ok); a wait on a peer that
                                    exceeds 20 s stores which wait (1..5) here instead of hanging; may be NULL */
  void* mc_ws;                   /* NVSwitch multicast alias of the workspace (ddp_mc_bind): a store to mc_ws + off
                                    lands at ws[r] + off of EVERY rank; NULL = push to each peer separately */
  const void* gather_ws;         /* attention backward only: this rank's workspace (of EITHER parity) that still holds
                                    the complete all-gather of exactly these B / V operands (the forward's); the
                                    kernel then neither pushes nor waits for them.  NULL = gather in this call */
} ddp_comm;

const char* ddp_last_error(void);
int ddp_version(void);
int ddp_num_sms(void);

/* Bytes of symmetric workspace / number of uint32 flags one ddp_gemm call needs (0 if world==1). */
int ddp_gemm_workspace(int op, const ddp_tensor* A, const ddp_tensor* B, int world, int chunk_rows,
                       size_t* ws_bytes, size_t* n_flags);

/* Fused collective GEMM on tcgen05/TMEM tiles fed by TMA.
 *   NT: A=[*,M,D]   B=[*,Tn,D]  C=[*,M,world*Tn]   all-gather(B) -> GEMM
 *   NN: A=[*,M,T]   B=[*,Tn,D]  C=[*,M,D]          all-gather(B) -> GEMM (rank sum in TMEM)
 *   TN: A=[*,R,T]   B=[*,R,D]   C=[*,Tn,D]         GEMM -> reduce-scatter into the owner's buffer
 * C = alpha * product (+ bias[col], NT/NN with world==1 only).  bias may be NULL; it has C's dtype.
 * dbg: 0 in production (bit 0 swaps LBO/SBO of MN-major descriptors; test harness only). */
int ddp_gemm(int op, const ddp_tensor* A, const ddp_tensor* B, const ddp_tensor* C, float alpha,
             const void* bias, const ddp_comm* comm, int dbg, void* stream);

/* ddp_gemm with a fused epilogue.  DDP_EPI_SOFTMAX_BWD (nt only) turns the dP = dO . gather(V)^T product of
 * the attention backward into dS directly:  C = scale * P * (alpha * A.B^T - delta[row]),  with P the saved
 * probabilities (same shape, dtype and batch layout as C; P tiles are TMA-prefetched into the epilogue) and
 * delta[b1][b0][rows] = rowsum(dO * O) in fp32 - the softmax backward (autograd of module.py:65-67) without a
 * separate pass over the slab. */
enum { DDP_EPI_NONE = 0, DDP_EPI_SOFTMAX_BWD = 1, DDP_EPI_SOFTMAX_STATS = 2, DDP_EPI_SOFTMAX_APPLY = 3 };
/* DDP_EPI_SOFTMAX_STATS / _APPLY (nt only) fuse module.py:65-67 (scale, masked_fill, softmax) into the
 * K.gather(Q)^T product so that the score slab never exists: the STATS launch stores, per row and column tile,
 * the online-softmax (max, sum) of scale*log2(e)*scores into stats[b1][b0][rows][ddp_gemm_stats_tiles()] and
 * writes nothing else; ddp_softmax_combine folds them into rowstat[b1][b0][rows] = (row max, 1/row sum); the
 * APPLY launch recomputes the tiles and writes C = exp2(t - max) / sum directly.  mask_bits is the mask packed
 * by ddp_pack_mask (bit set = masked out; NULL = no mask). */
typedef struct ddp_epilogue {
  int32_t mode;
  float scale;
  const ddp_tensor* P;       /* SOFTMAX_BWD */
  const float* delta;        /* SOFTMAX_BWD */
  const uint32_t* mask_bits; /* SOFTMAX_STATS / APPLY */
  int64_t mask_ld_words, mask_sb1_words;
  void* stats;               /* SOFTMAX_STATS out: float2 */
  const void* rowstat;       /* SOFTMAX_APPLY in: float2 */
} ddp_epilogue;
/* Pipeline watchdog: id of the last in-kernel mbarrier wait that gave up after 4 s (0 = none; 1 smem slot free,
 * 2 TMEM buffer free, 3 operands landed, 4 accumulator ready, 5 P chunk landed); synchronises, clears the id. */
int ddp_debug_timeout_site(void);
int ddp_debug_timeout_site_rowops(void); /* same, for the softmax kernels (site 6 = row landed) */
int ddp_gemm_stats_tiles(const ddp_tensor* A, const ddp_tensor* B, int world);
/* bits[b1][rows][ld_words] <- mask u8 [b1][rows][cols] (nonzero = masked); bits past `cols` are set. */
int ddp_pack_mask(const uint8_t* mask, int64_t b1, int64_t rows, int64_t cols, int64_t mask_ld, int64_t mask_sb1,
                  uint32_t* bits, int64_t ld_words, void* stream);
/* rowstat[r] = (max_t m, 1 / sum_t l * exp2(m_t - max)) over the nt column tiles of row r; n_rows = b1*b0*rows. */
int ddp_softmax_combine(const void* stats, int64_t n_rows, int32_t nt, void* rowstat, void* stream);
int ddp_gemm_ex(int op, const ddp_tensor* A, const ddp_tensor* B, const ddp_tensor* C, float alpha,
                const void* bias, const ddp_comm* comm, const ddp_epilogue* epi, int dbg, void* stream);

/* Slab-free attention forward (module.py:61-69 in ONE kernel, all-gather of B and V fused in):
 *     O = softmax(scale * A . gather(B)^T, mask) . gather(V)
 * A = [b1][b0][M][d] (this rank's rows; the reference's keys), B = [b1][b0][Tn][d] and V = [b1][b0][Tn][dv] (this
 * rank's share of the column-side operands; the reference's queries / values), O = [b1][b0][M][dv].  bf16 / fp16,
 * d and dv multiples of 8 and <= 64, scale > 0.  Scores stay in TMEM, probabilities go through shared memory
 * back into the tensor core (online softmax); the [M x T] slab is never written.  lse (optional, fp32
 * [b1][b0][M]) receives log2(sum_j exp(scale * s_ij)) = row max + log2(row sum) in the log2 domain.
 * mask_bits: packed by ddp_attn_pack_mask(transposed = 0) (bit set = masked out; NULL = no mask); a fully masked
 * row is NaN as in the reference.  comm as for ddp_gemm (workspace / flag sizes: ddp_attn_fwd_workspace).
 * Errors -6: shape or dtype outside what this kernel covers (callers fall back to the three-kernel path). */
int ddp_attn_fwd_workspace(const ddp_tensor* B, const ddp_tensor* V, int world, int chunk_rows, size_t* ws_bytes,
                           size_t* n_flags);
int ddp_attn_fwd(const ddp_tensor* A, const ddp_tensor* B, const ddp_tensor* V, const ddp_tensor* O, float* lse,
                 const uint32_t* mask_bits, int64_t mask_ld_words, int64_t mask_sb1_words, float scale,
                 const ddp_comm* comm, int dbg, void* stream);
/* Slab-free attention backward (autograd of module.py:61-69), two deterministic kernels that recompute the
 * probabilities from the forward's log-sum-exp instead of reading a saved slab:
 *   rows:  dA[b1][b0][M][d]  = sum_j dS_ij gather(B)_j                  (all-gather of B, V fused in)
 *   cols:  dV[b1][b0][Tn][dv] = sum_ranks sum_i P_ij dO_i,  dB likewise with dS and A
 *          (all-gather of B, V AND the reduce-scatter of the per-rank partial sums fused in)
 * with P_ij = exp2(scale*log2e*A_i.B_j + neg_lse_i), dS_ij = P_ij*(scale*dO_i.V_j + nds_i); G = dO.
 * neg_lse = -L and nds = -scale*rowsum(dO*O) are fp32 [b1][b0][stat_ld] arrays, stat_ld a multiple of 64 >= M
 * (rows past M are read but ignored).  mask_bits: as for ddp_attn_fwd; mask_bits_t: ddp_attn_pack_mask(transposed = 1).  Same shape limits as ddp_attn_fwd (error -6).
 * ddp_attn_bwd_rows uses the workspace of ddp_attn_fwd_workspace; ddp_attn_bwd_cols its own. */
int ddp_attn_bwd_rows(const ddp_tensor* A, const ddp_tensor* G, const ddp_tensor* B, const ddp_tensor* V,
                      const ddp_tensor* dA, const float* neg_lse, const float* nds, int64_t stat_ld,
                      const uint32_t* mask_bits, int64_t mask_ld_words, int64_t mask_sb1_words, float scale,
                      const ddp_comm* comm, int dbg, void* stream);
int ddp_attn_bwd_cols_workspace(const ddp_tensor* B, const ddp_tensor* V, int world, int chunk_rows, size_t* ws_bytes,
                                size_t* n_flags);
int ddp_attn_bwd_cols(const ddp_tensor* A, const ddp_tensor* G, const ddp_tensor* B, const ddp_tensor* V,
                      const ddp_tensor* dV, const ddp_tensor* dB, const float* neg_lse, const float* nds,
                      int64_t stat_ld, const uint32_t* mask_bits_t, int64_t mask_ld_words, int64_t mask_sb1_words,
                      float scale, const ddp_comm* comm, int dbg, void* stream);
/* Bit-packed masks of the attention kernels (bit set = masked out; elements outside the mask count as masked).
 * The kernels' threads own one row (ddp_attn_fwd, ddp_attn_bwd_rows) or one column (ddp_attn_bwd_cols) each and
 * read one 32-bit word per 32 elements of the other axis, so the words of 32 neighbouring rows (columns) are
 * interleaved and a warp's loads coalesce:
 *   transposed = 0: bits[b1][ceil(rows/32)][ld_words = ceil(cols/32)][32], bit k of word w of row r <- mask[r][32w+k]
 *   transposed = 1: bits[b1][ceil(cols/32)][ld_words = ceil(rows/32)][32], bit k of word w of col c <- mask[32w+k][c]
 * mask is uint8 [b1][rows][cols] (row stride mask_ld, batch stride mask_sb1).  ddp_attn_mask_words gives the
 * size of `bits` in 32-bit words; the kernels take ld_words and the per-b1 stride (mask_words / b1). */
size_t ddp_attn_mask_words(int64_t b1, int64_t rows, int64_t cols, int transposed);
int ddp_attn_pack_mask(const uint8_t* mask, int64_t b1, int64_t rows, int64_t cols, int64_t mask_ld, int64_t mask_sb1,
                       uint32_t* bits, int transposed, void* stream);
int ddp_debug_timeout_site_attn(void); /* pipeline watchdog of the attention kernels (sites 11-49) */
/* Row statistics for ddp_attn_bwd_rows / ddp_attn_bwd_cols in one launch (replaces the autograd of the reference's
 * softmax, module.py:65-67): neg_lse[b][h][r] = -lse, nds[b][h][r] = -scale * rowsum(dO * O), fp32 [b1*b0][stat_ld],
 * zero for r >= rows; non-finite values (fully masked rows) become 0.  dO / O: [b1][b0][rows][dv] views. */
int ddp_attn_bwd_stats(const ddp_tensor* dO, const ddp_tensor* O, const float* lse, float* neg_lse, float* nds,
                       int64_t stat_ld, float scale, void* stream);
/* Pipeline trace of ddp_attn_fwd (a profiling aid; the reference has only the `measure` printout,
 * multiplication/functions.py:24-41): while `device_buf` is non-null, CTA 0 of every forward launch stores
 * clock64() stamps of the first `steps` column steps of its first work item as [actors][steps][events] int64
 * (actors 0-2 and 5-7: softmax warps, 3-4: the two MMA issuers; events: see attn_fwd_sm100.cuh).  Pass NULL to
 * switch it off again.  Not thread safe; tools/trace_attn.py and `attn_harness trace` print the table. */
void ddp_attn_debug_trace(long long* device_buf);
int ddp_attn_debug_trace_dims(int* actors, int* steps, int* events);

/* Row softmax over the last dim of x[b1][b0][rows][cols] with fused scale and boolean mask:
 *   y = softmax(scale * x  masked_fill(mask, -inf))     (reference: module.py:65-67)
 * mask is uint8 [b1][rows][cols] (broadcast over b0 = heads), nonzero = masked out; may be NULL.
 * y may alias x (in place). */
int ddp_softmax_fwd(const ddp_tensor* x, const ddp_tensor* y, const uint8_t* mask, int64_t mask_ld,
                    int64_t mask_sb1, float scale, void* stream);

/* dx = scale * y * (dy - delta), delta = rowsum(dy*y) (computed in-kernel when delta == NULL, else read
 * from the fp32 array delta[b1][b0][rows]);  dx may alias dy.  (autograd of module.py:65-67) */
int ddp_softmax_bwd(const ddp_tensor* y, const ddp_tensor* dy, const ddp_tensor* dx, const float* delta,
                    float scale, void* stream);

/* out[r] = sum_c a[r][c] * b[r][c] in fp32; a, b are [rows][cols] views (row strides ld). */
int ddp_rowdot(const ddp_tensor* a, const ddp_tensor* b, float* out, void* stream);

/* out[c] = sum_r x[r][c]  (bias gradient; autograd of module.py:43-45,75 with add_bias=True).  x is
 * [rows][cols] with row stride ld; out has x's dtype.  `scratch` is caller-provided device memory of
 * `scratch_bytes` bytes (ddp_colsum_scratch_bytes gives the useful size) holding fp32 partial sums of the
 * row-split first stage; it may be NULL, then one CTA per 32 columns walks all the rows. */
size_t ddp_colsum_scratch_bytes(int64_t rows, int64_t cols);
int ddp_colsum(const ddp_tensor* x, void* out, void* scratch, size_t scratch_bytes, void* stream);
/* out[i] = sum over s < splits of part[s * n + i], fp32 accumulate: the split-K reduction of the weight gradients
 * (the reference lets cuBLAS/autograd do it inside nn.Linear, module.py:43-45).  n * sizeof(dtype) % 16 == 0. */
int ddp_sum_partials(const void* part, void* out, int dtype, int splits, int64_t n, void* stream);

/* Column sum for callers that cannot allocate (no counterpart in the reference).
 * Same result as ddp_colsum, but the fp32 partial sums go to a 4 MiB array built into the library instead of
 * caller scratch, and the call performs no cudaMalloc / cudaFree / other allocation.  Meant for code that has
 * no allocator to turn to: a server whose memory pool already owns the whole device, a latency-bound loop
 * that cannot afford the device-wide stall of cudaMalloc/cudaFree, a plug-in running under its host's
 * allocator lock.  Limits and rules:
 *   - x->cols <= ddp_colsum_noalloc_max_cols() (16384); wider inputs are refused with error -5.
 *   - Every GPU that loads the library holds the 4 MiB until the process exits.
 *   - The first call on each device looks up the array's device address; make that first call outside any
 *     stream capture.
 *   - All calls on a device share the array.  Nothing needs clearing between calls (every slot read is
 *     written first), but calls that can overlap on the GPU race: keep them on one stream or order them. */
size_t ddp_colsum_noalloc_max_cols(void);
int ddp_colsum_noalloc(const ddp_tensor* x, void* out, void* stream);

/* ---- symmetric heap (runtime) ------------------------------------------------------------ */
/* Allocates `bytes` of device memory suitable for peer mapping and returns a 64-byte IPC handle. */
/* Symmetric heap on the virtual-memory-management driver API with an NVSwitch multicast alias (replaces the
 * reference's Horovod / MPI transport, multiplication/functions.py:95,143,206; csrc/runtime/vmm_heap.cu):
 *   ddp_vmm_query      what the device offers: POSIX descriptor export, multicast objects, allocation granularity
 *   ddp_vmm_alloc      cuMemCreate + map + zero; *fd_out = descriptor to send to the peers (caller closes it)
 *   ddp_vmm_open       map the allocation behind a descriptor received from a peer
 *   ddp_vmm_release    unmap + drop the handle (own heap, peer mapping or multicast mapping)
 *   ddp_mc_create / ddp_mc_import / ddp_mc_add_device / ddp_mc_bind
 *                      one multicast object over every rank's heap; *mc_ptr + off addresses byte off of ALL heaps
 * Every failure names the driver call and its CUresult in ddp_last_error(). */
int ddp_vmm_query(int* fd_handles, int* multicast, size_t* granularity);
int ddp_vmm_alloc(size_t bytes, void** ptr, int* fd_out);
int ddp_vmm_open(int fd, size_t bytes, void** ptr);
int ddp_vmm_release(void* ptr);
int ddp_mc_create(size_t bytes, int n_devices, unsigned long long* object, int* fd_out);
int ddp_mc_import(int fd, unsigned long long* object);
int ddp_mc_add_device(unsigned long long object);
int ddp_mc_bind(unsigned long long object, void* heap_ptr, size_t bytes, void** mc_ptr);
/* cudaMalloc + CUDA IPC variant of the heap (fall-back when descriptors cannot be exported). */
int ddp_heap_alloc(size_t bytes, void** ptr, void* ipc_handle_out64);
int ddp_heap_open(const void* ipc_handle64, void** ptr);
int ddp_heap_close(void* ptr);
int ddp_heap_free(void* ptr);
int ddp_memset_async(void* ptr, int value, size_t bytes, void* stream);
/* Synchronous read of one device word (the sticky status word of ddp_comm). */
int ddp_read_u32(const void* dev_ptr, uint32_t* out);
/* Stream-ordered copy of one device word into pinned host memory (no device synchronisation): the runtime polls
 * the status word with it every few collective calls. */
int ddp_read_u32_async(const void* dev_ptr, uint32_t* pinned_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DDP_B200_H_ */
