/* ffemb.h — C ABI of libffemb.so: the B200 (sm_100a) embedding hot path for DLRM-FlexFlow.
 *
 * This is the drop-in boundary for ONE path of the reference (/root/reference, cited as
 * file:line): the FlexFlow Embedding operator.  Legion task bodies keep their reference
 * signatures (Embedding::forward_task / backward_task, src/ops/embedding.cu:238-272,322-354)
 * and call these entry points where they used to call the static raw-pointer wrappers
 *     Embedding::forward_kernel   include/model.h:1169-1177, src/ops/embedding.cu:219-231
 *     Embedding::backward_kernel  include/model.h:1178-1186, src/ops/embedding.cu:308-320
 * The reference-side binding is shown in INTEGRATION.md; dlrm-flexflow_b200/csrc/flexflow/
 * embedding.cu is that binding written out.
 *
 * Conventions:
 *   - plain C: pointers, int64 sizes, an opaque CUDA stream; no C++/torch types;
 *   - every pointer is a DEVICE-accessible pointer borrowed from the caller (Legion owns all
 *     regions); nothing is allocated or freed per call; scratch comes from the caller
 *     (FFHandler::workSpace or an EmbeddingMeta-owned buffer) and is sized by the *_workspace
 *     queries;
 *   - all work is ordered on `stream`: enqueued on it, or (the bag-major small-table kernels of the
 *     backward) forked from it onto a library-owned side stream and joined back before the call's
 *     work ends, CUDA-graph-capture style; no host synchronisation, no device selection (the
 *     caller's current device is used), re-entrant across devices and threads;
 *   - returns FFEMB_OK or an error code and never throws / exits; ffemb_last_error() gives the
 *     text for the calling thread.  (The reference asserts / exit(1)s instead,
 *     include/cuda_helper.h:6-46; the drop-in task bodies turn non-zero into that.)
 *   - layouts are the flat row-major buffers the reference kernels see:
 *     idx int64 [B][L], table float [R][D], pooled output / output-grad float [B][D].
 *   - indices are NOT range-checked on the hot entry points (the reference GPU kernels do not
 *     either); ffemb_check_indices is the debug aid.
 */
#ifndef FFEMB_H
#define FFEMB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FFEMB_VERSION 100
#if defined(__GNUC__)
#define FFEMB_API __attribute__((visibility("default")))
#else
#define FFEMB_API
#endif

/* status codes */
#define FFEMB_OK 0
#define FFEMB_EINVAL 1     /* bad argument (NULL pointer, negative size, AGGR_MODE_NONE, ...) */
#define FFEMB_ECUDA 2      /* a CUDA call or launch failed */
#define FFEMB_EWORKSPACE 3 /* workspace missing, misaligned or smaller than *_workspace said */
#define FFEMB_ERANGE 4     /* ffemb_check_indices found an index outside [0, R) */

/* AggrMode values of include/ffconst.h:12-16, passed through as ints */
#define FFEMB_AGGR_NONE 20 /* accepted by FFModel::embedding but asserts in the kernels -> EINVAL */
#define FFEMB_AGGR_SUM 21
#define FFEMB_AGGR_AVG 22   /* REF_LITERAL: divide inside the bag loop, src/ops/embedding.cu:180-188 */
#define FFEMB_AGGR_MEAN 122 /* opt-in true mean (sum * (1/L)); not a reference enum value */

#define FFEMB_MAX_TABLES 64 /* tables per batched call */
#define FFEMB_MAX_PEERS 8   /* GPUs of one NVSwitch box */

/* ABI-compatible with cudaStream_t */
typedef struct CUstream_st *ffemb_stream_t;

/* One table of a batched call.  All tables of a call share B and D; L (bag width) and R
 * (rows) are per table.  `out` is the pooled output in forward and the output-grad in
 * backward.  `table` is the weight (forward: read; backward_sgd: read-write) or the dense
 * weight-grad (backward_dense*).                                                           */
typedef struct ffemb_table {
  const int64_t *idx; /* [B][L] */
  float *out;         /* [B][D] */
  float *table;       /* [R][D] */
  int64_t R;
  int32_t L;
  int32_t reserved;
} ffemb_table_t;

FFEMB_API const char *ffemb_last_error(void);
FFEMB_API int ffemb_version(void);

/* ---- forward: replaces Embedding::forward_kernel ------------------------------------------
 * out[b][d] = pool_j table[idx[b][j]][d]; every element of out is written (region is
 * WRITE_ONLY in the reference).  Accumulation is sequential in j per output element, so SUM
 * and AVG(REF_LITERAL) are bit-identical to embed_forward.  ffemb_forward has no row count (neither
 * has the wrapper it replaces): row numbers are carried as 32 bits there, and an index >= 2^32 (or
 * negative) TRAPS the kernel rather than gathering a wrong row; tables of more than 2^32 rows go
 * through ffemb_forward_batched with R set, which then takes the 64-bit path.                 */
FFEMB_API int ffemb_forward(const int64_t *idx, float *out, const float *table, int64_t L, int64_t D,
                  int64_t B, int aggr, ffemb_stream_t stream);

/* all tables of one GPU in one launch */
FFEMB_API int ffemb_forward_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                          int aggr, ffemb_stream_t stream);

/* ---- forward fused with the exchange (multi-GPU, table-wise sharding) ------------------------
 * The reference has no collective here: pooled outputs live in zero-copy host memory and Legion
 * copies each batch slice to the GPU that owns those samples when the data-parallel Concat
 * re-partitions them (MAP_TO_ZC_MEMORY, src/ops/embedding.cu:295-298; src/ops/concat.cu:88-97).
 * Here the gather-pool kernel stores every pooled row STRAIGHT into the Concat input of the rank
 * that owns the sample, over NVLink peer memory — the all-to-all happens tile by tile inside the
 * compute kernel, in the Concat layout of src/ops/concat.cu:211-251:
 *     sample b  ->  rank h = b / batch_per_peer (Legion's blocked split, src/runtime/model.cc:876-895)
 *     peer_x[h][(b - h*batch_per_peer) * row_stride + tables[t].reserved + d]   d in [0, D)
 * tables[t].reserved = column offset (in floats) of table t's block inside a sample row;
 * tables[t].out is ignored.  peer_x[] are device pointers valid on THIS device (peer-mapped:
 * cudaIpc / symmetric memory; peer_x[own rank] is plain local memory).  The caller orders the
 * consumers after all ranks' kernels (a cross-rank barrier on the stream).                    */
typedef struct ffemb_exchange {
  float *peer_x[FFEMB_MAX_PEERS];
  int32_t num_peers;
  int32_t reserved;
  int64_t batch_per_peer;
  int64_t row_stride;
} ffemb_exchange_t;
FFEMB_API int ffemb_forward_exchange(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                           int aggr, const ffemb_exchange_t *ex, ffemb_stream_t stream);

/* The transposed route for the output-grads: this rank holds dx [n_samples][row_stride] (the
 * gradient of its Concat input, samples [first_sample, first_sample + n_samples) of the batch)
 * and pushes column block tables[t].reserved of every sample into the output-grad buffer of the
 * table's owner: tables[t].out = that owner's [B][D] output-grad (peer-mapped pointer),
 * rows first_sample .. first_sample + n_samples - 1.  idx / table / R / L are ignored.          */
FFEMB_API int ffemb_exchange_backward(const ffemb_table_t *tables, int num_tables, int64_t D,
                            const float *dx, int64_t n_samples, int64_t first_sample,
                            int64_t row_stride, ffemb_stream_t stream);

/* The same push for PARTS of mixed width in one launch.  A part is a whole table or a column range
 * of one (a heavy table split by columns over two ranks: same indices on both, each pools / updates
 * its own columns — bit-identical results, no extra exchange bytes; sharding.plan_shards):
 *   parts[p].out      = the owner's output-grad buffer of the part, [B][w_p] floats (peer-mapped)
 *   parts[p].reserved = first column of the part inside a row of dx
 *   parts[p].L        = w_p, the part's width in floats (a multiple of 4)
 * idx / table are ignored.  parts[0].R, when in 1..16, is a hint: warps the push keeps resident per SM
 * (default 4) — few (2) when it runs UNDER compute kernels of a neighbouring table group, many (16) when
 * it runs alone; R of the other parts is ignored.  Rows travel as tiles of up to 8 KB: cp.async.bulk global -> shared,
 * then ONE contiguous cp.async.bulk shared -> peer global per tile (large posted NVLink writes issued
 * by a few dozen resident warps, so the push runs under the compute kernels of the neighbouring table
 * group); unaligned shapes fall back to 16-byte load/store per lane.                            */
FFEMB_API int ffemb_exchange_backward_parts(const ffemb_table_t *parts, int num_parts, const float *dx,
                                  int64_t n_samples, int64_t first_sample, int64_t row_stride,
                                  ffemb_stream_t stream);

/* ---- cross-rank ordering of the fused exchange ------------------------------
 * The stores of ffemb_forward_exchange / ffemb_exchange_backward* land in ANOTHER GPU's memory; the
 * consumer there must be ordered after them, and the producer's next step after the consumer.  In the
 * reference that ordering is Legion's (region dependences between the embedding tasks and Concat's,
 * src/ops/concat.cu:88-97); here it is a set of epoch flags in a small pad per rank that every rank
 * can address — raw device pointers valid on the calling device: cudaDeviceEnablePeerAccess inside one
 * Legion process (one process drives all GPUs of a node, src/runtime/model.cc:679-706 builds its
 * ncclComm_t the same way), cudaIpcOpenMemHandle across processes, or symmetric memory.  No NCCL call
 * and no host synchronisation is involved.
 *   pad: ffemb_sync_pad_bytes() bytes per rank, 8-byte aligned, ZERO-FILLED once before first use.
 *   ffemb_sync_signal(sync, slot): everything enqueued on `stream` before this call is visible to every
 *       peer that later passes ffemb_sync_wait on the same slot (fence.sys + st.release.sys of this
 *       rank's next epoch for the slot into every peer's pad, its own included).
 *   ffemb_sync_wait(sync, slot)  : work enqueued on `stream` after this call runs only once EVERY rank
 *       has signalled the slot as many times as this rank has now waited on it (a one-warp kernel that
 *       spins with ld.acquire.sys; it traps after 20 s rather than hanging the device).
 * Epochs are kept in the pad, so the calls carry no step number and can be captured in a CUDA graph.
 * Every rank must call signal and wait on a slot equally often.                                  */
#define FFEMB_SYNC_SLOTS 16
typedef struct ffemb_sync {
  uint64_t *peer_pad[FFEMB_MAX_PEERS]; /* pad of rank h as mapped on THIS device */
  int32_t num_peers;
  int32_t rank;
} ffemb_sync_t;
FFEMB_API size_t ffemb_sync_pad_bytes(void);
FFEMB_API int ffemb_sync_signal(const ffemb_sync_t *sync, int slot, ffemb_stream_t stream);
FFEMB_API int ffemb_sync_wait(const ffemb_sync_t *sync, int slot, ffemb_stream_t stream);

/* ---- row-wise sharded tables (a table too large for one GPU) --------------------------------
 * Not in the reference (Embedding asserts sample-dim-only partitioning, src/ops/embedding.cu:84-85).
 * Shard map = Legion's blocked restriction on the row dimension (src/runtime/model.cc:880-892):
 * rows_per_shard = ceil(R / G), owner(row) = row / rows_per_shard, local = row % rows_per_shard.
 * forward_rowwise pools THIS rank's samples (idx [B][L] holds global row ids) and reads every row
 * from the owner's shard over NVLink peer memory (peer_shards[g] = device pointer to shard g,
 * valid on this device); adds are sequential in j, so results are bit-identical to the unsharded
 * table.  Backward, SURVEY 8e (bucketise by owner, route indices + the needed output-grad rows to the
 * owners as remote stores, owners reduce their own lookups only): see ffemb_rowwise_route below.
 * ffemb_rowwise_localize is the round-1 form kept as the measured comparison (every owner walks the
 * WHOLE all-gathered batch; rows of other shards map to a scratch row `rows_per_shard` appended to the
 * shard, cleared here and never read by anyone).                                                  */
FFEMB_API int ffemb_forward_rowwise(const int64_t *idx, float *out, const float *const *peer_shards,
                          int num_peers, int64_t rows_per_shard, int64_t L, int64_t D, int64_t B,
                          int aggr, ffemb_stream_t stream);
FFEMB_API int ffemb_rowwise_localize(const int64_t *idx, int64_t *local_idx, int64_t n,
                           int64_t rows_per_shard, int shard, float *scratch_row, int64_t D,
                           ffemb_stream_t stream);

/* Row-wise backward, step 1: ROUTE.  Rank `rank` holds the indices of ITS samples (tables[t].idx =
 * [n_samples][L_t] global row ids) and dx [n_samples][row_stride], the gradient of its interaction
 * input (table t's block at column tables[t].reserved).  Every owner g keeps, for all tables back to
 * back, a slot array shaped like the GLOBAL index tensors and a routed output-grad buffer:
 *     peer_slots[g] : int64, table t at entry  S_t = sum_{s<t} G*ext*L_s,  shaped [G*ext][L_t]
 *     peer_rows[g]  : float, table t at row    G*ext*t,                    shaped [G*ext][D]
 * (ext = samples_per_peer = ceil(B / G), Legion's blocked batch split, src/runtime/model.cc:876-895;
 * pointers as mapped on THIS device).  The kernel fills the slots of its samples — global bags
 * [rank*ext, rank*ext + n_samples) — on EVERY owner:  row - g*rows_per_shard  where g owns the row,
 * -1 (empty slot) elsewhere; and stores the bag's output-grad row to every owner of at least one of
 * its lookups.  rows_per_shard = ceil(tables[t].R / G) with tables[t].R the rows of the WHOLE table;
 * tables[t].out / table are ignored.  Slot positions depend on (rank, sample, j) only: nothing is
 * counted or compacted, no counts are exchanged, and an owner meets equal rows in global (sample, j)
 * order — the order of the unsharded backward.  NVLink bytes per rank: 8*G per lookup (slots) + one
 * output-grad row per (bag, owner) pair, against the whole batch's gradient per rank for an
 * all-gather.  Order the owners' backward after all ranks' routes with ffemb_sync_signal / _wait.  */
typedef struct ffemb_route {
  int64_t *peer_slots[FFEMB_MAX_PEERS];
  float *peer_rows[FFEMB_MAX_PEERS];
  int32_t num_peers;
  int32_t rank;
  int64_t samples_per_peer;
} ffemb_route_t;
FFEMB_API int ffemb_rowwise_route(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t n_samples,
                        const float *dx, int64_t row_stride, const ffemb_route_t *route,
                        ffemb_stream_t stream);
/* Row-wise backward, step 2: the owner's fused backward + SGD over its slot arrays.  Exactly
 * ffemb_backward_sgd_batched (tables[t].idx = the table's slot array [B][L_t] with B = G*ext,
 * tables[t].out = its routed output-grad rows [B][D], tables[t].table = the shard, R = its rows)
 * except that an index < 0 is an EMPTY SLOT: it gets the key one past the last row, sorts to the
 * end, and the segmented sum neither loads an output-grad row nor touches a table row for it.  Every
 * table takes the sorted path.  ffemb_backward_workspace sizes the scratch as for any batched call. */
FFEMB_API int ffemb_backward_sgd_padded(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                              int aggr, float lr, void *workspace, size_t workspace_bytes,
                              ffemb_stream_t stream);

/* ---- backward, API-exact (unfused): replaces Embedding::backward_kernel -------------------
 * Accumulates into the dense, caller-zeroed weight-grad like embed_backward
 * (src/ops/embedding.cu:192-217).
 *   _atomic : red.global.add scatter (unordered, like the reference; the measured comparison)
 *   (plain) : deterministic sort + segmented reduce, needs workspace                         */
FFEMB_API int ffemb_backward_dense_atomic(const int64_t *idx, const float *outgrad, float *wgrad, int64_t L,
                                int64_t D, int64_t B, int aggr, ffemb_stream_t stream);
FFEMB_API int ffemb_backward_dense(const int64_t *idx, const float *outgrad, float *wgrad, int64_t L,
                         int64_t D, int64_t B, int64_t R, int aggr, void *workspace,
                         size_t workspace_bytes, ffemb_stream_t stream);

/* ---- backward fused with the sparse SGD row update ----------------------------------------
 * Replaces, for one table and momentum == weight_decay == 0:
 *   Op::zero_grad on the weight-grad   src/runtime/model.cc:466-490
 *   Embedding::backward_kernel          src/ops/embedding.cu:308-320
 *   SGDOptimizer::update -> sgd_update  src/runtime/optimizer_kernel.cu:23-41
 * table[r] = fma(-lr, sum_of_grads(r), table[r]) for every touched row r, once; untouched
 * rows are not read or written.  Deterministic: radix sort of (row, bag) pairs, then a
 * fixed-shape segmented sum.                                                                 */
FFEMB_API int ffemb_backward_sgd(const int64_t *idx, const float *outgrad, float *table, int64_t L,
                       int64_t D, int64_t B, int64_t R, int aggr, float lr, void *workspace,
                       size_t workspace_bytes, ffemb_stream_t stream);

/* The same step for callers that cannot allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scratch of the sorted backward lives in a file-scope device
 * array of FFEMB_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it
 * until the process exits) instead of a caller-supplied workspace; the call allocates nothing (no
 * cudaMalloc / cudaFree, no allocating library routine), runs the launch sequence of
 * ffemb_backward_sgd and gives bit-identical results.  Shapes whose
 * ffemb_backward_workspace(B*L, R, D) exceeds the array are refused with FFEMB_EWORKSPACE and a
 * message naming both sizes.  All calls on a device share the array, which keeps the last call's
 * contents: the span a call uses is cleared on `stream` first, and calls that overlap on the GPU
 * (different streams, or host threads, on one device) RACE — serialise them.                   */
#define FFEMB_STATIC_SCRATCH_BYTES (64u << 20)
FFEMB_API size_t ffemb_static_scratch_bytes(void);
FFEMB_API int ffemb_backward_sgd_static(const int64_t *idx, const float *outgrad, float *table, int64_t L,
                              int64_t D, int64_t B, int64_t R, int aggr, float lr, ffemb_stream_t stream);

FFEMB_API int ffemb_backward_sgd_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                               int aggr, float lr, void *workspace, size_t workspace_bytes,
                               ffemb_stream_t stream);
/* The same, split in two: the sort of the lookups depends on the INDICES only, so a caller can
 * enqueue it early — on a side stream, under the forward pass, the exchange or the dense MLPs
 * (in FlexFlow: a task launched next to the forward task of the same step; the indices region is
 * READ_ONLY in both) — and run only the segmented sum + row update once the output-grads exist.
 * _prepare ignores tables[t].out; _prepared must get the same tables / D / B / aggr / workspace
 * and be ordered after _prepare (stream order or an event).  Results are bit-identical to
 * ffemb_backward_sgd_batched.                                                                 */
FFEMB_API int ffemb_backward_prepare(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                           int aggr, void *workspace, size_t workspace_bytes, ffemb_stream_t stream);
FFEMB_API int ffemb_backward_sgd_prepared(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                                int aggr, float lr, void *workspace, size_t workspace_bytes,
                                ffemb_stream_t stream);
FFEMB_API int ffemb_backward_dense_batched(const ffemb_table_t *tables, int num_tables, int64_t D,
                                 int64_t B, int aggr, void *workspace, size_t workspace_bytes,
                                 ffemb_stream_t stream);

/* ---- backward fused with the reference's FULL sgd_update ------------------------
 * ffemb_backward_sgd* is exact only for momentum == weight_decay == 0; with either set, sgd_update
 * (src/runtime/optimizer_kernel.cu:23-41) changes EVERY row every step:
 *     gt = WGrad + weight_decay*W;  if (momentum > 0) { V = V*momentum + gt;
 *     gt = nesterov ? gt + momentum*V : V; }   W -= lr*gt
 * This entry point keeps those semantics without a dense gradient: rows the step touches get the
 * formula with their summed gradient inside the sorted segmented reduction (and are marked in a
 * bitmap), every other row gets it with WGrad = 0 in one streaming pass over W (and V) that also
 * clears the bitmap.  HBM traffic 8 (16 with momentum) bytes per table element instead of the
 * reference's 4 (zero-fill) + 12 (20), plus the atomics.  Same FFMA contraction as the reference's
 * kernel, so single-lookup rows and untouched rows are bit-identical to it.
 *   V[t]       : momentum buffer of table t, [R][D] floats (ignored when momentum == 0; the reference
 *                allocates it zero-initialised, src/runtime/optimizer.cc:50-87)
 *   touched[t] : ffemb_sgd_touched_words(R) 32-bit words, all zero on entry, all zero on return
 * Every table goes through the sorted path; no split (prepare) form.
 * Stream capture: unlike the other entry points this one (and the Adam form below) must NOT be
 * recorded into a CUDA graph — the optimizer arguments are uploaded from a host copy that lives
 * only for the duration of the call (and Adam's alpha_t changes every step anyway).             */
typedef struct ffemb_sgd {
  float lr, weight_decay, momentum;
  int32_t nesterov;
} ffemb_sgd_t;
FFEMB_API int64_t ffemb_sgd_touched_words(int64_t R);
FFEMB_API int ffemb_backward_sgdx_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                                int aggr, const ffemb_sgd_t *opt, float *const *V,
                                uint32_t *const *touched, void *workspace, size_t workspace_bytes,
                                ffemb_stream_t stream);

/* The same for AdamOptimizer (src/runtime/optimizer_kernel.cu:206-226; the reference's Adam is not
 * lazy either — with a zero gradient M and V still decay and W still moves):
 *     gt = WGrad + weight_decay*W;  M = beta1*M + (1-beta1)*gt;  V = beta2*V + (1-beta2)*gt*gt;
 *     W -= alpha_t * M / (sqrt(V) + epsilon)
 * alpha_t is the bias-corrected step the host advances once per iteration (AdamOptimizer::next,
 * src/runtime/optimizer.cc:248-254).  M[t], V[t]: [R][D] floats, zero-initialised by the caller
 * (AdamOptimizer::init).  Operation-for-operation the reference kernel's SASS (which products are
 * fused, IEEE sqrt and division), so single-lookup and untouched rows are bit-identical to it.
 * 24 bytes per table element per step (W, M, V read + written) — the semantics' minimum.          */
typedef struct ffemb_adam {
  float alpha_t, beta1, beta2, weight_decay, epsilon;
} ffemb_adam_t;
FFEMB_API int ffemb_backward_adam_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                                int aggr, const ffemb_adam_t *opt, float *const *M, float *const *V,
                                uint32_t *const *touched, void *workspace, size_t workspace_bytes,
                                ffemb_stream_t stream);

/* ---- straight into / out of the Concat operator's tensors (one GPU) ----------------
 * The reference's DLRM feeds every pooled output through Concat (interaction "cat",
 * examples/cpp/DLRM/dlrm.cc:49-61): forward copies each [B][D] input into its column block of the
 * [B][row_stride] output, one copy_with_stride launch per input (src/ops/concat.cu:211-251);
 * backward adds each column block of the output's gradient into the input's zero-filled gradient
 * region, one add_with_stride launch per input (src/ops/concat.cu:325-366; the zero-fill is
 * Op::zero_grad, src/runtime/model.cc:481-488).  These two entry points remove both passes for the
 * embedding inputs: forward_concat stores pooled row (b, t) at x[b*row_stride + tables[t].reserved
 * + d] (it is ffemb_forward_exchange with one peer); backward_sgd_concat reads table t's
 * output-grad rows in place at dx[b*row_stride + tables[t].reserved + d] — no per-table [B][D]
 * output / output-grad buffer exists.  tables[t].reserved = column offset in floats (a multiple of
 * 4), tables[t].out is ignored; x / dx 16-byte aligned, row_stride a multiple of 4.  prepared != 0:
 * ffemb_backward_prepare already ran on this workspace.  Results are bit-identical to
 * ffemb_forward_batched / ffemb_backward_sgd_batched on separate buffers.                         */
FFEMB_API int ffemb_forward_concat(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                         int aggr, float *x, int64_t row_stride, ffemb_stream_t stream);
FFEMB_API int ffemb_backward_sgd_concat(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                              int aggr, float lr, const float *dx, int64_t row_stride, int prepared,
                              void *workspace, size_t workspace_bytes, ffemb_stream_t stream);

/* Bytes of scratch the sorted backward needs for `num_lookups` = sum_t B*L_t lookups over
 * tables holding `total_rows` = sum_t R_t rows of D floats.  0 on invalid arguments.         */
FFEMB_API size_t ffemb_backward_workspace(int64_t num_lookups, int64_t total_rows, int64_t D);

/* The same plus room for the AVG / MEAN backward's pre-scaled output-grads.  embed_backward divides the
 * output-grad by in_dim for EVERY lookup (src/ops/embedding.cu:206-211: gradient = output_grad[i] /
 * in_dim inside the per-element thread, re-done for each of the bag's rows); the quotient only depends
 * on the bag, so with this much scratch the sorted backward divides ONCE per bag (IEEE division, the
 * same bits) into the workspace and the segmented sum adds the quotients — the per-lookup division was
 * 42 % of the segmented sum at bag width 64.  A workspace sized by ffemb_backward_workspace still
 * works: the division then stays in the walk.  B = batch, num_tables = tables of the call.      */
FFEMB_API size_t ffemb_backward_workspace_avg(int64_t num_lookups, int64_t total_rows, int64_t D, int64_t B,
                                    int num_tables);

/* Which backward algorithm a table of R rows, bag width L takes at batch B (informational; the
 * accounting in bench.py uses it): 0 = radix sort + segmented sum, 1 = "tiny" (R <= 16, register
 * accumulators), 2 = "small" (whole [R][D] gradient in shared memory).  All are deterministic.  */
FFEMB_API int ffemb_backward_path(int64_t R, int64_t L, int64_t D, int64_t B);

/* ---- sparse-input delivery: replaces DataLoader::load_sparse_input ------------------------
 * examples/cpp/DLRM/dlrm.cu:19-58, run by the reference once per table and step
 * (DataLoader::next_batch, examples/cpp/DLRM/dlrm.cc:482-520): a pinned staging allocation, a host
 * gather loop, a synchronous cudaMemcpy, cudaFreeHost and cudaDeviceSynchronize.  Here ONE launch
 * cuts the index batches of all tables out of the sample-major dataset
 *     full[sample][row_len]  (int64; the reference's full_sparse_input, `row_len` =
 *                             num_sparse_inputs * in_dim there)
 * `full` is any device-accessible pointer: device memory, or the reference's own placement —
 * zero-copy (mapped pinned) host memory, dlrm.cc:505-509 — which the kernel then reads over PCIe
 * itself, coalesced.  Batch sample i is dataset row sample_ids[i] (device-accessible int32, the
 * reference's SampleIdxs::idxs, examples/cpp/DLRM/dlrm.h:17-21), or first_sample + i when
 * sample_ids is NULL (the reference only ever fills consecutive ids, dlrm.cc:497-498).
 * Per table: tables[t].idx = DESTINATION [B][L] (written), tables[t].L = bag width (in_dim),
 * tables[t].reserved = first column of the table's block inside a sample row (the reference's
 * my_input_idx * in_dim); out / table / R are ignored.  Column blocks must not overlap.        */
FFEMB_API int ffemb_load_sparse_inputs(const ffemb_table_t *tables, int num_tables, const int64_t *full,
                             int64_t row_len, const int32_t *sample_ids, int64_t first_sample,
                             int64_t B, ffemb_stream_t stream);

/* ---- debug aid (not on the timed path) -----------------------------------------------------
 * Writes the number of indices outside [0, R) to *bad_count (device int64).                   */
FFEMB_API int ffemb_check_indices(const int64_t *idx, int64_t n, int64_t R, int64_t *bad_count,
                        ffemb_stream_t stream);

/* ---- profiling (the OpMeta::profiling analogue, src/ops/embedding.cu:266-271) -------------
 * When enabled, every entry point brackets its internal phases with CUDA events on the
 * caller's stream.  ffemb_profile_collect synchronises on those events, adds the elapsed
 * milliseconds and the number of kernel launches per phase into the two arrays
 * (FFEMB_NUM_PHASES entries each, either may be NULL) and resets.  Off by default; the hot
 * path then records nothing.  ffemb_launch_count is the running number of kernels this
 * library has launched in this process (its own kernels and the radix-sort passes).          */
#define FFEMB_PHASE_FORWARD 0
#define FFEMB_PHASE_KEYGEN 1
#define FFEMB_PHASE_SORT 2
#define FFEMB_PHASE_REDUCE1 3   /* level-1 segmented sum + row update */
#define FFEMB_PHASE_REDUCE_UP 4 /* upper levels */
#define FFEMB_PHASE_ATOMIC 5    /* atomic comparison backward */
#define FFEMB_PHASE_SMALL 6     /* bag-major backward of tables whose gradient fits on chip */
#define FFEMB_PHASE_EXCHANGE 7  /* output-grad push over peer memory */
#define FFEMB_PHASE_INPUT 8     /* sparse-input delivery (ffemb_load_sparse_inputs) */
#define FFEMB_PHASE_DENSE 9     /* general SGD: the pass over the rows a step did not touch */
#define FFEMB_PHASE_SYNC 10     /* cross-rank flag kernels (ffemb_sync_signal / ffemb_sync_wait) */
#define FFEMB_PHASE_PRESCALE 11 /* AVG / MEAN backward: output-grad / in_dim, once per bag, ahead of the segmented sum */
#define FFEMB_PHASE_HOT 12      /* skewed indices: rows with thousands of lookups summed bag-major (select, mark, sum, apply) */
#define FFEMB_NUM_PHASES 13
FFEMB_API int ffemb_profile_enable(int on);
FFEMB_API int ffemb_profile_collect(float *ms_per_phase, int64_t *launches_per_phase);
FFEMB_API int64_t ffemb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* FFEMB_H */
