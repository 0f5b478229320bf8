/* oracle/ffemb_oracle.c — TEST INFRASTRUCTURE (the parity checker), not product code.
 *
 * A plain-C, single-threaded CPU restatement of the arithmetic on the DLRM-FlexFlow
 * Embedding hot path, for the cases where the reference has no CPU code of its own
 * (multi-hot bags, AGGR_MODE_AVG, the dense zero-fill + sgd_update, the shard maps).
 * Every function cites the reference file:line (relative to /root/reference) it follows.
 *
 * Pinning: the reference holds NO test, golden vector or fixture for the embedding op
 * (SURVEY.md §4, §8c).  This restatement is therefore pinned against the reference's own
 * code compiled unmodified here (oracle/_ref/libffref_cpu.so, built by oracle/Makefile from
 * /root/reference/src/ops/embedding.cc) and, on the GPU box, against
 * oracle/_ref/libffref_gpu.so (/root/reference/src/ops/embedding.cu et al. for sm_100a);
 * outputs of those runs are committed under tests/golden/ with the generating script.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (no
 * contraction: every fused multiply-add below is spelled out with fmaf on purpose).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* include/ffconst.h:12-16 */
#define AGGR_MODE_NONE 20
#define AGGR_MODE_SUM 21
#define AGGR_MODE_AVG 22
/* not a reference enum value: the true-mean forward (sum, then one multiply by 1/L), which is
 * what the reference's CPU lookup computes with normalize_by_lengths (src/ops/embedding.cc:
 * 302-313) and what its own backward differentiates (src/ops/embedding.cu:206-211). */
#define FFEMB_AGGR_MEAN 122

#define ORACLE_OK 0
#define ORACLE_EINVAL 1
#define ORACLE_ERANGE 2

/* ---- embed_forward: src/ops/embedding.cu:166-190 ------------------------------------------
 * for each output element i=(b,d): out[i]=0; for j<L: out[i]+=W[idx[b*L+j]*D+d];
 * AVG divides INSIDE the j loop.  R>0 enables a bounds check (the GPU
 * reference has none; the CPU one asserts, src/ops/embedding.cc:71-73).                       */
int oracle_embed_forward(const int64_t *idx, float *out, const float *W, int64_t L, int64_t D,
                         int64_t B, int aggr, int64_t R) {
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG && aggr != FFEMB_AGGR_MEAN)
    return ORACLE_EINVAL; /* AGGR_MODE_NONE trips assert(aggr==AVG), embedding.cu:185 */
  for (int64_t b = 0; b < B; b++) {
    for (int64_t d = 0; d < D; d++) {
      volatile float acc = 0.0f; /* volatile: keep fp32 rounding at every step */
      for (int64_t j = 0; j < L; j++) {
        int64_t w = idx[b * L + j];
        if (R > 0 && (w < 0 || w >= R)) return ORACLE_ERANGE;
        acc = acc + W[w * D + d];
        if (aggr == AGGR_MODE_AVG) acc = acc / (float)L; /* embedding.cu:186 */
      }
      if (aggr == FFEMB_AGGR_MEAN && L > 0) {
        float len_inv = 1.0f / (float)L; /* embedding.cc:302-313 */
        acc = acc * len_inv;
      }
      out[b * D + d] = acc;
    }
  }
  return ORACLE_OK;
}

/* ---- embed_backward: src/ops/embedding.cu:192-217 -----------------------------------------
 * g = outgrad[i] (SUM) or outgrad[i]/L (AVG); for j<L: atomicAdd(&Wgrad[idx*D+d], g).
 * The atomics are unordered on the GPU; this walks i ascending, j ascending — one legal
 * order.  Wgrad is accumulated into (embedding.cu:214), not overwritten.                      */
int oracle_embed_backward(const int64_t *idx, const float *outgrad, float *Wgrad, int64_t L,
                          int64_t D, int64_t B, int aggr, int64_t R) {
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG && aggr != FFEMB_AGGR_MEAN)
    return ORACLE_EINVAL;
  for (int64_t b = 0; b < B; b++)
    for (int64_t j = 0; j < L; j++) {
      int64_t w = idx[b * L + j];
      if (R > 0 && (w < 0 || w >= R)) return ORACLE_ERANGE;
      for (int64_t d = 0; d < D; d++) {
        float g = outgrad[b * D + d];
        if (aggr != AGGR_MODE_SUM) g = g / (float)L; /* embedding.cu:210 */
        volatile float s = Wgrad[w * D + d] + g;
        Wgrad[w * D + d] = s;
      }
    }
  return ORACLE_OK;
}

/* Same sums in fp64 (tolerance analysis): g is still formed in fp32 like the
 * reference, only the accumulation across lookups is exact-ish.                               */
int oracle_embed_backward_f64(const int64_t *idx, const float *outgrad, double *Wgrad,
                              int64_t L, int64_t D, int64_t B, int aggr, int64_t R) {
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG && aggr != FFEMB_AGGR_MEAN)
    return ORACLE_EINVAL;
  for (int64_t b = 0; b < B; b++)
    for (int64_t j = 0; j < L; j++) {
      int64_t w = idx[b * L + j];
      if (R > 0 && (w < 0 || w >= R)) return ORACLE_ERANGE;
      for (int64_t d = 0; d < D; d++) {
        float g = outgrad[b * D + d];
        if (aggr != AGGR_MODE_SUM) g = g / (float)L;
        Wgrad[w * D + d] += (double)g;
      }
    }
  return ORACLE_OK;
}

/* ---- zero-fill: src/runtime/cuda_helper.cu:51-59 via initializer_kernel.cu:209-241 -------- */
void oracle_zero(float *p, int64_t n) {
  for (int64_t i = 0; i < n; i++) p[i] = 0.0f;
}

/* ---- sgd_update: src/runtime/optimizer_kernel.cu:23-41 ------------------------------------
 *   gt = WGrad[i] + weight_decay*W[i]; (momentum…); W[i] -= lr*gt
 * nvcc's default -fmad=true contracts both statements into FFMA (checked in the SASS of the
 * reference object: oracle/Makefile target `sass-check`); use_fma=1 reproduces that rounding,
 * use_fma=0 rounds the product first (what a -fmad=false build or a CPU port would do).      */
void oracle_sgd_update(int64_t count, float lr, float weight_decay, float momentum, int nesterov,
                       const float *WGrad, float *V, float *W, int use_fma) {
  for (int64_t i = 0; i < count; i++) {
    float gt;
    if (use_fma) gt = fmaf(weight_decay, W[i], WGrad[i]);
    else { volatile float p = weight_decay * W[i]; gt = WGrad[i] + p; }
    if (momentum > 0.0f) {
      if (use_fma) V[i] = fmaf(V[i], momentum, gt);
      else { volatile float p = V[i] * momentum; V[i] = p + gt; }
      if (nesterov) {
        if (use_fma) gt = fmaf(momentum, V[i], gt);
        else { volatile float p = momentum * V[i]; gt = gt + p; }
      } else gt = V[i];
    }
    if (use_fma) W[i] = fmaf(-lr, gt, W[i]);
    else { volatile float p = lr * gt; W[i] = W[i] - p; }
  }
}

/* ---- adam_update: src/runtime/optimizer_kernel.cu:206-226 -----------------------------------
 *   gt = WGrad[i] + weight_decay*W[i]; mt = beta1*M[i] + (1-beta1)*gt;
 *   vt = beta2*V[i] + (1-beta2)*gt*gt; M[i] = mt; V[i] = vt; W[i] -= alpha_t*mt/(sqrt(vt)+epsilon)
 * Rounding follows the SASS nvcc emits for the reference object (cuobjdump of
 * oracle/_ref/optimizer_kernel_cu.o): gt = FFMA(W, wd, WGrad); t1 = gt*(1-beta1); t2 = gt*(1-beta2);
 * gg = gt*t2; mt = FFMA(M, beta1, t1); vt = FFMA(V, beta2, gg); num = mt*alpha_t;
 * W = W - num / (sqrtf(vt) + epsilon) with IEEE sqrt and division, (1-beta) formed in fp32.      */
void oracle_adam_update(int64_t count, float alpha_t, float beta1, float beta2, float weight_decay,
                        float epsilon, const float *WGrad, float *M, float *V, float *W) {
  volatile float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
  for (int64_t i = 0; i < count; i++) {
    float gt = fmaf(W[i], weight_decay, WGrad[i]);
    volatile float t1 = gt * omb1;
    volatile float t2 = gt * omb2;
    volatile float gg = gt * t2;
    float mt = fmaf(M[i], beta1, t1);
    float vt = fmaf(V[i], beta2, gg);
    M[i] = mt;
    V[i] = vt;
    volatile float num = mt * alpha_t;
    volatile float den = sqrtf(vt) + epsilon;
    volatile float q = num / den;
    W[i] = W[i] - q;
  }
}

/* One reference training step of one table, as FFModel runs it:
 * zero_gradients (model.cc:1993-1996) -> Embedding::backward -> SGDOptimizer::update.
 * Wgrad is caller-provided scratch of R*D floats.                                             */
int oracle_reference_step(const int64_t *idx, const float *outgrad, float *W, float *Wgrad,
                          int64_t L, int64_t D, int64_t B, int64_t R, int aggr, float lr,
                          int use_fma) {
  oracle_zero(Wgrad, R * D);
  int rc = oracle_embed_backward(idx, outgrad, Wgrad, L, D, B, aggr, R);
  if (rc) return rc;
  oracle_sgd_update(R * D, lr, 0.0f, 0.0f, 0, Wgrad, NULL, W, use_fma);
  return ORACLE_OK;
}

/* ======================= placement / routing (integer work: bit-exact) ==================== */

/* Table-wise placement: table i whole on GPU i % G.
 * examples/cpp/DLRM/strategies/dlrm_strategy.cc:252-256 (add_embed_config(..., i % ngpus)),
 * grid [1,1] per table (:110-114).                                                           */
int oracle_table_owner(int64_t table, int64_t G) { return (int)(table % G); }

/* Data-parallel batch split: Legion create_partition_by_restriction with blocked extents
 * ext = (B + G - 1) / G  (src/runtime/model.cc:876-895).  Rank g owns [lo, hi).               */
void oracle_batch_block(int64_t B, int64_t G, int64_t g, int64_t *lo, int64_t *hi) {
  int64_t ext = (B + G - 1) / G;
  int64_t l = g * ext, h = l + ext;
  if (l > B) l = B;
  if (h > B) h = B;
  *lo = l;
  *hi = h;
}

/* Row-wise sharding does not exist in the reference (src/ops/embedding.cu:84-85 asserts
 * sample-dim-only partitioning).  Definition adopted: the same blocked extents
 * applied to the row dimension: rows_per_shard = ceil(R/G), owner = row / rows_per_shard,
 * local row = row % rows_per_shard.                                                          */
int64_t oracle_rows_per_shard(int64_t R, int64_t G) { return (R + G - 1) / G; }
int oracle_row_owner(int64_t row, int64_t R, int64_t G) {
  return (int)(row / oracle_rows_per_shard(R, G));
}
int64_t oracle_row_local(int64_t row, int64_t R, int64_t G) {
  return row % oracle_rows_per_shard(R, G);
}

/* Stable bucketing of an index list by row owner (row-wise routing).  For each owner o the
 * positions p (ascending) with owner(idx[p])==o are written to perm[offsets[o] ..
 * offsets[o+1]) and their local rows to local[]; offsets has G+1 entries.                    */
int oracle_route_rows(const int64_t *idx, int64_t n, int64_t R, int64_t G, int64_t *offsets,
                      int64_t *perm, int64_t *local) {
  int64_t rps = oracle_rows_per_shard(R, G);
  for (int64_t o = 0; o <= G; o++) offsets[o] = 0;
  for (int64_t p = 0; p < n; p++) {
    if (idx[p] < 0 || idx[p] >= R) return ORACLE_ERANGE;
    offsets[idx[p] / rps + 1]++;
  }
  for (int64_t o = 0; o < G; o++) offsets[o + 1] += offsets[o];
  int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (size_t)(G > 0 ? G : 1));
  if (!cur) return ORACLE_EINVAL;
  for (int64_t o = 0; o < G; o++) cur[o] = offsets[o];
  for (int64_t p = 0; p < n; p++) {
    int64_t o = idx[p] / rps;
    perm[cur[o]] = p;
    local[cur[o]] = idx[p] % rps;
    cur[o]++;
  }
  free(cur);
  return ORACLE_OK;
}

/* ---- Concat forward along axis 1 of [batch][features]: src/ops/concat.cu:211-251 with
 * copy_with_stride (src/runtime/cuda_helper.cu:127-144): output row b =
 * [in_0(b), in_1(b), ...], input i contributing widths[i] floats.  This is the destination
 * layout of the pooled-embedding exchange.                                      */
void oracle_concat_forward(float *out, const float *const *inputs, const int64_t *widths,
                           int64_t num_inputs, int64_t batch) {
  int64_t out_w = 0;
  for (int64_t i = 0; i < num_inputs; i++) out_w += widths[i];
  int64_t off = 0;
  for (int64_t i = 0; i < num_inputs; i++) {
    for (int64_t b = 0; b < batch; b++)
      for (int64_t d = 0; d < widths[i]; d++)
        out[b * out_w + off + d] = inputs[i][b * widths[i] + d];
    off += widths[i];
  }
}

/* ---- Concat backward: src/ops/concat.cu:325-366 with add_with_stride
 * (src/runtime/cuda_helper.cu:110-125): in_grad_i(b) += out_grad(b)[off_i : off_i+w_i].     */
void oracle_concat_backward(const float *out_grad, float *const *in_grads,
                            const int64_t *widths, int64_t num_inputs, int64_t batch) {
  int64_t out_w = 0;
  for (int64_t i = 0; i < num_inputs; i++) out_w += widths[i];
  int64_t off = 0;
  for (int64_t i = 0; i < num_inputs; i++) {
    for (int64_t b = 0; b < batch; b++)
      for (int64_t d = 0; d < widths[i]; d++) {
        volatile float s = in_grads[i][b * widths[i] + d] + out_grad[b * out_w + off + d];
        in_grads[i][b * widths[i] + d] = s;
      }
    off += widths[i];
  }
}

/* ---- DataLoader::load_sparse_input: examples/cpp/DLRM/dlrm.cu:19-58 (the host loop :38-45) ----
 * batch[i*in_dim + j] = full[idxs[i] * num_sparse_inputs*in_dim + my_input_idx*in_dim + j], generalised
 * to a per-table column block [col, col + width) of a sample row of row_len int64 (the reference
 * has col = my_input_idx*in_dim, width = in_dim, row_len = num_sparse_inputs*in_dim); idxs == NULL
 * means the consecutive ids first_sample + i that DataLoader::next_batch fills
 * (examples/cpp/DLRM/dlrm.cc:497-498).                                                         */
void oracle_load_sparse_input(const int64_t *full, int64_t row_len, int64_t col, int64_t width,
                              const int32_t *idxs, int64_t first_sample, int64_t B, int64_t *batch) {
  for (int64_t i = 0; i < B; i++) {
    int64_t row = idxs ? (int64_t)idxs[i] : first_sample + i;
    int64_t full_offset = row * row_len + col;
    int64_t batch_offset = i * width;
    for (int64_t j = 0; j < width; j++) batch[batch_offset + j] = full[full_offset + j];
  }
}

/* Number of distinct rows touched (U in the algorithmic-bytes formula).
 * Sort-and-count; scratch is allocated here.                                                 */
static int cmp_i64(const void *a, const void *b) {
  int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
  return (x > y) - (x < y);
}
int64_t oracle_count_unique(const int64_t *idx, int64_t n) {
  if (n <= 0) return 0;
  int64_t *t = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
  if (!t) return -1;
  memcpy(t, idx, sizeof(int64_t) * (size_t)n);
  qsort(t, (size_t)n, sizeof(int64_t), cmp_i64);
  int64_t u = 1;
  for (int64_t i = 1; i < n; i++) u += (t[i] != t[i - 1]);
  free(t);
  return u;
}
