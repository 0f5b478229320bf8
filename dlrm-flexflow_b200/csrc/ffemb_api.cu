// ffemb_api.cu — the extern "C" surface of libffemb.so (include/ffemb.h): argument checking,
// descriptor packing, error text.  No allocation, no synchronisation, no device selection.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "ffemb_internal.cuh"

namespace ffemb {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
  if (e == cudaSuccess) return FFEMB_OK;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return FFEMB_ECUDA;
}

const Tuning &tuning() {
  static const Tuning t = [] {
    auto num = [](const char *name) -> long long {
      const char *e = getenv(name);
      return (e && *e) ? atoll(e) : -1;
    };
    Tuning x;
    x.stream_mb = num("FFEMB_STREAM_MB");
    x.l1_kb = num("FFEMB_L1_KB");
    x.fwd_perm = (int)num("FFEMB_FWD_PERM");
    x.fwd_avg_div = (int)num("FFEMB_FWD_AVG_DIV");
    x.fwd_target = (int)num("FFEMB_FWD_TARGET");
    x.r1_mix = (int)num("FFEMB_R1_MIX");
    x.r1_flags = (int)num("FFEMB_R1_FLAGS");
    x.r1_pipe = (int)num("FFEMB_R1_PIPE");
    x.r1_ring = (int)num("FFEMB_R1_RING");
    x.small = (int)num("FFEMB_SMALL");
    x.small_fork = (int)num("FFEMB_SMALL_FORK");
    x.up_fused = (int)num("FFEMB_UP_FUSED");
    x.hot = (int)num("FFEMB_HOT");
    x.slab_mb = (int)num("FFEMB_SLAB_MB");
    x.own_sort = (int)num("FFEMB_OWN_SORT");
    x.xpush = (int)num("FFEMB_XPUSH");
    x.prescale = (int)num("FFEMB_PRESCALE");
    x.fwd_bulk = (int)num("FFEMB_FWD_BULK");
    x.sort_aligned = (int)num("FFEMB_SORT_ALIGNED");
    return x;
  }();
  return t;
}

// ---- profiling ------------------------------------------------------------------------------
struct ProfRec { int phase; cudaEvent_t a, b; };
static std::mutex g_prof_mu;
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_open;  // begin events awaiting their end, per nesting
static std::atomic<long long> g_launches{0};
static long long g_phase_launches[FFEMB_NUM_PHASES] = {0};

void prof_begin(int phase, cudaStream_t s) {
  (void)phase;
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, s);
  g_prof_open.push_back(e);
}

void prof_end(int phase, int launches, cudaStream_t s) {
  g_launches.fetch_add(launches, std::memory_order_relaxed);
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (g_prof_open.empty()) return;
  ProfRec r;
  r.phase = phase;
  r.a = g_prof_open.back();
  g_prof_open.pop_back();
  if (cudaEventCreate(&r.b) != cudaSuccess) { cudaEventDestroy(r.a); return; }
  cudaEventRecord(r.b, s);
  g_prof_recs.push_back(r);
  g_phase_launches[phase] += launches;
}

// Validate and pack a batched call.  Mirrors the shape checks of the reference's task bodies
// (src/ops/embedding.cu:251-258: input/output batch agree, weight/output width agree) as far
// as raw pointers allow, plus what the reference leaves to UB.
static int pack(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B, int aggr,
                TableBatch &tb, bool need_out = true, bool need_table = true) {
  if (!aggr_valid(aggr)) {
    // AGGR_MODE_NONE reaches assert(aggr == AGGR_MODE_AVG) in the reference kernels
    // (src/ops/embedding.cu:185,209)
    set_error("invalid aggr %d (expected 21 SUM, 22 AVG, 122 MEAN)", aggr);
    return FFEMB_EINVAL;
  }
  if (!tables || num_tables < 0 || num_tables > FFEMB_MAX_TABLES || D <= 0 || B < 0 ||
      D > (1 << 20) || B >= (1ll << 31)) {
    set_error("invalid batched call (tables=%p num_tables=%d D=%lld B=%lld)", (const void *)tables,
              num_tables, (long long)D, (long long)B);
    return FFEMB_EINVAL;
  }
  memset(&tb, 0, sizeof(tb));
  tb.num_tables = num_tables;
  uint64_t rows = 0, lookups = 0;
  for (int t = 0; t < num_tables; t++) {
    const ffemb_table_t &x = tables[t];
    if (x.L < 0 || x.R < 0 ||
        (B > 0 && ((need_out && !x.out) || (need_table && (!x.table || (x.L > 0 && !x.idx)))))) {
      set_error("table %d: invalid descriptor (idx=%p out=%p table=%p R=%lld L=%d)", t,
                (const void *)x.idx, (void *)x.out, (void *)x.table, (long long)x.R, x.L);
      return FFEMB_EINVAL;
    }
    tb.t[t].idx = x.idx;
    tb.t[t].out = x.out;
    tb.t[t].table = x.table;
    tb.t[t].R = x.R;
    tb.t[t].L = x.L;
    tb.t[t].lookup_base = (uint32_t)lookups;
    tb.row_base[t] = rows;
    rows += (uint64_t)x.R;
    lookups += (uint64_t)B * (uint64_t)x.L;
    if (lookups >= (1ull << 31)) {
      set_error("more than 2^31 lookups in one call; split the tables");
      return FFEMB_EINVAL;
    }
  }
  tb.row_base[num_tables] = rows;
  tb.og_stride = D;
  return FFEMB_OK;
}

static int pack1(const int64_t *idx, const float *out, const float *table, int64_t L, int64_t D,
                 int64_t B, int64_t R, int aggr, TableBatch &tb) {
  if (L < 0 || L >= (1ll << 31)) {
    set_error("invalid bag width %lld", (long long)L);
    return FFEMB_EINVAL;
  }
  ffemb_table_t x;
  x.idx = idx;
  x.out = const_cast<float *>(out);
  x.table = const_cast<float *>(table);
  x.R = R;
  x.L = (int32_t)L;
  x.reserved = 0;
  return pack(&x, 1, D, B, aggr, tb);
}

__global__ void check_indices_kernel(const int64_t *idx, int64_t n, int64_t R,
                                     unsigned long long *bad) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long mine = 0;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) mine += (idx[i] < 0 || idx[i] >= R);
  if (mine) atomicAdd(bad, mine);
}

}  // namespace ffemb

using namespace ffemb;

extern "C" {

const char *ffemb_last_error(void) { return g_err; }
int ffemb_version(void) { return FFEMB_VERSION; }

int ffemb_forward(const int64_t *idx, float *out, const float *table, int64_t L, int64_t D,
                  int64_t B, int aggr, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack1(idx, out, table, L, D, B, 0, aggr, tb);
  if (rc) return rc;
  OutMap om;
  memset(&om, 0, sizeof(om));
  return launch_forward(tb, D, B, aggr, om, (cudaStream_t)stream);
}

int ffemb_forward_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                          int aggr, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb);
  if (rc) return rc;
  OutMap om;
  memset(&om, 0, sizeof(om));
  return launch_forward(tb, D, B, aggr, om, (cudaStream_t)stream);
}

int ffemb_forward_exchange(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                           int aggr, const ffemb_exchange_t *ex, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb, /*need_out=*/false);
  if (rc) return rc;
  if (!ex || ex->num_peers < 1 || ex->num_peers > FFEMB_MAX_PEERS || ex->batch_per_peer < 1 ||
      ex->row_stride < D || ex->batch_per_peer * ex->num_peers < B) {
    set_error("forward_exchange: invalid exchange descriptor");
    return FFEMB_EINVAL;
  }
  OutMap om;
  memset(&om, 0, sizeof(om));
  for (int h = 0; h < ex->num_peers; h++) {
    if (!ex->peer_x[h] && B > (int64_t)h * ex->batch_per_peer) {
      set_error("forward_exchange: peer_x[%d] is NULL", h);
      return FFEMB_EINVAL;
    }
    om.peer[h] = ex->peer_x[h];
  }
  om.npeers = ex->num_peers;
  om.ext = ex->batch_per_peer;
  om.stride = ex->row_stride;
  for (int t = 0; t < num_tables; t++) {
    if (tables[t].reserved < 0 || (int64_t)tables[t].reserved + D > ex->row_stride) {
      set_error("forward_exchange: table %d column offset %d outside a row of %lld floats", t,
                tables[t].reserved, (long long)ex->row_stride);
      return FFEMB_EINVAL;
    }
    tb.t[t].lookup_base = (uint32_t)tables[t].reserved;  // column offset of the table's block
  }
  return launch_forward(tb, D, B, aggr, om, (cudaStream_t)stream);
}

int ffemb_exchange_backward(const ffemb_table_t *tables, int num_tables, int64_t D, const float *dx,
                            int64_t n_samples, int64_t first_sample, int64_t row_stride,
                            ffemb_stream_t stream) {
  if (!tables || num_tables < 0 || num_tables > FFEMB_MAX_TABLES || D <= 0 || n_samples < 0 ||
      first_sample < 0 || row_stride < D || (n_samples > 0 && !dx)) {
    set_error("exchange_backward: invalid arguments");
    return FFEMB_EINVAL;
  }
  TableBatch tb;
  memset(&tb, 0, sizeof(tb));
  tb.num_tables = num_tables;
  for (int t = 0; t < num_tables; t++) {
    if (!tables[t].out || tables[t].reserved < 0 || (int64_t)tables[t].reserved + D > row_stride) {
      set_error("exchange_backward: table %d: invalid destination or column offset", t);
      return FFEMB_EINVAL;
    }
    tb.t[t].out = tables[t].out;
    tb.t[t].lookup_base = (uint32_t)tables[t].reserved;
  }
  return launch_exchange_backward(tb, D, dx, n_samples, first_sample, row_stride, (cudaStream_t)stream);
}

int ffemb_backward_dense_atomic(const int64_t *idx, const float *outgrad, float *wgrad, int64_t L,
                                int64_t D, int64_t B, int aggr, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack1(idx, outgrad, wgrad, L, D, B, 0, aggr, tb);
  if (rc) return rc;
  return launch_backward_atomic(tb, D, B, aggr, (cudaStream_t)stream);
}

int ffemb_backward_dense(const int64_t *idx, const float *outgrad, float *wgrad, int64_t L,
                         int64_t D, int64_t B, int64_t R, int aggr, void *workspace,
                         size_t workspace_bytes, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack1(idx, outgrad, wgrad, L, D, B, R, aggr, tb);
  if (rc) return rc;
  if (R <= 0 && B * L > 0) {
    set_error("backward_dense needs the row count R > 0");
    return FFEMB_EINVAL;
  }
  return launch_backward_sorted(tb, D, B, aggr, 0.f, APPLY_GRAD, workspace, workspace_bytes,
                                (cudaStream_t)stream);
}

int ffemb_backward_sgd(const int64_t *idx, const float *outgrad, float *table, int64_t L,
                       int64_t D, int64_t B, int64_t R, int aggr, float lr, void *workspace,
                       size_t workspace_bytes, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack1(idx, outgrad, table, L, D, B, R, aggr, tb);
  if (rc) return rc;
  if (R <= 0 && B * L > 0) {
    set_error("backward_sgd needs the row count R > 0");
    return FFEMB_EINVAL;
  }
  return launch_backward_sorted(tb, D, B, aggr, lr, APPLY_SGD, workspace, workspace_bytes,
                                (cudaStream_t)stream);
}

int ffemb_backward_sgd_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                               int aggr, float lr, void *workspace, size_t workspace_bytes,
                               ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb);
  if (rc) return rc;
  return launch_backward_sorted(tb, D, B, aggr, lr, APPLY_SGD, workspace, workspace_bytes,
                                (cudaStream_t)stream);
}

int ffemb_backward_sgd_padded(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                              int aggr, float lr, void *workspace, size_t workspace_bytes,
                              ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb);
  if (rc) return rc;
  tb.padded = 1;
  return launch_backward_sorted(tb, D, B, aggr, lr, APPLY_SGD, workspace, workspace_bytes,
                                (cudaStream_t)stream);
}

int ffemb_backward_prepare(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B, int aggr,
                           void *workspace, size_t workspace_bytes, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb, /*need_out=*/false);
  if (rc) return rc;
  return launch_backward_sorted(tb, D, B, aggr, 0.f, APPLY_SGD, workspace, workspace_bytes,
                                (cudaStream_t)stream, STAGE_SORT);
}

int ffemb_backward_sgd_prepared(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                                int aggr, float lr, void *workspace, size_t workspace_bytes,
                                ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb);
  if (rc) return rc;
  return launch_backward_sorted(tb, D, B, aggr, lr, APPLY_SGD, workspace, workspace_bytes,
                                (cudaStream_t)stream, STAGE_APPLY);
}

int64_t ffemb_sgd_touched_words(int64_t R) { return R <= 0 ? 0 : (R + 31) / 32; }

int ffemb_backward_sgdx_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                                int aggr, const ffemb_sgd_t *opt, float *const *V,
                                uint32_t *const *touched, void *workspace, size_t workspace_bytes,
                                ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb);
  if (rc) return rc;
  if (!opt || !touched || opt->momentum < 0.f || (opt->momentum > 0.f && !V)) {
    set_error("backward_sgdx: optimizer arguments missing (opt=%p V=%p touched=%p)", (const void *)opt,
              (const void *)V, (const void *)touched);
    return FFEMB_EINVAL;
  }
  OptArgs o;
  memset(&o, 0, sizeof(o));
  o.lr = opt->lr;
  o.wd = opt->weight_decay;
  o.mom = opt->momentum;
  o.nesterov = opt->nesterov;
  for (int t = 0; t < num_tables; t++) {
    if (tables[t].R <= 0) continue;
    if (!touched[t] || (uintptr_t)touched[t] % 4 != 0 ||
        (o.mom > 0.f && (!V[t] || (uintptr_t)V[t] % 16 != 0))) {
      set_error("backward_sgdx: table %d needs a touched bitmap%s", t,
                o.mom > 0.f ? " and a 16-byte aligned momentum buffer" : "");
      return FFEMB_EINVAL;
    }
    o.V[t] = o.mom > 0.f ? V[t] : nullptr;
    o.touched[t] = touched[t];
  }
  return launch_backward_sorted(tb, D, B, aggr, o.lr, APPLY_SGDX, workspace, workspace_bytes,
                                (cudaStream_t)stream, STAGE_ALL, &o);
}

int ffemb_backward_adam_batched(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                                int aggr, const ffemb_adam_t *opt, float *const *M, float *const *V,
                                uint32_t *const *touched, void *workspace, size_t workspace_bytes,
                                ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb);
  if (rc) return rc;
  if (!opt || !touched || !M || !V) {
    set_error("backward_adam: optimizer arguments missing (opt=%p M=%p V=%p touched=%p)", (const void *)opt,
              (const void *)M, (const void *)V, (const void *)touched);
    return FFEMB_EINVAL;
  }
  OptArgs o;
  memset(&o, 0, sizeof(o));
  o.adam = 1;
  o.alpha_t = opt->alpha_t;
  o.beta1 = opt->beta1;
  o.beta2 = opt->beta2;
  o.wd = opt->weight_decay;
  o.eps = opt->epsilon;
  for (int t = 0; t < num_tables; t++) {
    if (tables[t].R <= 0) continue;
    if (!touched[t] || (uintptr_t)touched[t] % 4 != 0 || !M[t] || !V[t] ||
        ((uintptr_t)M[t] | (uintptr_t)V[t]) % 16 != 0) {
      set_error("backward_adam: table %d needs a touched bitmap and 16-byte aligned M and V buffers", t);
      return FFEMB_EINVAL;
    }
    o.M[t] = M[t];
    o.V[t] = V[t];
    o.touched[t] = touched[t];
  }
  return launch_backward_sorted(tb, D, B, aggr, 0.f, APPLY_SGDX, workspace, workspace_bytes,
                                (cudaStream_t)stream, STAGE_ALL, &o);
}

int ffemb_forward_concat(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B, int aggr,
                         float *x, int64_t row_stride, ffemb_stream_t stream) {
  ffemb_exchange_t ex;
  memset(&ex, 0, sizeof(ex));
  ex.peer_x[0] = x;
  ex.num_peers = 1;
  ex.batch_per_peer = B > 0 ? B : 1;
  ex.row_stride = row_stride;
  return ffemb_forward_exchange(tables, num_tables, D, B, aggr, &ex, stream);
}

int ffemb_backward_sgd_concat(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t B,
                              int aggr, float lr, const float *dx, int64_t row_stride, int prepared,
                              void *workspace, size_t workspace_bytes, ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb, /*need_out=*/false);
  if (rc) return rc;
  if ((B > 0 && !dx) || row_stride < D || row_stride % 4 != 0 || (uintptr_t)dx % 16 != 0) {
    set_error("backward_sgd_concat: dx must be 16-byte aligned, row_stride >= D and a multiple of 4 "
              "(dx=%p row_stride=%lld D=%lld)", (const void *)dx, (long long)row_stride, (long long)D);
    return FFEMB_EINVAL;
  }
  for (int t = 0; t < num_tables; t++) {
    if (tables[t].reserved < 0 || tables[t].reserved % 4 != 0 || (int64_t)tables[t].reserved + D > row_stride) {
      set_error("backward_sgd_concat: table %d column offset %d (must be a multiple of 4, inside a row of %lld)",
                t, tables[t].reserved, (long long)row_stride);
      return FFEMB_EINVAL;
    }
    tb.t[t].out = const_cast<float *>(dx) + tables[t].reserved;
  }
  tb.og_stride = row_stride;
  return launch_backward_sorted(tb, D, B, aggr, lr, APPLY_SGD, workspace, workspace_bytes,
                                (cudaStream_t)stream, prepared ? STAGE_APPLY : STAGE_ALL);
}

int ffemb_backward_dense_batched(const ffemb_table_t *tables, int num_tables, int64_t D,
                                 int64_t B, int aggr, void *workspace, size_t workspace_bytes,
                                 ffemb_stream_t stream) {
  TableBatch tb;
  int rc = pack(tables, num_tables, D, B, aggr, tb);
  if (rc) return rc;
  return launch_backward_sorted(tb, D, B, aggr, 0.f, APPLY_GRAD, workspace, workspace_bytes,
                                (cudaStream_t)stream);
}

int ffemb_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = on != 0;
  return FFEMB_OK;
}

int ffemb_profile_collect(float *ms_per_phase, int64_t *launches_per_phase) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  float ms[FFEMB_NUM_PHASES] = {0};
  int rc = FFEMB_OK;
  for (auto &r : g_prof_recs) {
    float t = 0.f;
    cudaError_t e = cudaEventSynchronize(r.b);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&t, r.a, r.b);
    if (e != cudaSuccess) rc = cuda_fail(e, "profile collect");
    else ms[r.phase] += t;
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  g_prof_recs.clear();
  for (auto e : g_prof_open) cudaEventDestroy(e);
  g_prof_open.clear();
  for (int p = 0; p < FFEMB_NUM_PHASES; p++) {
    if (ms_per_phase) ms_per_phase[p] = ms[p];
    if (launches_per_phase) launches_per_phase[p] = g_phase_launches[p];
    g_phase_launches[p] = 0;
  }
  return rc;
}

int64_t ffemb_launch_count(void) { return (int64_t)g_launches.load(); }

int ffemb_backward_path(int64_t R, int64_t L, int64_t D, int64_t B) { return backward_path(R, L, D, B); }

size_t ffemb_backward_workspace(int64_t num_lookups, int64_t total_rows, int64_t D) {
  return backward_workspace_bytes(num_lookups, total_rows, D);
}

size_t ffemb_backward_workspace_avg(int64_t num_lookups, int64_t total_rows, int64_t D, int64_t B, int num_tables) {
  const size_t base = backward_workspace_bytes(num_lookups, total_rows, D);
  if (base == 0 || B < 0 || num_tables < 0 || num_tables > FFEMB_MAX_TABLES) return 0;
  return base + (size_t)num_tables * prescale_bytes(B, D);
}

int ffemb_load_sparse_inputs(const ffemb_table_t *tables, int num_tables, const int64_t *full,
                             int64_t row_len, const int32_t *sample_ids, int64_t first_sample,
                             int64_t B, ffemb_stream_t stream) {
  if (!tables || num_tables < 0 || num_tables > FFEMB_MAX_TABLES || row_len <= 0 || row_len >= (1ll << 31) ||
      B < 0 || first_sample < 0 || (B > 0 && !full)) {
    set_error("load_sparse_inputs: invalid arguments (tables=%p num_tables=%d row_len=%lld B=%lld)",
              (const void *)tables, num_tables, (long long)row_len, (long long)B);
    return FFEMB_EINVAL;
  }
  for (int t = 0; t < num_tables; t++) {
    const ffemb_table_t &x = tables[t];
    if (x.L < 0 || x.reserved < 0 || (int64_t)x.reserved + x.L > row_len || (B > 0 && x.L > 0 && !x.idx)) {
      set_error("load_sparse_inputs: table %d: columns [%d, %d) outside a row of %lld, or no destination", t,
                x.reserved, x.reserved + x.L, (long long)row_len);
      return FFEMB_EINVAL;
    }
  }
  return launch_load_sparse(tables, num_tables, full, row_len, sample_ids, first_sample, B, (cudaStream_t)stream);
}

int ffemb_check_indices(const int64_t *idx, int64_t n, int64_t R, int64_t *bad_count,
                        ffemb_stream_t stream) {
  if (!bad_count || n < 0 || (n > 0 && !idx)) {
    set_error("check_indices: invalid arguments");
    return FFEMB_EINVAL;
  }
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(bad_count, 0, sizeof(int64_t), s);
  if (e != cudaSuccess) return cuda_fail(e, "memset");
  if (n > 0) {
    int blocks = (int)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
    check_indices_kernel<<<blocks, 256, 0, s>>>(idx, n, R, (unsigned long long *)bad_count);
  }
  return cuda_fail(cudaGetLastError(), "check_indices launch");
}

}  // extern "C"
