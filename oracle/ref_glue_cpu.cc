/* oracle/ref_glue_cpu.cc — TEST INFRASTRUCTURE, not product code.
 *
 * C entry points over the reference's OWN CPU embedding code.  oracle/Makefile compiles
 * /root/reference/src/ops/embedding.cc UNMODIFIED, where it lies, against oracle/shim/legion.h
 * and the reference's own include/ tree, and links it with this file into
 * oracle/_ref/libffref_cpu.so.  Nothing here restates arithmetic: the wrappers only dress
 * plain pointers up as the (Task, PhysicalRegion) arguments the reference's Legion task
 * bodies expect, the way Legion would on a LOC_PROC.
 *
 * Reference entry points driven (all in src/ops/embedding.cc):
 *   Embedding::forward_task_cpu   :377-415   Embedding::backward_task_cpu  :417-453
 *   EmbeddingLookup_int64_t_float_float__avx2_fma :23-319 (called directly for bags > 1 and
 *   for the normalize_by_lengths mean, which the task wrapper never requests)
 */
#include "model.h"

#include <chrono>
#include <map>
#include <string>
#include <thread>

/* defined (non-static) in the reference translation unit */
void EmbeddingLookup_int64_t_float_float__avx2_fma(const int block_size, const int output_size,
                                                   const int index_size, const int data_size,
                                                   const float *input, const int64_t *indices,
                                                   const int *lengths, const float *weight,
                                                   bool normalize_by_lengths, float *out);

namespace {

/* Legion logical layout of the three regions (dim0 fastest), src/ops/embedding.cc:387-398:
 * input [L][B], output [D][B], weight rect dim1 = D (rect_weight.hi[1]-lo[1]+1 == out_dim). */
struct TaskFrame {
  Legion::Task task;
  std::vector<Legion::PhysicalRegion> regions;
  Legion::Runtime runtime;
  void add(void *base, Legion::coord_t d0, Legion::coord_t d1) {
    Legion::Rect<2> r(Legion::Point<2>(0, 0), Legion::Point<2>(d0 - 1, d1 - 1));
    Legion::Domain dom(r);
    Legion::LogicalRegion lr{Legion::IndexSpace(dom)};
    Legion::RegionRequirement req(lr, Legion::READ_WRITE, Legion::EXCLUSIVE, lr);
    task.regions.push_back(req);
    regions.push_back(Legion::PhysicalRegion(base, dom));
  }
};

}  // namespace

extern "C" {

/* Embedding::forward_task_cpu on one table: idx int64 [B][1], out float [B][D], W [R][D].
 * The reference requires L == 1 (assert at :408) and, unless built with NDEBUG, every
 * index < 1,000,000 (:71-73 with data_size hard-coded at :404).                              */
void ffref_cpu_forward_task(const int64_t *idx, float *out, const float *W, int64_t L, int64_t D,
                            int64_t B, int64_t R) {
  TaskFrame f;
  f.add((void *)idx, L, B);
  f.add((void *)out, D, B);
  f.add((void *)W, R, D);
  Embedding::forward_task_cpu(&f.task, f.regions, NULL, &f.runtime);
}

/* Embedding::backward_task_cpu: Wgrad[idx[b]][d] += outgrad[b][d], serial, bag 1 only.       */
void ffref_cpu_backward_task(const int64_t *idx, const float *outgrad, float *Wgrad, int64_t L,
                             int64_t D, int64_t B, int64_t R) {
  TaskFrame f;
  f.add((void *)idx, L, B);
  f.add((void *)outgrad, D, B);
  f.add((void *)Wgrad, R, D);
  Embedding::backward_task_cpu(&f.task, f.regions, NULL, &f.runtime);
}

/* The AVX2 lookup called directly with fixed-width bags of L and optional mean
 * normalisation (normalize_by_lengths) — the reference's CPU definition of SUM / true mean.  */
void ffref_cpu_lookup(const int64_t *idx, float *out, const float *W, int64_t L, int64_t D,
                      int64_t B, int64_t R, int normalize) {
  std::vector<int> lengths((size_t)B, (int)L);
  EmbeddingLookup_int64_t_float_float__avx2_fma((int)D, (int)B, (int)(B * L), (int)R, W, idx,
                                                lengths.data(), NULL, normalize != 0, out);
}

/* CPU baseline driver: the hetero strategy gives every embedding a [1,1] grid on CPU i % P
 * (examples/cpp/DLRM/strategies/dlrm_strategy_hetero.cc:28-36); Legion runs one task per
 * table on P LOC_PROCs (-ll:cpu P).  Here: P worker threads, table t on thread t % P, each
 * running forward_task_cpu then backward_task_cpu for its tables, `steps` times over
 * `nbatches` pre-generated index batches (idx[t] points at nbatches consecutive [B][L]
 * batches).  Returns wall seconds for all steps; fwd_only=1 skips the backward tasks.        */
double ffref_cpu_run_tables(int T, const int64_t *const *idx, float *const *out,
                            const float *const *W, const float *const *outgrad,
                            float *const *Wgrad, const int64_t *R, int64_t L, int64_t D,
                            int64_t B, int P, int steps, int nbatches, int fwd_only) {
  if (P < 1) P = 1;
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> workers;
  for (int p = 0; p < P; p++) {
    workers.emplace_back([=]() {
      for (int s = 0; s < steps; s++) {
        size_t boff = (size_t)(s % nbatches) * (size_t)(B * L);
        for (int t = p; t < T; t += P) {
          ffref_cpu_forward_task(idx[t] + boff, out[t], W[t], L, D, B, R[t]);
          if (!fwd_only)
            ffref_cpu_backward_task(idx[t] + boff, outgrad[t], Wgrad[t], L, D, B, R[t]);
        }
      }
    });
  }
  for (auto &w : workers) w.join();
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

/* ---- strategy files: src/runtime/strategy.cc:95-189 (declared in include/config.h) --------
 * ffref_strategy_load_op: the reference's load_strategies_from_file on `file`, then the entry of
 * op `name` (keyed by FFConfig::get_hash_id, strategy.cc:20-26).  out = {device_type, nDims,
 * dim[0..nDims), device_ids[0..num_parts)}; returns the number of ints written, -1 when the file
 * does not load, -2 when the op is not in it.                                                 */
int ffref_strategy_load_op(const char *file, const char *name, int *out, int max_out) {
  std::map<MappingTagID, ParallelConfig> strategies;
  if (!load_strategies_from_file(file, strategies)) return -1;
  std::map<MappingTagID, ParallelConfig>::const_iterator it = strategies.find(FFConfig::get_hash_id(name));
  if (it == strategies.end()) return -2;
  const ParallelConfig &c = it->second;
  int n = 1, k = 0;
  for (int j = 0; j < c.nDims; j++) n *= c.dim[j];
  if (2 + c.nDims + n > max_out) return -3;
  out[k++] = (int)c.device_type;
  out[k++] = c.nDims;
  for (int j = 0; j < c.nDims; j++) out[k++] = c.dim[j];
  for (int j = 0; j < n; j++) out[k++] = c.device_ids[j];
  return k;
}

/* ffref_strategy_save_tablewise: the strategy the reference's DLRM generator describes
 * (examples/cpp/DLRM/strategies/dlrm_strategy.cc:97-117,252-256: "embedding<i>" whole on GPU
 * i % G with a [1,1] grid; "concat" data-parallel over the G GPUs) written by the reference's own
 * save_strategies_to_file (strategy.cc:147-189).                                               */
int ffref_strategy_save_tablewise(const char *file, int num_tables, int G) {
  std::map<std::string, ParallelConfig> strategies;
  for (int i = 0; i < num_tables; i++) {
    ParallelConfig c;
    c.device_type = ParallelConfig::GPU;
    c.nDims = 2;
    c.dim[0] = 1;
    c.dim[1] = 1;
    c.device_ids[0] = i % G;
    strategies["embedding" + std::to_string(i)] = c;
  }
  ParallelConfig c;
  c.device_type = ParallelConfig::GPU;
  c.nDims = 2;
  c.dim[0] = 1;
  c.dim[1] = G;
  for (int g = 0; g < G; g++) c.device_ids[g] = g;
  strategies["concat"] = c;
  return save_strategies_to_file(file, strategies) ? 0 : -1;
}

int ffref_cpu_asserts_enabled(void) {
#ifdef NDEBUG
  return 0;
#else
  return 1;
#endif
}

}  /* extern "C" */
