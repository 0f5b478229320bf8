/* oracle/ref_glue_gpu.cu — TEST INFRASTRUCTURE, not product code.
 *
 * C entry points over the reference's OWN GPU code for the embedding path, compiled
 * UNMODIFIED for sm_100a by oracle/Makefile from where it lies under /root/reference:
 *   src/ops/embedding.cu              Embedding::forward_task/backward_task, forward_kernel/
 *                                     backward_kernel, embed_forward/embed_backward
 *   src/runtime/accessor_kernel.cu    TensorAccessorR/W (region -> dense pointer)
 *   src/runtime/cuda_helper.cu        get_legion_stream, assign_kernel, print_tensor
 *   src/runtime/initializer_kernel.cu ZeroInitializer::init_task (the per-step zero-fill)
 *   src/runtime/optimizer_kernel.cu   SGDOptimizer::ps_update_task -> sgd_update
 * against oracle/shim/legion.h + the reference's include/ tree -> oracle/_ref/libffref_gpu.so.
 * It is (1) the GPU parity oracle (the shipping behaviour of the path, incl. multi-hot and the
 * AVG-in-loop quirk) and (2) the "reference on B200" timing row.  Nothing here restates
 * arithmetic; it only builds the (Task, PhysicalRegion) frames Legion would hand the task
 * bodies, with device pointers, and defines the link-time leftovers of FlexFlow's control
 * plane (Op/Tensor/FFModel/Simulator members the embedding TU references but this path never
 * calls) as aborting stubs.
 */
#include "model.h"
#include "cuda_helper.h"
#ifndef FFEMB_DROPIN
#include "dlrm.h" /* examples/cpp/DLRM: DataLoader, SampleIdxs (only in libffref_gpu.so) */
#endif

#include <cstdio>
#include <cstdlib>

#define FFREF_DEAD(what)                                                        \
  do {                                                                          \
    fprintf(stderr, "oracle/ref_glue_gpu: control-plane stub called: %s\n", what); \
    abort();                                                                    \
  } while (0)

/* ---- link-time leftovers (never executed on this path) ---------------------------------- */
OpMeta::OpMeta(FFHandler _handle) : handle(_handle), profiling(false) {}
Tensor::Tensor() { numDim = 0; owner_op = NULL; owner_idx = 0; }
Tensor &Tensor::operator=(const Tensor &) { FFREF_DEAD("Tensor::operator="); }
size_t Tensor::get_volume() const { FFREF_DEAD("Tensor::get_volume"); }
bool Tensor::get_input_sub_tensor(const ParallelConfig &, Tensor &, OperatorType) {
  FFREF_DEAD("Tensor::get_input_sub_tensor");
}
bool Tensor::get_output_sub_tensor(const ParallelConfig &, Tensor &, OperatorType) {
  FFREF_DEAD("Tensor::get_output_sub_tensor");
}
Op::Op(FFModel &, OperatorType t, const Op *, const char *, const Tensor &) : op_type(t) {
  FFREF_DEAD("Op::Op");
}
ParallelConfig Op::get_random_parallel_config(const FFModel &) const { FFREF_DEAD("Op"); }
ParallelConfig Op::get_data_parallel_config(const FFModel &) const { FFREF_DEAD("Op"); }
bool Op::is_valid_parallel_config(const FFModel &, const ParallelConfig &) const { FFREF_DEAD("Op"); }
bool Op::is_adoptable_parallel_config(FFModel const &, ParallelConfig const &) const { FFREF_DEAD("Op"); }
Domain Op::get_input_tensor_shape(const ParallelConfig &, int, int) { FFREF_DEAD("Op"); }
Domain Op::get_output_tensor_shape(const ParallelConfig &, int, int) { FFREF_DEAD("Op"); }
Domain Op::get_weight_tensor_shape(const ParallelConfig &, int, int) { FFREF_DEAD("Op"); }
bool Op::can_inplace_output() { return false; }
bool Op::has_inplace_output() { return false; }
void Op::do_inplace_output() { FFREF_DEAD("Op::do_inplace_output"); }
void Op::inner_measure_operator_cost(Simulator *, std::function<void()> const &,
                                     std::function<void()> const &, CostMetrics &) {
  FFREF_DEAD("Op::inner_measure_operator_cost");
}
void *Simulator::allocate(size_t, DataType) { FFREF_DEAD("Simulator::allocate"); }
void Simulator::free_all() { FFREF_DEAD("Simulator::free_all"); }
IndexSpace FFModel::get_or_create_task_is(int, const std::string &) { FFREF_DEAD("FFModel"); }
template <>
Tensor FFModel::create_tensor<2>(const int *, DataType, const Op *, bool) { FFREF_DEAD("FFModel"); }
template <>
Parameter FFModel::create_linear_weight<2, 2>(Op *, const int *, DataType, Initializer *, bool,
                                              ParameterSyncType) { FFREF_DEAD("FFModel"); }
template <>
void FFModel::create_disjoint_partition<2>(const Tensor &, const IndexSpaceT<2> &,
                                           LogicalPartition &, LogicalPartition &) {
  FFREF_DEAD("FFModel");
}
size_t FFConfig::get_hash_id(const std::string &) { return 0; }
bool FFConfig::find_parallel_config(int, const std::string &, ParallelConfig &) const {
  FFREF_DEAD("FFConfig::find_parallel_config");
}
Optimizer::Optimizer(const FFModel *_model) : model(_model) {}
SGDOptimizer::SGDOptimizer(const FFModel *_model, double _lr, double _momentum, bool _nesterov,
                           double _weight_decay)
    : Optimizer(_model), lr(_lr), momentum(_momentum), nesterov(_nesterov),
      weight_decay(_weight_decay) {}
AdamOptimizer::AdamOptimizer(const FFModel *_model, double _alpha, double _beta1, double _beta2,
                             double _weight_decay, double _epsilon)
    : Optimizer(_model), alpha(_alpha), beta1(_beta1), beta2(_beta2), weight_decay(_weight_decay),
      epsilon(_epsilon), alpha_t(_alpha), beta1_t(1.0f), beta2_t(1.0f) {}
void AdamOptimizer::init(void) {}
void AdamOptimizer::next(void) {}
void AdamOptimizer::update(const Parameter *) { FFREF_DEAD("AdamOptimizer::update"); }
void AdamOptimizer::set_weight_decay(double w) { weight_decay = w; }
void SGDOptimizer::init(void) {}
void SGDOptimizer::next(void) {}
void SGDOptimizer::update(const Parameter *) { FFREF_DEAD("SGDOptimizer::update"); }

namespace {
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
void *g_workspace = NULL;  /* FFHandler::workSpace stand-in (set by ffref_gpu_set_workspace) */
size_t g_workspace_bytes = 0;
EmbeddingMeta *make_meta(int aggr) {
  FFHandler h;
  memset(&h, 0, sizeof(h));
  h.workSpace = g_workspace;
  h.workSpaceSize = g_workspace_bytes;
  EmbeddingMeta *m = new EmbeddingMeta(h);
  m->profiling = false;
  m->aggr = (AggrMode)aggr;
  return m;
}
}  // namespace

#ifdef FFEMB_DROPIN
/* the fused leaf body of dlrm-flexflow_b200/csrc/flexflow/embedding_ffemb.cu */
void ffemb_embedding_backward_sgd_task(const Legion::Task *task,
                                       const std::vector<Legion::PhysicalRegion> &regions,
                                       Legion::Context ctx, Legion::Runtime *runtime);
/* the per-GPU batched leaf bodies of dlrm-flexflow_b200/csrc/flexflow/embedding_group_ffemb.cu */
struct FFEmbGroupArgs { int num_tables; float lr; int prepared; int parity; };
void ffemb_embedding_group_forward_task(const Legion::Task *, const std::vector<Legion::PhysicalRegion> &,
                                        Legion::Context, Legion::Runtime *);
void ffemb_embedding_group_prepare_task(const Legion::Task *, const std::vector<Legion::PhysicalRegion> &,
                                        Legion::Context, Legion::Runtime *);
void ffemb_embedding_group_backward_sgd_task(const Legion::Task *, const std::vector<Legion::PhysicalRegion> &,
                                             Legion::Context, Legion::Runtime *);
#endif

extern "C" {

/* The per-GPU scratch Legion would hand every task through FFHandler (1 GB by default,
 * src/runtime/model.cc:2219, src/runtime/model.cu:97-109).                                   */
void ffref_gpu_set_workspace(void *ws, size_t bytes) {
  g_workspace = ws;
  g_workspace_bytes = bytes;
}

#ifdef FFEMB_DROPIN
/* Only in libffdropin_gpu.so (the drop-in's task bodies linked in place of the reference's):
 * the opt-in fused backward + SGD leaf task; regions[2] is the table itself.                 */
void ffdropin_gpu_backward_sgd_task(const int64_t *idx, const float *outgrad, float *W, int64_t L,
                                    int64_t D, int64_t B, int64_t R, int aggr, float lr) {
  TaskFrame f;
  f.add((void *)idx, L, B);
  f.add((void *)outgrad, D, B);
  f.add((void *)W, R, D);
  EmbeddingMeta *m = make_meta(aggr);
  f.task.local_args = &m;
  f.task.local_arglen = sizeof(m);
  f.task.args = &lr;
  f.task.arglen = sizeof(lr);
  ffemb_embedding_backward_sgd_task(&f.task, f.regions, NULL, &f.runtime);
  delete m;
}
#endif

#ifdef FFEMB_DROPIN
/* The per-GPU batched leaf tasks behind the frames Legion would hand them: one task, the regions of
 * `n` tables listed back to back.  what: 0 forward (idx, out, W), 1 prepare (idx, W), 2 backward + SGD
 * (idx, outgrad, W RW; prepared = the prepare task ran before).                                   */
void ffdropin_gpu_group_task(int what, int n, const int64_t *const *idx, float *const *out, float *const *W,
                             const int64_t *L, int64_t D, int64_t B, const int64_t *R, int aggr, float lr,
                             int prepared) {
  TaskFrame f;
  for (int t = 0; t < n; t++) {
    f.add((void *)idx[t], L[t], B);
    if (what != 1) f.add((void *)out[t], D, B);
    f.add((void *)W[t], R[t], D);
  }
  EmbeddingMeta *m = make_meta(aggr);
  f.task.local_args = &m;
  f.task.local_arglen = sizeof(m);
  FFEmbGroupArgs a = {n, lr, prepared, 0};
  f.task.args = &a;
  f.task.arglen = sizeof(a);
  if (what == 0) ffemb_embedding_group_forward_task(&f.task, f.regions, NULL, &f.runtime);
  else if (what == 1) ffemb_embedding_group_prepare_task(&f.task, f.regions, NULL, &f.runtime);
  else ffemb_embedding_group_backward_sgd_task(&f.task, f.regions, NULL, &f.runtime);
  delete m;
}
#endif

/* Embedding::forward_task (src/ops/embedding.cu:238-272) on device pointers; stream 0
 * (built with DISABLE_LEGION_CUDA_HIJACK, src/runtime/cuda_helper.cu:22-30).               */
void ffref_gpu_forward_task(const int64_t *idx, float *out, const float *W, int64_t L, int64_t D,
                            int64_t B, int64_t R, int aggr) {
  TaskFrame f;
  f.add((void *)idx, L, B);
  f.add((void *)out, D, B);
  f.add((void *)W, R, D);
  EmbeddingMeta *m = make_meta(aggr);
  f.task.local_args = &m;
  f.task.local_arglen = sizeof(m);
  Embedding::forward_task(&f.task, f.regions, NULL, &f.runtime);
  delete m;
}

/* Embedding::backward_task (src/ops/embedding.cu:322-354): atomicAdd into Wgrad.            */
void ffref_gpu_backward_task(const int64_t *idx, const float *outgrad, float *Wgrad, int64_t L,
                             int64_t D, int64_t B, int64_t R, int aggr) {
  TaskFrame f;
  f.add((void *)idx, L, B);
  f.add((void *)outgrad, D, B);
  f.add((void *)Wgrad, R, D);
  EmbeddingMeta *m = make_meta(aggr);
  f.task.local_args = &m;
  f.task.local_arglen = sizeof(m);
  Embedding::backward_task(&f.task, f.regions, NULL, &f.runtime);
  delete m;
}

/* The static raw-pointer wrappers (the seam, include/model.h:1169-1186), stream-explicit,
 * for timing without the task frame.                                                        */
void ffref_gpu_forward_kernel(const int64_t *idx, float *out, const float *W, int L, int D, int B,
                              int aggr, void *stream) {
  Embedding::forward_kernel(idx, out, W, L, D, B, (AggrMode)aggr, B * D, (cudaStream_t)stream);
}
void ffref_gpu_backward_kernel(const int64_t *idx, const float *outgrad, float *Wgrad, int L,
                               int D, int B, int aggr, void *stream) {
  Embedding::backward_kernel(idx, outgrad, Wgrad, L, D, B, (AggrMode)aggr, B * D,
                             (cudaStream_t)stream);
}

#ifndef FFEMB_DROPIN
/* DataLoader::load_sparse_input (examples/cpp/DLRM/dlrm.cu:19-58) as DataLoader::next_batch launches
 * it for table `my_input_idx` (examples/cpp/DLRM/dlrm.cc:482-520): regions[0] = the full dataset
 * [num_samples][num_sparse_inputs*in_dim] in HOST memory (zero-copy there; the task reads it on the
 * host), regions[1] = the table's batch input [B][in_dim] in device memory; task->args = the int
 * hash num_sparse_inputs*MAX_NUM_EMB + my_input_idx; local_args = SampleIdxs with consecutive ids
 * from first_sample.  B <= MAX_NUM_SAMPLES (65536, dlrm.h:17).                                  */
int ffref_gpu_load_sparse_input(const int64_t *full_host, int64_t num_samples, int num_sparse_inputs,
                                int in_dim, int my_input_idx, int first_sample, int B,
                                int64_t *batch_dev) {
  if (B > MAX_NUM_SAMPLES) return 1;
  TaskFrame f;
  f.add((void *)full_host, (Legion::coord_t)num_sparse_inputs * in_dim, num_samples);
  f.add((void *)batch_dev, in_dim, B);
  int hash = num_sparse_inputs * MAX_NUM_EMB + my_input_idx;
  SampleIdxs *meta = new SampleIdxs;
  meta->num_samples = B;
  for (int i = 0; i < B; i++) meta->idxs[i] = first_sample + i;
  f.task.args = &hash;
  f.task.arglen = sizeof(hash);
  f.task.local_args = meta;
  f.task.local_arglen = sizeof(SampleIdxs);
  DataLoader::load_sparse_input(&f.task, f.regions, NULL, &f.runtime);
  delete meta;
  return 0;
}
#endif

/* ZeroInitializer::init_task (src/runtime/initializer_kernel.cu:209-241) over one region —
 * what Op::zero_grad launches on the weight-grad every step (src/runtime/model.cc:466-490).
 * Note: the reference's GET_BLOCKS takes int (include/cuda_helper.h:58); n must be < 2^31.   */
void ffref_gpu_zero_task(float *p, int64_t d0, int64_t d1) {
  TaskFrame f;
  f.add((void *)p, d0, d1);
  ZeroInitializer::init_task(&f.task, f.regions, NULL, &f.runtime);
}

/* SGDOptimizer::ps_update_task (src/runtime/optimizer_kernel.cu:43-111) with momentum == 0:
 * regions[0] = weight-grad, regions[1] = weight; launches sgd_update over all d0*d1 floats.  */
void ffref_gpu_sgd_task(const float *Wgrad, float *W, int64_t d0, int64_t d1, double lr,
                        double weight_decay) {
  TaskFrame f;
  f.add((void *)Wgrad, d0, d1);
  f.add((void *)W, d0, d1);
  SGDOptimizer opt(NULL, lr, 0.0, false, weight_decay);
  f.task.args = &opt;
  f.task.arglen = sizeof(opt);
  SGDOptimizer::ps_update_task(&f.task, f.regions, NULL, &f.runtime);
}

/* The same task with the optimizer's momentum / nesterov / weight decay: with momentum > 0 the task
 * takes a third region, the momentum buffer V (src/runtime/optimizer_kernel.cu:49-56,84-90).      */
void ffref_gpu_sgd_task_full(const float *Wgrad, float *W, float *V, int64_t d0, int64_t d1, double lr,
                             double weight_decay, double momentum, int nesterov) {
  TaskFrame f;
  f.add((void *)Wgrad, d0, d1);
  f.add((void *)W, d0, d1);
  if (momentum > 0.0) f.add((void *)V, d0, d1);
  SGDOptimizer opt(NULL, lr, momentum, nesterov != 0, weight_decay);
  f.task.args = &opt;
  f.task.arglen = sizeof(opt);
  SGDOptimizer::ps_update_task(&f.task, f.regions, NULL, &f.runtime);
}

/* AdamOptimizer::ps_update_task (src/runtime/optimizer_kernel.cu:228-290): regions = weight-grad,
 * weight, V, M; alpha_t as AdamOptimizer::next left it (passed in).                              */
void ffref_gpu_adam_task(const float *Wgrad, float *W, float *V, float *M, int64_t d0, int64_t d1,
                         double alpha_t, double beta1, double beta2, double weight_decay, double epsilon) {
  TaskFrame f;
  f.add((void *)Wgrad, d0, d1);
  f.add((void *)W, d0, d1);
  f.add((void *)V, d0, d1);
  f.add((void *)M, d0, d1);
  AdamOptimizer *opt = new AdamOptimizer(NULL, alpha_t, beta1, beta2, weight_decay, epsilon);
  opt->alpha_t = alpha_t;
  f.task.args = opt;
  f.task.arglen = sizeof(AdamOptimizer);
  AdamOptimizer::ps_update_task(&f.task, f.regions, NULL, &f.runtime);
  cudaDeviceSynchronize(); /* the task reads *opt only on the host, but keep the frame alive anyway */
  delete opt;
}

}  /* extern "C" */
