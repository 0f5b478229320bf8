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
#include <cstring>
#include <map>
#include <vector>
#include <cstdlib>

#define FFREF_DEAD(what)                                                        \
  do {                                                                          \
    fprintf(stderr, "oracle/ref_glue_gpu: control-plane stub called: %s\n", what); \
    abort();                                                                    \
  } while (0)

/* ---- link-time leftovers (never executed on this path) ---------------------------------- */
OpMeta::OpMeta(FFHandler _handle) : handle(_handle), profiling(false) {}
Tensor::Tensor() { numDim = 0; owner_op = NULL; owner_idx = 0; }
/* (plain member-wise assignment, as src/runtime/model.cc:46-61 does it: needed by the weight-creation
 * record below, where `weights[0] = model.create_linear_weight(...)` assigns a Parameter) */
Tensor &Tensor::operator=(const Tensor &o) {
  if (this == &o) return *this;
  numDim = o.numDim;
  memcpy(adim, o.adim, sizeof(adim));
  data_type = o.data_type; sync_type = o.sync_type;
  owner_op = o.owner_op; owner_idx = o.owner_idx;
  region = o.region; region_grad = o.region_grad;
  part = o.part; part_grad = o.part_grad;
  physical_region = o.physical_region;
  return *this;
}
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
/* The two FFModel entry points weight creation goes through abort like the rest of the control plane —
 * unless a test has installed a record (ff*_create_weights_record below): then they write their
 * arguments down and hand back a [1,1] task space / a bare 2-D parameter. */
namespace {
struct WeightsRecord {
  int task_is_calls, task_is_dims;
  char task_is_name[MAX_OPNAME];
  int weight_calls, dims[2], data_type, create_grad, sync_type;
  const void *op, *initializer;
};
WeightsRecord *g_wrec = NULL;
}  // namespace
IndexSpace FFModel::get_or_create_task_is(int n, const std::string &name) {
  if (!g_wrec) FFREF_DEAD("FFModel");
  g_wrec->task_is_calls++;
  g_wrec->task_is_dims = n;
  snprintf(g_wrec->task_is_name, sizeof(g_wrec->task_is_name), "%s", name.c_str());
  return Legion::IndexSpaceT<2>(Legion::Rect<2>(Legion::Point<2>(0, 0), Legion::Point<2>(0, 0)));
}
template <>
Tensor FFModel::create_tensor<2>(const int *, DataType, const Op *, bool) { FFREF_DEAD("FFModel"); }
template <>
Parameter FFModel::create_linear_weight<2, 2>(Op *op, const int *dims, DataType dt, Initializer *init,
                                              bool create_grad, ParameterSyncType sync) {
  if (!g_wrec) FFREF_DEAD("FFModel");
  g_wrec->weight_calls++;
  g_wrec->dims[0] = dims[0]; g_wrec->dims[1] = dims[1];
  g_wrec->data_type = (int)dt; g_wrec->create_grad = create_grad; g_wrec->sync_type = (int)sync;
  g_wrec->op = op; g_wrec->initializer = init;
  Parameter p;
  p.numDim = 2;
  p.adim[0] = dims[1]; p.adim[1] = dims[0];  /* [num_entries][out_channels], like Embedding's ctor */
  p.data_type = dt; p.sync_type = sync; p.owner_op = op;
  return p;
}
template <>
void FFModel::create_disjoint_partition<2>(const Tensor &, const IndexSpaceT<2> &,
                                           LogicalPartition &, LogicalPartition &) {
  FFREF_DEAD("FFModel");
}
/* (the mapping tag of an op: the hash of its name, as src/runtime/strategy.cc:21-24 forms it) */
size_t FFConfig::get_hash_id(const std::string &pcname) { return std::hash<std::string>()(pcname); }
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
/* ... and the launcher side of the same file (what replaces Embedding::forward() / backward()) */
struct FFEmbGroup { std::vector<Embedding *> ops; };
void ffemb_fused_create_weights(Embedding *op, FFModel &model);
void ffemb_group_forward(const FFModel &ff, const FFEmbGroup &g);
void ffemb_group_prepare(const FFModel &ff, const FFEmbGroup &g);
void ffemb_group_backward_sgd(const FFModel &ff, const FFEmbGroup &g, bool prepared);
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

/* The LAUNCHER side, driven without Legion and without a GPU.  libffdropin_gpu.so: the per-GPU group
 * launchers ffemb_group_forward / _prepare / _backward_sgd over `n` tables (what = 0 / 1 / 2);
 * libffref_gpu.so: the reference's own Embedding::forward / backward (src/ops/embedding.cu:274-306,
 * 356-388; what = 0 / 2, n = 1) from its object code.  The ops are laid out the way FFModel::compile
 * leaves them for the table-wise strategy (a [1,1] task grid, one logical region per input / output /
 * output-grad / table / table-grad, the op's meta), the launcher runs against a Runtime whose launch hook
 * writes the launch down, and the hook also turns the region requirements into the (pointer, rectangle)
 * frame a leaf task would be handed.  `rec` gets
 *   [0] task id  [1] #requirements  [2] mapping tag  [3] mapper id  [4] must-parallelism  [5] points
 *   [6] arglen  [7] args.num_tables  [8] bits of args.lr  [9] args.prepared  [10] args.parity
 *       ([7..10] = 0 for the reference, whose tasks take no by-value argument)
 *   [11] #per-point arguments  [12] 1 if the point's argument is the lead op's OpMeta*  [13] its length
 *   [14..16] READ_ONLY / WRITE_ONLY / READ_WRITE  [17] EXCLUSIVE  [18] FID_DATA  [19] MAP_TO_ZC_MEMORY
 *   [20..22] CUSTOM_GPU_TASK_ID_1..3  [23] EMBED_FWD_TASK_ID  [24] EMBED_BWD_TASK_ID  [25..31] 0
 *   then per requirement: privilege, coherence, tag, #fields, field, base pointer, extent 0, extent 1
 * — the last three being what the hand-made frames of ffref_gpu_*_task / ffdropin_gpu_group_task give the
 * leaf bodies in the GPU tests.  Returns the number of entries written, < 0 on error.  Nothing is
 * launched and no device memory is touched: the pointers are only carried.                        */
namespace {
enum { kRecHead = 32 };
struct LaunchRecord {
  int64_t *rec;
  int max_rec, n;
  const std::map<unsigned, void *> *base_of;
  OpMeta *lead_meta;
};
void record_index_launch(void *user, const Legion::IndexLauncher &l) {
  LaunchRecord &R = *(LaunchRecord *)user;
  const int need = kRecHead + 8 * (int)l.region_requirements.size();
  if (need > R.max_rec) { R.n = -2; return; }
  int64_t *r = R.rec;
  for (int i = 0; i < kRecHead; i++) r[i] = 0;
  r[0] = l.task_id; r[1] = (int64_t)l.region_requirements.size(); r[2] = (int64_t)l.tag; r[3] = l.map_id;
  r[4] = l.must_parallelism; r[5] = (int64_t)l.launch_domain.get_volume();
  r[6] = (int64_t)l.global_arg.len;
#ifdef FFEMB_DROPIN
  if (l.global_arg.len != sizeof(FFEmbGroupArgs)) { R.n = -5; return; }
  const FFEmbGroupArgs &a = *(const FFEmbGroupArgs *)l.global_arg.ptr;
  int32_t lr_bits;
  memcpy(&lr_bits, &a.lr, 4);
  r[7] = a.num_tables; r[8] = lr_bits; r[9] = a.prepared; r[10] = a.parity;
#endif
  r[11] = (int64_t)l.argument_map.size();
  if (!l.argument_map.bytes.empty()) {
    const std::vector<char> &pa = l.argument_map.bytes.begin()->second;
    r[13] = (int64_t)pa.size();
    OpMeta *given = NULL;
    if (pa.size() == sizeof(OpMeta *)) memcpy(&given, pa.data(), sizeof(given));
    r[12] = given == R.lead_meta;
  }
  r[14] = Legion::READ_ONLY; r[15] = Legion::WRITE_ONLY; r[16] = Legion::READ_WRITE; r[17] = Legion::EXCLUSIVE;
  r[18] = FID_DATA; r[19] = (int64_t)MAP_TO_ZC_MEMORY;
  r[20] = CUSTOM_GPU_TASK_ID_1; r[21] = CUSTOM_GPU_TASK_ID_2; r[22] = CUSTOM_GPU_TASK_ID_3;
  r[23] = EMBED_FWD_TASK_ID; r[24] = EMBED_BWD_TASK_ID;
  /* the frame: one PhysicalRegion per requirement — the single colour of a [1,1] partition is the whole
   * parent region — with the pointer the region was registered with */
  std::vector<Legion::PhysicalRegion> regions;
  for (size_t i = 0; i < l.region_requirements.size(); i++) {
    const Legion::RegionRequirement &q = l.region_requirements[i];
    std::map<unsigned, void *>::const_iterator it = R.base_of->find(q.parent.id);
    if (it == R.base_of->end()) { R.n = -3; return; }
    regions.push_back(Legion::PhysicalRegion(it->second, q.parent.get_index_space().dom));
    const Legion::Domain &d = regions.back().dom;
    int64_t *e = r + kRecHead + 8 * i;
    e[0] = q.privilege; e[1] = q.prop; e[2] = (int64_t)q.tag; e[3] = (int64_t)q.privilege_fields.size();
    e[4] = q.privilege_fields.empty() ? -1 : (int64_t)*q.privilege_fields.begin();
    e[5] = (int64_t)(uintptr_t)regions.back().base;
    e[6] = d.dim == 2 ? d.rect_data[2] - d.rect_data[0] + 1 : -1;
    e[7] = d.dim == 2 ? d.rect_data[3] - d.rect_data[1] + 1 : -1;
  }
  R.n = need;
}
}  // namespace

int ffref_launch_record(int what, int n, void *const *idx, void *const *out, void *const *W, void *const *Wgrad,
                        const int64_t *L, int64_t D, int64_t B, const int64_t *Rows, int aggr, float lr,
                        int prepared, int64_t *rec, int max_rec) {
  if (n < 1 || what < 0 || what > 2) return -1;
#ifndef FFEMB_DROPIN
  if (n != 1 || what == 1) return -1;  /* the reference launches one op at a time and has no prepare task */
#endif
  Legion::Runtime runtime;
  Legion::ContextStub ctx;
  /* FFModel and Embedding cannot be constructed here (their constructors belong to the control plane,
   * which is stubbed out above): zero-filled storage with the fields the launchers read filled in */
  std::vector<char> ff_store(sizeof(FFModel) + alignof(FFModel), 0);
  FFModel &ff = *(FFModel *)(((uintptr_t)ff_store.data() + alignof(FFModel) - 1) / alignof(FFModel) * alignof(FFModel));
  ff.config.lg_ctx = &ctx;
  ff.config.lg_hlr = &runtime;
  SGDOptimizer opt(&ff, (double)lr, 0.0, false, 0.0);
  ff.optimizer = &opt;
  std::map<unsigned, void *> base_of;
  unsigned next_id = 100;
  std::vector<std::vector<char> > op_store(n);
  std::vector<Embedding *> ops;
  EmbeddingMeta *meta = make_meta(aggr);
  const Legion::Rect<2> one(Legion::Point<2>(0, 0), Legion::Point<2>(0, 0));
  for (int t = 0; t < n; t++) {
    op_store[t].assign(sizeof(Embedding) + alignof(Embedding), 0);
    Embedding *op = (Embedding *)(((uintptr_t)op_store[t].data() + alignof(Embedding) - 1) / alignof(Embedding) * alignof(Embedding));
    snprintf(op->name, sizeof(op->name), "embedding%d", t);
    op->task_is = Legion::IndexSpaceT<2>(one);
    op->meta[0] = meta;
    op->numWeights = 1;
    op->num_entries = (int)Rows[t];
    op->out_channels = (int)D;
    struct Reg { Legion::LogicalRegion *lr; void *base; int64_t d0, d1; };
    Legion::LogicalPartition part;
    part.ip.color_space = Legion::Domain(one);
    Reg regs[5] = {{&op->inputs[0].region, idx[t], L[t], B},
                   {&op->outputs[0].region, out ? out[t] : NULL, D, B},
                   {&op->outputs[0].region_grad, out ? out[t] : NULL, D, B},
                   {&op->weights[0].region, W[t], Rows[t], D},
                   {&op->weights[0].region_grad, Wgrad ? Wgrad[t] : NULL, Rows[t], D}};
    for (int k = 0; k < 5; k++) {
      Legion::Rect<2> r(Legion::Point<2>(0, 0), Legion::Point<2>(regs[k].d0 - 1, regs[k].d1 - 1));
      *regs[k].lr = Legion::LogicalRegion(Legion::IndexSpace(Legion::Domain(r)));
      regs[k].lr->id = next_id;
      base_of[next_id++] = regs[k].base;
    }
    op->input_lps[0] = part;
    op->outputs[0].part = part;
    op->outputs[0].part_grad = part;
    op->weights[0].part = part;
    op->weights[0].part_grad = part;
    ops.push_back(op);
  }
  LaunchRecord R = {rec, max_rec, -4, &base_of, meta};
  runtime.index_launch_hook = record_index_launch;
  runtime.index_launch_user = &R;
#ifdef FFEMB_DROPIN
  FFEmbGroup group;
  group.ops = ops;
  if (what == 0) ffemb_group_forward(ff, group);
  else if (what == 1) ffemb_group_prepare(ff, group);
  else ffemb_group_backward_sgd(ff, group, prepared != 0);
#else
  (void)prepared;
  if (what == 0) ops[0]->Embedding::forward(ff);
  else ops[0]->Embedding::backward(ff);
#endif
  delete meta;
  return R.n;
}

/* Weight creation, written down: libffref_gpu.so runs the reference's own Embedding::create_weights
 * (src/ops/embedding.cu:58-74), libffdropin_gpu.so the fused mode's ffemb_fused_create_weights, both on
 * an op with `rows` entries of `dim` floats named `name`; FFModel's two entry points record what they
 * were asked for.  rec: [0] get_or_create_task_is calls [1] its dimension argument
 *   [2] create_linear_weight calls [3] dims[0] [4] dims[1] [5] data type [6] create_grad [7] sync type
 *   [8] 1 if the op passed is this op [9] 1 if the initializer passed is the op's kernel_initializer
 *   [10] DT_FLOAT [11..13] ParameterSyncType NONE / PS / NCCL [14] points of op->task_is afterwards
 *   [15] weights[0].numDim afterwards.  name_out (MAX_OPNAME bytes): the name the task space was asked by. */
int ffref_create_weights_record(const char *name, int rows, int dim, int64_t *rec, char *name_out) {
  std::vector<char> ff_store(sizeof(FFModel) + alignof(FFModel), 0), op_store(sizeof(Embedding) + alignof(Embedding), 0);
  FFModel &ff = *(FFModel *)(((uintptr_t)ff_store.data() + alignof(FFModel) - 1) / alignof(FFModel) * alignof(FFModel));
  Embedding *op = (Embedding *)(((uintptr_t)op_store.data() + alignof(Embedding) - 1) / alignof(Embedding) * alignof(Embedding));
  snprintf(op->name, sizeof(op->name), "%s", name);
  op->num_entries = rows;
  op->out_channels = dim;
  op->numWeights = 1;
  op->kernel_initializer = (Initializer *)(uintptr_t)0x1234;  /* only carried */
  WeightsRecord w;
  memset(&w, 0, sizeof(w));
  g_wrec = &w;
#ifdef FFEMB_DROPIN
  ffemb_fused_create_weights(op, ff);
#else
  op->Embedding::create_weights(ff);
#endif
  g_wrec = NULL;
  rec[0] = w.task_is_calls; rec[1] = w.task_is_dims; rec[2] = w.weight_calls; rec[3] = w.dims[0]; rec[4] = w.dims[1];
  rec[5] = w.data_type; rec[6] = w.create_grad; rec[7] = w.sync_type; rec[8] = w.op == (const void *)(Op *)op;
  rec[9] = w.initializer == (const void *)op->kernel_initializer;
  rec[10] = DT_FLOAT; rec[11] = (int)ParameterSyncType::NONE; rec[12] = (int)ParameterSyncType::PS;
  rec[13] = (int)ParameterSyncType::NCCL;
  rec[14] = (int64_t)op->task_is.dom.get_volume(); rec[15] = op->weights[0].numDim;
  snprintf(name_out, MAX_OPNAME, "%s", w.task_is_name);
  return 16;
}

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
