// embedding_group_ffemb.cu — the Legion side of the FAST path: one leaf task per GPU for ALL the
// embedding tables placed on it, and the launcher-side edits of the fused mode, written out as
// code that compiles against the reference's own headers (cited as file:line under
// /root/reference) and the Legion stand-in of oracle/shim.
//
// Why a group: the reference launches one index task per Embedding op and direction
// (src/ops/embedding.cu:274-306,356-388), i.e. 26 forward + 26 backward leaf tasks per step for
// the Criteo shapes, each a kernel launch of its own (plus, in this library, a sort).  The numbers
// bench.py reports come from the table-batched entry points — ffemb_forward_batched,
// ffemb_backward_prepare, ffemb_backward_sgd_prepared: one launch sequence for all tables of a GPU,
// the index sort enqueued next to the forward.  The table-wise strategy gives every embedding a
// [1,1] task grid pinned to GPU i % ngpu (examples/cpp/DLRM/strategies/dlrm_strategy.cc:252-256), so
// "the tables of one GPU" is a well-defined set of ops whose partitions all have the single colour
// (0,0): one index launch over that one point can list the regions of all of them.
//
// Leaf tasks (registered like the reference's, src/runtime/model.cc:2510-2531, under the custom GPU
// task ids include/model.h keeps free, CUSTOM_GPU_TASK_ID_1..):
//   ffemb_embedding_group_forward_task       regions[3t] input_t RO, [3t+1] output_t WO, [3t+2] table_t RO
//   ffemb_embedding_group_prepare_task       regions[2t] input_t RO, [2t+1] table_t RO   (index sort only)
//   ffemb_embedding_group_backward_sgd_task  regions[3t] input_t RO, [3t+1] output-grad_t RO, [3t+2] table_t RW
// task->args is an FFEmbGroupArgs; task->local_args the group's OpMeta* (an EmbeddingMeta: aggr +
// FFHandler with the per-GPU workSpace), exactly as the per-table tasks get theirs.
// The multi-GPU forms (pooled rows stored into the peers' Concat inputs, output-grads pushed back,
// epoch flags instead of Legion dependences for the peer stores) take the peer pointers from an
// FFEmbGroupExchange kept next to the meta — see ffemb_embedding_group_forward_exchange_task.
//
// Launcher side (what replaces Embedding::create_weights / forward() / backward() for the ops of a
// group; Legion itself is absent from the reference checkout: the three group launchers are run on the CPU
// against the stand-in's recording Runtime, and create_weights against a recording FFModel next to the
// reference's own (tests/test_dropin_launchers.py) — the leaf bodies above are what the GPU tests drive):
//   ffemb_fused_create_weights   src/ops/embedding.cu:58-74   create_grad = false, ParameterSyncType::NONE
//   ffemb_group_forward          src/ops/embedding.cu:274-306 one launch for the group, no MAP_TO_ZC_MEMORY
//   ffemb_group_prepare          (new) the index sort, launched right after the forward
//   ffemb_group_backward_sgd     src/ops/embedding.cu:356-388 region 2 = the TABLE (RW), lr in the arguments
#include "model.h"
#include "cuda_helper.h"

#include "ffemb.h"

#include <sstream>
#include <vector>

struct FFEmbGroupArgs {
  int num_tables;
  float lr;       // SGDOptimizer::lr (include/optimizer.h:55); momentum == weight_decay == 0
  int prepared;   // backward: ffemb_embedding_group_prepare_task already ran on this GPU's workspace
  int parity;     // exchange: which of the two Concat-input buffers this step writes (x is double-buffered)
};

// Peer view of one GPU for the fused exchange, made once at init time (the Concat inputs and the
// output-grad buffers are allocated peer-mapped; inside one Legion process — one process drives all
// GPUs of a node, src/runtime/model.cc:679-706 — that is cudaDeviceEnablePeerAccess on the
// framebuffer instances' pointers).
struct FFEmbGroupExchange {
  ffemb_exchange_t x[2];            // the peers' Concat inputs, by step parity
  ffemb_sync_t sync;                // the peers' flag pads
  int col[FFEMB_MAX_TABLES];        // column of table t's block inside a Concat input row
  float *peer_outgrad[FFEMB_MAX_TABLES];  // (push side) owner's output-grad buffer of part p, peer-mapped
};

namespace {

void die_unless_ok(int rc, const char *entry) {
  if (rc == FFEMB_OK) return;
  std::stringstream msg;
  msg << "libffemb failure in " << entry << " (code " << rc << "): " << ffemb_last_error();
  FatalError(msg.str());
}

cudaStream_t task_stream() {
  cudaStream_t s;
  checkCUDA(get_legion_stream(&s));  // src/runtime/cuda_helper.cu:22-30
  return s;
}

const EmbeddingMeta *meta_of(const Task *task) { return *((EmbeddingMeta **)task->local_args); }

const FFEmbGroupArgs &args_of(const Task *task) {
  assert(task->arglen >= sizeof(FFEmbGroupArgs));
  const FFEmbGroupArgs &a = *(const FFEmbGroupArgs *)task->args;
  assert(a.num_tables >= 1 && a.num_tables <= FFEMB_MAX_TABLES);
  return a;
}

// One table's descriptor from its three regions; the same shape agreement the reference asserts per
// table (src/ops/embedding.cu:251-256, 336-341).  `dim` / `batch` must agree across the group.
template <typename OutAcc>
void describe(ffemb_table_t &d, const TensorAccessorR<int64_t, 2> &in, const OutAcc &out, float *table,
              const Rect<2> &tab_rect, int64_t &dim, int64_t &batch, bool first) {
  const int64_t bag = in.rect.hi[0] - in.rect.lo[0] + 1;
  const int64_t D = out.rect.hi[0] - out.rect.lo[0] + 1, B = out.rect.hi[1] - out.rect.lo[1] + 1;
  assert(in.rect.lo[1] == out.rect.lo[1] && in.rect.hi[1] == out.rect.hi[1]);
  assert(tab_rect.hi[1] - tab_rect.lo[1] + 1 == D);
  if (first) { dim = D; batch = B; }
  assert(D == dim && B == batch);  // one launch: the group shares batch and width
  d.idx = in.ptr;
  d.out = const_cast<float *>((const float *)out.ptr);
  d.table = table;
  d.R = tab_rect.hi[0] - tab_rect.lo[0] + 1;
  d.L = (int32_t)bag;
  d.reserved = 0;
}

}  // namespace

// ---- leaf bodies ---------------------------------------------------------------------------------
void ffemb_embedding_group_forward_task(const Task *task, const std::vector<PhysicalRegion> &regions,
                                        Context ctx, Runtime *runtime) {
  const FFEmbGroupArgs &a = args_of(task);
  assert((int)regions.size() == 3 * a.num_tables && (int)task->regions.size() == 3 * a.num_tables);
  const EmbeddingMeta *m = meta_of(task);
  ffemb_table_t t[FFEMB_MAX_TABLES];
  int64_t dim = 0, batch = 0;
  for (int i = 0; i < a.num_tables; i++) {
    TensorAccessorR<int64_t, 2> in(regions[3 * i], task->regions[3 * i], FID_DATA, ctx, runtime);
    TensorAccessorW<float, 2> out(regions[3 * i + 1], task->regions[3 * i + 1], FID_DATA, ctx, runtime, false);
    TensorAccessorR<float, 2> tab(regions[3 * i + 2], task->regions[3 * i + 2], FID_DATA, ctx, runtime);
    describe(t[i], in, out, const_cast<float *>(tab.ptr), tab.rect, dim, batch, i == 0);
  }
  die_unless_ok(ffemb_forward_batched(t, a.num_tables, dim, batch, (int)m->aggr, task_stream()),
                "ffemb_forward_batched");
  if (m->profiling) checkCUDA(cudaDeviceSynchronize());
}

// The index sort of the backward depends on the indices only: launched right after the forward task
// of the same step (both read the input regions), it runs under the dense layers; the backward task
// then only does the segmented sum + row update.  regions[2t] input, regions[2t+1] table (row count).
void ffemb_embedding_group_prepare_task(const Task *task, const std::vector<PhysicalRegion> &regions,
                                        Context ctx, Runtime *runtime) {
  const FFEmbGroupArgs &a = args_of(task);
  assert((int)regions.size() == 2 * a.num_tables);
  const EmbeddingMeta *m = meta_of(task);
  ffemb_table_t t[FFEMB_MAX_TABLES];
  int64_t dim = 0, batch = 0;
  for (int i = 0; i < a.num_tables; i++) {
    TensorAccessorR<int64_t, 2> in(regions[2 * i], task->regions[2 * i], FID_DATA, ctx, runtime);
    TensorAccessorR<float, 2> tab(regions[2 * i + 1], task->regions[2 * i + 1], FID_DATA, ctx, runtime);
    const int64_t D = tab.rect.hi[1] - tab.rect.lo[1] + 1, B = in.rect.hi[1] - in.rect.lo[1] + 1;
    if (i == 0) { dim = D; batch = B; }
    assert(D == dim && B == batch);
    t[i].idx = in.ptr;
    t[i].out = NULL;
    t[i].table = const_cast<float *>(tab.ptr);
    t[i].R = tab.rect.hi[0] - tab.rect.lo[0] + 1;
    t[i].L = (int32_t)(in.rect.hi[0] - in.rect.lo[0] + 1);
    t[i].reserved = 0;
  }
  die_unless_ok(ffemb_backward_prepare(t, a.num_tables, dim, batch, (int)m->aggr, m->handle.workSpace,
                                       m->handle.workSpaceSize, task_stream()),
                "ffemb_backward_prepare");
}

// Replaces, for every table of the group: Op::zero_grad on the weight-grad (src/runtime/model.cc:
// 466-490), Embedding::backward_task (src/ops/embedding.cu:322-354) and SGDOptimizer::update ->
// sgd_update (src/runtime/optimizer_kernel.cu:23-41).  regions[3t+2] is the TABLE (RW).
void ffemb_embedding_group_backward_sgd_task(const Task *task, const std::vector<PhysicalRegion> &regions,
                                             Context ctx, Runtime *runtime) {
  const FFEmbGroupArgs &a = args_of(task);
  assert((int)regions.size() == 3 * a.num_tables && (int)task->regions.size() == 3 * a.num_tables);
  const EmbeddingMeta *m = meta_of(task);
  ffemb_table_t t[FFEMB_MAX_TABLES];
  int64_t dim = 0, batch = 0;
  for (int i = 0; i < a.num_tables; i++) {
    TensorAccessorR<int64_t, 2> in(regions[3 * i], task->regions[3 * i], FID_DATA, ctx, runtime);
    TensorAccessorR<float, 2> og(regions[3 * i + 1], task->regions[3 * i + 1], FID_DATA, ctx, runtime);
    TensorAccessorW<float, 2> tab(regions[3 * i + 2], task->regions[3 * i + 2], FID_DATA, ctx, runtime, true);
    describe(t[i], in, og, tab.ptr, tab.rect, dim, batch, i == 0);
    t[i].out = const_cast<float *>(og.ptr);
  }
  int rc = a.prepared
               ? ffemb_backward_sgd_prepared(t, a.num_tables, dim, batch, (int)m->aggr, a.lr, m->handle.workSpace,
                                             m->handle.workSpaceSize, task_stream())
               : ffemb_backward_sgd_batched(t, a.num_tables, dim, batch, (int)m->aggr, a.lr, m->handle.workSpace,
                                            m->handle.workSpaceSize, task_stream());
  die_unless_ok(rc, a.prepared ? "ffemb_backward_sgd_prepared" : "ffemb_backward_sgd_batched");
  if (m->profiling) checkCUDA(cudaDeviceSynchronize());
}

// Multi-GPU forward: regions[2t] input_t RO, regions[2t+1] table_t RO; the pooled rows do not go to an
// output region of this task but straight into the Concat input of the GPU that owns each sample
// (ffemb_forward_exchange), then the group's flag is raised (ffemb_sync_signal): the consumer on
// every GPU — Concat's task, or the push task below — waits on it (ffemb_sync_wait) where the
// reference relies on Legion copying the zero-copy output regions (src/ops/embedding.cu:295-298,
// src/ops/concat.cu:88-97).  `ex` travels in the task arguments behind FFEmbGroupArgs.
void ffemb_embedding_group_forward_exchange_task(const Task *task, const std::vector<PhysicalRegion> &regions,
                                                 Context ctx, Runtime *runtime) {
  const FFEmbGroupArgs &a = args_of(task);
  assert(task->arglen == sizeof(FFEmbGroupArgs) + sizeof(FFEmbGroupExchange *));
  const FFEmbGroupExchange *ex = *(FFEmbGroupExchange *const *)((const char *)task->args + sizeof(FFEmbGroupArgs));
  assert((int)regions.size() == 2 * a.num_tables);
  const EmbeddingMeta *m = meta_of(task);
  ffemb_table_t t[FFEMB_MAX_TABLES];
  int64_t dim = 0, batch = 0;
  for (int i = 0; i < a.num_tables; i++) {
    TensorAccessorR<int64_t, 2> in(regions[2 * i], task->regions[2 * i], FID_DATA, ctx, runtime);
    TensorAccessorR<float, 2> tab(regions[2 * i + 1], task->regions[2 * i + 1], FID_DATA, ctx, runtime);
    const int64_t D = tab.rect.hi[1] - tab.rect.lo[1] + 1, B = in.rect.hi[1] - in.rect.lo[1] + 1;
    if (i == 0) { dim = D; batch = B; }
    assert(D == dim && B == batch);
    t[i].idx = in.ptr;
    t[i].out = NULL;
    t[i].table = const_cast<float *>(tab.ptr);
    t[i].R = tab.rect.hi[0] - tab.rect.lo[0] + 1;
    t[i].L = (int32_t)(in.rect.hi[0] - in.rect.lo[0] + 1);
    t[i].reserved = ex->col[i];
  }
  cudaStream_t s = task_stream();
  die_unless_ok(ffemb_forward_exchange(t, a.num_tables, dim, batch, (int)m->aggr, &ex->x[a.parity & 1], s),
                "ffemb_forward_exchange");
  die_unless_ok(ffemb_sync_signal(&ex->sync, 0, s), "ffemb_sync_signal");
}

// ---- launcher side (see the header comment) --------------------------------------------------------
// The ops of one GPU.  `meta` is the OpMeta* the FIRST op's init_task made on that GPU (Op::meta[0] of
// a [1,1] grid): its FFHandler carries the GPU's workSpace.
struct FFEmbGroup {
  std::vector<Embedding *> ops;
};

// Embedding::create_weights in fused mode: no weight-grad region, and the table stays out of
// FFModel::parameters (src/runtime/model.cc:1643-1645), so neither SGDOptimizer::update nor the
// gradient all-reduce (src/runtime/optimizer_kernel.cu:170-171) nor Op::zero_grad touches it.
void ffemb_fused_create_weights(Embedding *op, FFModel &model) {
  std::string pcname = op->name;
  op->task_is = IndexSpaceT<2>(model.get_or_create_task_is(2, pcname));
  const int dims[2] = {op->out_channels, op->num_entries};
  op->weights[0] = model.create_linear_weight<2, 2>(op, dims, DT_FLOAT, op->kernel_initializer,
                                                    false /*create_grad*/, ParameterSyncType::NONE);
  assert(op->numWeights == 1);
}

namespace {
IndexLauncher group_launcher(const FFModel &ff, const FFEmbGroup &g, TaskID id, const FFEmbGroupArgs &args,
                             ArgumentMap &argmap, OpMeta **mp) {
  Context ctx = ff.config.lg_ctx;
  Runtime *runtime = ff.config.lg_hlr;
  Embedding *lead = g.ops[0];
  Rect<2> rect = runtime->get_index_space_domain(ctx, lead->task_is);
  assert(rect.lo[0] == rect.hi[0] && rect.lo[1] == rect.hi[1]);  // a [1,1] grid: the table-wise strategy
  *mp = lead->meta[0];
  for (PointInRectIterator<2> it(rect); it(); it++) argmap.set_point(*it, TaskArgument(mp, sizeof(OpMeta *)));
  // the mapping tag of the lead op: FFMapper::slice_task places the point on that op's device
  // (src/mapper/mapper.cc:346-438), which is the device of every op of the group
  return IndexLauncher(id, lead->task_is, TaskArgument(&args, sizeof(args)), argmap, Predicate::TRUE_PRED,
                       false /*must*/, 0 /*mapper_id*/, FFConfig::get_hash_id(std::string(lead->name)));
}
}  // namespace

void ffemb_group_forward(const FFModel &ff, const FFEmbGroup &g) {
  FFEmbGroupArgs args = {(int)g.ops.size(), 0.f, 0, 0};
  ArgumentMap argmap;
  OpMeta *mp = NULL;
  IndexLauncher launcher = group_launcher(ff, g, CUSTOM_GPU_TASK_ID_1, args, argmap, &mp);
  int r = 0;
  for (Embedding *op : g.ops) {
    launcher.add_region_requirement(
        RegionRequirement(op->input_lps[0], 0 /*projection*/, READ_ONLY, EXCLUSIVE, op->inputs[0].region));
    launcher.add_field(r++, FID_DATA);
    // (no MAP_TO_ZC_MEMORY: the pooled rows stay in framebuffer memory; the exchange form does not
    // list an output region at all)
    launcher.add_region_requirement(
        RegionRequirement(op->outputs[0].part, 0 /*projection*/, WRITE_ONLY, EXCLUSIVE, op->outputs[0].region));
    launcher.add_field(r++, FID_DATA);
    launcher.add_region_requirement(
        RegionRequirement(op->weights[0].part, 0 /*projection*/, READ_ONLY, EXCLUSIVE, op->weights[0].region));
    launcher.add_field(r++, FID_DATA);
  }
  ff.config.lg_hlr->execute_index_space(ff.config.lg_ctx, launcher);
}

void ffemb_group_prepare(const FFModel &ff, const FFEmbGroup &g) {
  FFEmbGroupArgs args = {(int)g.ops.size(), 0.f, 0, 0};
  ArgumentMap argmap;
  OpMeta *mp = NULL;
  IndexLauncher launcher = group_launcher(ff, g, CUSTOM_GPU_TASK_ID_2, args, argmap, &mp);
  int r = 0;
  for (Embedding *op : g.ops) {
    launcher.add_region_requirement(
        RegionRequirement(op->input_lps[0], 0 /*projection*/, READ_ONLY, EXCLUSIVE, op->inputs[0].region));
    launcher.add_field(r++, FID_DATA);
    launcher.add_region_requirement(
        RegionRequirement(op->weights[0].part, 0 /*projection*/, READ_ONLY, EXCLUSIVE, op->weights[0].region));
    launcher.add_field(r++, FID_DATA);
  }
  ff.config.lg_hlr->execute_index_space(ff.config.lg_ctx, launcher);
}

// Embedding::backward() of every op of the group in one launch.  Region 2 of each table is the table
// itself, READ_WRITE (the reference lists weights[0].part_grad, src/ops/embedding.cu:382-386); the
// learning rate travels in the task arguments.
void ffemb_group_backward_sgd(const FFModel &ff, const FFEmbGroup &g, bool prepared) {
  const SGDOptimizer *opt = (const SGDOptimizer *)ff.optimizer;
  assert(opt->momentum == 0.0 && opt->weight_decay == 0.0);  // else: ffemb_backward_sgdx_batched
  FFEmbGroupArgs args = {(int)g.ops.size(), (float)opt->lr, prepared ? 1 : 0, 0};
  ArgumentMap argmap;
  OpMeta *mp = NULL;
  IndexLauncher launcher = group_launcher(ff, g, CUSTOM_GPU_TASK_ID_3, args, argmap, &mp);
  int r = 0;
  for (Embedding *op : g.ops) {
    launcher.add_region_requirement(
        RegionRequirement(op->input_lps[0], 0 /*projection*/, READ_ONLY, EXCLUSIVE, op->inputs[0].region));
    launcher.add_field(r++, FID_DATA);
    launcher.add_region_requirement(RegionRequirement(op->outputs[0].part_grad, 0 /*projection*/, READ_ONLY,
                                                      EXCLUSIVE, op->outputs[0].region_grad));
    launcher.add_field(r++, FID_DATA);
    launcher.add_region_requirement(
        RegionRequirement(op->weights[0].part, 0 /*projection*/, READ_WRITE, EXCLUSIVE, op->weights[0].region));
    launcher.add_field(r++, FID_DATA);
  }
  ff.config.lg_hlr->execute_index_space(ff.config.lg_ctx, launcher);
}

// Registration (what src/runtime/model.cc:2510-2531 gains; Legion's preregister_task_variant is not in
// the stand-in, so this is the text a maintainer adds):
//   { TaskVariantRegistrar r(CUSTOM_GPU_TASK_ID_1, "Embedding group forward");
//     r.add_constraint(ProcessorConstraint(Processor::TOC_PROC)); r.set_leaf();
//     Runtime::preregister_task_variant<ffemb_embedding_group_forward_task>(r, "Embedding group forward"); }
//   ... CUSTOM_GPU_TASK_ID_2 -> ffemb_embedding_group_prepare_task,
//       CUSTOM_GPU_TASK_ID_3 -> ffemb_embedding_group_backward_sgd_task,
//       CUSTOM_GPU_TASK_ID_4 -> ffemb_embedding_group_forward_exchange_task.
