// embedding_ffemb.cu — the reference-side binding of libffemb.so, written out.
//
// This translation unit is what a DLRM-FlexFlow maintainer adds to the reference tree
// (cited as file:line under /root/reference).  It REPLACES these definitions of
// src/ops/embedding.cu and nothing else:
//     embed_forward / embed_backward                  :166-217   (the two naive kernels)
//     Embedding::forward_kernel / backward_kernel     :219-231, :308-320   (the raw-pointer seam)
//     Embedding::forward_task  / backward_task        :238-272, :322-354   (the Legion leaf bodies)
// Everything Legion-facing that stays in the reference file keeps working unchanged: the
// constructor, create_weights, create_output_and_partition, init / init_task, the index
// launchers forward() / backward() and measure_operator_cost() (which calls the two static
// wrappers below).  INTEGRATION.md has the three-line patch and the optional fused mode.
//
// It is compiled here against the reference's own headers (include/model.h et al.) and the
// Legion stand-in under oracle/shim (Legion itself is absent from the reference checkout), and
// linked with the reference's own accessor/helper objects into oracle/_ref/libffdropin_gpu.so so
// that the parity tests can drive the SAME (Task, PhysicalRegion) frames through the reference's
// task bodies and through these.
//
// Error convention: libffemb returns codes and never exits; the reference prints and exit(1)s
// (include/cuda_helper.h:6-46).  die_unless_ok() turns one into the other.
#include "model.h"
#include "cuda_helper.h"

#include "ffemb.h"

namespace {

void die_unless_ok(int rc, const char *entry) {
  if (rc == FFEMB_OK) return;
  std::stringstream msg;
  msg << "libffemb failure in " << entry << " (code " << rc << "): " << ffemb_last_error();
  FatalError(msg.str());
}

// Shape of an embedding leaf task, derived from the rectangles of its three regions (dim 0 is the
// fastest-varying one in FlexFlow's tensors): input [bag][batch], output [dim][batch],
// table [rows][dim] as Legion declares it (flat buffer = row-major [rows][dim]).
struct Shape {
  int64_t bag, dim, batch, rows;
};
template <typename InAcc, typename OutAcc, typename TabAcc>
Shape shape_of(const InAcc &in, const OutAcc &out, const TabAcc &tab) {
  Shape s;
  s.bag = in.rect.hi[0] - in.rect.lo[0] + 1;
  s.dim = out.rect.hi[0] - out.rect.lo[0] + 1;
  s.batch = out.rect.hi[1] - out.rect.lo[1] + 1;
  s.rows = tab.rect.hi[0] - tab.rect.lo[0] + 1;
  // the same agreement the reference asserts (src/ops/embedding.cu:251-256, 336-341):
  // samples of input and output coincide; the table is as wide as the output
  assert(in.rect.lo[1] == out.rect.lo[1] && in.rect.hi[1] == out.rect.hi[1]);
  assert(tab.rect.hi[1] - tab.rect.lo[1] + 1 == s.dim);
  return s;
}

cudaStream_t task_stream() {
  cudaStream_t s;
  checkCUDA(get_legion_stream(&s));  // src/runtime/cuda_helper.cu:22-30
  return s;
}

const EmbeddingMeta *meta_of(const Task *task) { return *((EmbeddingMeta **)task->local_args); }

}  // namespace

// ---- the seam: same signatures as include/model.h:1169-1186 -------------------------------------
void Embedding::forward_kernel(int64_t const *input_ptr, float *output_ptr, float const *weight_ptr,
                               int in_dim, int out_dim, int batch_size, AggrMode aggr,
                               int outputSize, cudaStream_t stream) {
  assert((int64_t)outputSize == (int64_t)batch_size * out_dim);
  die_unless_ok(ffemb_forward(input_ptr, output_ptr, weight_ptr, in_dim, out_dim, batch_size,
                              (int)aggr, stream),
                "ffemb_forward");
}

// API-exact backward: accumulate into the dense, already zeroed weight-grad.  This wrapper has no
// OpMeta (the simulator calls it with bare pointers, src/ops/embedding.cu:417-431), hence no
// workspace: it is the unordered red.global.add variant, which is what embed_backward is.
void Embedding::backward_kernel(int64_t const *input_ptr, float const *output_ptr,
                                float *weight_grad_ptr, int in_dim, int out_dim, int batch_size,
                                AggrMode aggr, int outputSize, cudaStream_t stream) {
  assert((int64_t)outputSize == (int64_t)batch_size * out_dim);
  die_unless_ok(ffemb_backward_dense_atomic(input_ptr, output_ptr, weight_grad_ptr, in_dim, out_dim,
                                            batch_size, (int)aggr, stream),
                "ffemb_backward_dense_atomic");
}

// ---- the Legion leaf bodies --------------------------------------------------------------------
// regions[0] input (RO), regions[1] output (WO), regions[2] table (RO)
void Embedding::forward_task(const Task *task, const std::vector<PhysicalRegion> &regions,
                             Context ctx, Runtime *runtime) {
  assert(regions.size() == 3 && task->regions.size() == 3);
  const EmbeddingMeta *m = meta_of(task);
  TensorAccessorR<int64_t, 2> in(regions[0], task->regions[0], FID_DATA, ctx, runtime);
  TensorAccessorW<float, 2> out(regions[1], task->regions[1], FID_DATA, ctx, runtime, false);
  TensorAccessorR<float, 2> tab(regions[2], task->regions[2], FID_DATA, ctx, runtime);
  const Shape f = shape_of(in, out, tab);
  // the task knows the row count from the table's rectangle, the raw-pointer wrapper above does not:
  // with it the library can pick the 64-bit row path for tables of more than 2^32 rows
  ffemb_table_t one;
  one.idx = in.ptr;
  one.out = out.ptr;
  one.table = const_cast<float *>(tab.ptr);
  one.R = f.rows;
  one.L = (int32_t)f.bag;
  one.reserved = 0;
  die_unless_ok(ffemb_forward_batched(&one, 1, f.dim, f.batch, (int)m->aggr, task_stream()),
                "ffemb_forward_batched");
  if (m->profiling) checkCUDA(cudaDeviceSynchronize());
}

// regions[0] input (RO), regions[1] output-grad (RO), regions[2] weight-grad (RW, arrives zeroed).
// Deterministic when the handler's workspace (FFHandler::workSpace, 1 GB by default,
// src/runtime/model.cc:2219) is large enough for the sort, else the atomic variant.
void Embedding::backward_task(const Task *task, const std::vector<PhysicalRegion> &regions,
                              Context ctx, Runtime *runtime) {
  assert(regions.size() == 3 && task->regions.size() == 3);
  const EmbeddingMeta *m = meta_of(task);
  TensorAccessorR<int64_t, 2> in(regions[0], task->regions[0], FID_DATA, ctx, runtime);
  TensorAccessorR<float, 2> og(regions[1], task->regions[1], FID_DATA, ctx, runtime);
  TensorAccessorW<float, 2> tab(regions[2], task->regions[2], FID_DATA, ctx, runtime, true /*readOutput*/);
  const Shape f = shape_of(in, og, tab);
  const size_t need = ffemb_backward_workspace(f.batch * f.bag, f.rows, f.dim);
  void *ws = m->handle.workSpace;
  const bool sorted = ws != NULL && need != 0 && need <= m->handle.workSpaceSize &&
                      ((uintptr_t)ws & 255) == 0 && f.dim % 4 == 0 && f.dim <= 512;
  if (sorted)
    die_unless_ok(ffemb_backward_dense(in.ptr, og.ptr, tab.ptr, f.bag, f.dim, f.batch,
                                       f.rows, (int)m->aggr, ws, m->handle.workSpaceSize,
                                       task_stream()),
                  "ffemb_backward_dense");
  else
    die_unless_ok(ffemb_backward_dense_atomic(in.ptr, og.ptr, tab.ptr, f.bag, f.dim, f.batch,
                                              (int)m->aggr, task_stream()),
                  "ffemb_backward_dense_atomic");
  if (m->profiling) checkCUDA(cudaDeviceSynchronize());
}

// ---- fused mode (opt-in) -------------------------------------------------------
// regions[0] input (RO), regions[1] output-grad (RO), regions[2] the TABLE itself (RW).
// task->args carries the SGD learning rate as one float (SGDOptimizer::lr, include/optimizer.h:55).
// Valid when momentum == weight_decay == 0 (src/runtime/optimizer_kernel.cu:31-39): replaces
// Op::zero_grad + backward_task + SGDOptimizer::update for this table.
void ffemb_embedding_backward_sgd_task(const Task *task, const std::vector<PhysicalRegion> &regions,
                                       Context ctx, Runtime *runtime) {
  assert(regions.size() == 3 && task->regions.size() == 3);
  assert(task->arglen == sizeof(float));
  const float lr = *(const float *)task->args;
  const EmbeddingMeta *m = meta_of(task);
  TensorAccessorR<int64_t, 2> in(regions[0], task->regions[0], FID_DATA, ctx, runtime);
  TensorAccessorR<float, 2> og(regions[1], task->regions[1], FID_DATA, ctx, runtime);
  TensorAccessorW<float, 2> tab(regions[2], task->regions[2], FID_DATA, ctx, runtime, true /*readOutput*/);
  const Shape f = shape_of(in, og, tab);
  die_unless_ok(ffemb_backward_sgd(in.ptr, og.ptr, tab.ptr, f.bag, f.dim, f.batch, f.rows,
                                   (int)m->aggr, lr, m->handle.workSpace, m->handle.workSpaceSize,
                                   task_stream()),
                "ffemb_backward_sgd");
  if (m->profiling) checkCUDA(cudaDeviceSynchronize());
}
