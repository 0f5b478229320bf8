// ffemb_static.cu — ffemb_backward_sgd_static: the fused backward + SGD of ONE table with its scratch in
// a file-scope device array instead of a caller-supplied workspace.  No counterpart in the reference:
// added for callers that cannot allocate device memory (code
// under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  It runs the same launch
// sequence as ffemb_backward_sgd (keygen, sort, segmented sum + row update) — only where the sort
// buffers, partial rows and flags live differs — so its results are bit-identical to that entry point.
//
// The array is held by every GPU that loads the library until the process exits, hence the modest cap.
// Every call on a device shares it and it keeps the last call's contents: the span a call uses is
// cleared on the call's stream first, and calls that overlap on the GPU (two streams, two host threads
// on one device) race — serialise them.
#include "ffemb_internal.cuh"

#include <cuda_runtime.h>

#include <atomic>

namespace {

constexpr size_t kStaticScratchBytes = (size_t)FFEMB_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// Device code names the array directly; this (never launched) kernel also keeps the symbol referenced
// from device code, so no whole-program pass can drop it as unused.
__global__ void static_scratch_anchor_kernel(unsigned char **where) { *where = g_static_scratch; }

// g_static_scratch is a host-side stand-in here, not a device address: the kernels get the address
// cudaGetSymbolAddress reports for the CURRENT device, looked up once per device.
constexpr int kMaxDevices = 64;
std::atomic<void *> g_dev_ptr[kMaxDevices];  // zero-initialised: not looked up yet

int static_scratch(void **out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return ffemb::cuda_fail(e, "cudaGetDevice (static scratch)");
  if (dev < 0 || dev >= kMaxDevices) {
    ffemb::set_error("static scratch: device ordinal %d out of range", dev);
    return FFEMB_EINVAL;
  }
  (void)&static_scratch_anchor_kernel;
  void *p = g_dev_ptr[dev].load(std::memory_order_acquire);
  if (!p) {  // first call on this device (two host threads racing here store the same address)
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return ffemb::cuda_fail(e, "cudaGetSymbolAddress(static scratch)");
    if (!p) {
      ffemb::set_error("cudaGetSymbolAddress(static scratch) returned a null address");
      return FFEMB_ECUDA;
    }
    g_dev_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return FFEMB_OK;
}

}  // namespace

extern "C" {

size_t ffemb_static_scratch_bytes(void) { return kStaticScratchBytes; }

int ffemb_backward_sgd_static(const int64_t *idx, const float *outgrad, float *table, int64_t L, int64_t D,
                              int64_t B, int64_t R, int aggr, float lr, ffemb_stream_t stream) {
  if (L < 0 || B < 0 || D <= 0 || R < 0) {
    ffemb::set_error("backward_sgd_static: invalid shape (L %lld, D %lld, B %lld, R %lld)", (long long)L,
                     (long long)D, (long long)B, (long long)R);
    return FFEMB_EINVAL;
  }
  if (B * L == 0) return FFEMB_OK;
  if (R <= 0) {
    ffemb::set_error("backward_sgd_static needs the row count R > 0");
    return FFEMB_EINVAL;
  }
  if (B * L >= (1ll << 31)) {
    ffemb::set_error("backward_sgd_static: too many lookups for one call (%lld)", (long long)(B * L));
    return FFEMB_EINVAL;
  }
  const size_t need = ffemb_backward_workspace(B * L, R, D);
  if (need == 0 || need > kStaticScratchBytes) {
    ffemb::set_error("backward_sgd_static: %lld lookups over %lld rows of %lld floats need %zu bytes of scratch, "
                     "the static array holds %zu (FFEMB_STATIC_SCRATCH_BYTES): use ffemb_backward_sgd with a "
                     "workspace of your own",
                     (long long)(B * L), (long long)R, (long long)D, need, kStaticScratchBytes);
    return FFEMB_EWORKSPACE;
  }
  void *ws = nullptr;
  int rc = static_scratch(&ws);
  if (rc) return rc;
  // the array keeps the previous call's sort buffers, partial rows and flags: start from zero
  cudaError_t e = cudaMemsetAsync(ws, 0, need, (cudaStream_t)stream);
  if (e != cudaSuccess) return ffemb::cuda_fail(e, "clearing the static scratch");
  return ffemb_backward_sgd(idx, outgrad, table, L, D, B, R, aggr, lr, ws, need, stream);
}

}  // extern "C"
