// ffemb_input.cu — sparse-input delivery: the index batches of ALL tables of a GPU cut out of the
// sample-major dataset in ONE launch.
//
// Replaces DataLoader::load_sparse_input (/root/reference/examples/cpp/DLRM/dlrm.cu:19-58), which
// the reference runs once per table and step (DataLoader::next_batch,
// examples/cpp/DLRM/dlrm.cc:482-520): cudaHostAlloc of a staging buffer, a host loop
//     input_zc[i*in_dim + j] = full[idxs[i] * num_sparse_inputs*in_dim + my_input_idx*in_dim + j]
// a synchronous cudaMemcpy to the table's [B][in_dim] input, cudaFreeHost, cudaDeviceSynchronize.
// The full dataset already sits in zero-copy (mapped, pinned) host memory there (MAP_TO_ZC_MEMORY,
// dlrm.cc:505-509), so the GPU can read it directly: here one kernel walks the B sample rows of
// the batch — contiguous row_len*8-byte pieces of the dataset, read coalesced whether the pointer
// is device memory or mapped host memory (then the reads are the PCIe transfer itself, issued by
// the copy kernel, many in flight) — and scatters every column block to its table's index tensor.
// No staging buffer, no per-table launch, no host synchronisation.
// Tables may have different bag widths (column offset and width per table); columns that belong
// to no table are skipped.
#include <string.h>

#include "ffemb_internal.cuh"

namespace ffemb {

struct ColumnMap {                        // sorted by column offset, passed by value
  int64_t *dst[FFEMB_MAX_TABLES];         // [B][width]
  int32_t col[FFEMB_MAX_TABLES + 1];      // first column of the block; col[n] = sentinel (INT_MAX)
  int32_t width[FFEMB_MAX_TABLES];
  int32_t n;
};

static constexpr int kInThreads = 256;
static constexpr int kInUnroll = 4;

__global__ void __launch_bounds__(kInThreads)
load_sparse_kernel(const int64_t *__restrict__ full, int64_t row_len, const int32_t *__restrict__ sample_ids,
                   int64_t first_sample, int64_t B, const __grid_constant__ ColumnMap cm) {
  const int64_t total = B * row_len;
  const int64_t stride = (int64_t)gridDim.x * kInThreads;
  const uint64_t pol = policy_evict_first();  // read once: no L1 allocation, first out of L2
  for (int64_t e0 = (int64_t)blockIdx.x * kInThreads + threadIdx.x; e0 < total; e0 += kInUnroll * stride) {
    int64_t v[kInUnroll];
    int64_t i[kInUnroll];
    int c[kInUnroll];
#pragma unroll
    for (int u = 0; u < kInUnroll; u++) {  // all loads first: kInUnroll PCIe / HBM reads in flight per thread
      const int64_t e = e0 + u * stride;
      i[u] = e / row_len;
      c[u] = (int)(e - i[u] * row_len);
      v[u] = 0;
      if (e < total) {
        const int64_t row = sample_ids ? (int64_t)sample_ids[i[u]] : first_sample + i[u];
        v[u] = ldg_idx_stream(full + row * row_len + c[u], pol);
      }
    }
#pragma unroll
    for (int u = 0; u < kInUnroll; u++) {
      if (e0 + u * stride >= total) continue;
      int lo = 0, hi = cm.n - 1;  // last block whose first column <= c
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (cm.col[mid] <= c[u]) lo = mid; else hi = mid - 1;
      }
      const int j = c[u] - cm.col[lo];
      if (j >= 0 && j < cm.width[lo]) cm.dst[lo][i[u] * cm.width[lo] + j] = v[u];
    }
  }
}

int launch_load_sparse(const ffemb_table_t *tables, int num_tables, const int64_t *full, int64_t row_len,
                       const int32_t *sample_ids, int64_t first_sample, int64_t B, cudaStream_t s) {
  ColumnMap cm;
  memset(&cm, 0, sizeof(cm));
  int order[FFEMB_MAX_TABLES];
  int n = 0;
  for (int t = 0; t < num_tables; t++)
    if (tables[t].L > 0) order[n++] = t;
  for (int a = 1; a < n; a++)  // insertion sort by column offset (n <= 64)
    for (int b = a; b > 0 && tables[order[b]].reserved < tables[order[b - 1]].reserved; b--) {
      const int x = order[b]; order[b] = order[b - 1]; order[b - 1] = x;
    }
  for (int k = 0; k < n; k++) {
    const ffemb_table_t &x = tables[order[k]];
    if (k > 0 && x.reserved < cm.col[k - 1] + cm.width[k - 1]) {
      set_error("load_sparse_inputs: column blocks of tables %d and %d overlap", order[k - 1], order[k]);
      return FFEMB_EINVAL;
    }
    cm.dst[k] = const_cast<int64_t *>(x.idx);
    cm.col[k] = x.reserved;
    cm.width[k] = x.L;
  }
  cm.col[n] = 0x7fffffff;
  cm.n = n;
  if (n == 0 || B == 0) return FFEMB_OK;
  ProfScope prof(FFEMB_PHASE_INPUT, 1, s);
  const int64_t total = B * row_len;
  int64_t blocks = (total + (int64_t)kInThreads * kInUnroll - 1) / ((int64_t)kInThreads * kInUnroll);
  if (blocks > 148 * 8) blocks = 148 * 8;  // 8 CTAs per SM, grid-stride beyond that
  load_sparse_kernel<<<(unsigned)blocks, kInThreads, 0, s>>>(full, row_len, sample_ids, first_sample, B, cm);
  return cuda_fail(cudaGetLastError(), "load_sparse_inputs launch");
}

}  // namespace ffemb
