// ffemb_exchange.cu — the cross-GPU side of the table-wise exchange: ordering flags in peer-mapped
// memory and the output-grad push over NVLink, for sm_100a.
//
// The reference has no collective on this path: pooled outputs and output-grads live in zero-copy
// host memory and Legion copies batch slices between the table owner and the data-parallel Concat
// (/root/reference/src/ops/embedding.cu:295-298,377-380 MAP_TO_ZC_MEMORY; src/ops/concat.cu:88-97),
// ordered by Legion's dependence analysis.  Here the transfers are stores into peer memory issued by
// the kernels themselves, so the ORDERING has to be explicit too:
//   * ffemb_sync_signal / ffemb_sync_wait: per-slot epoch flags in a small pad that every rank maps
//     (raw peer pointers: cudaIpc, cudaDeviceEnablePeerAccess inside one Legion process, or symmetric
//     memory).  signal = "everything enqueued before me on this stream is visible to you" (fence.sys +
//     st.release.sys into every peer's pad); wait = a one-warp kernel that spins (ld.acquire.sys) until
//     every peer's flag for the slot reached this rank's next epoch.  The epochs live in the pad itself,
//     so neither call takes a step number and both can be captured in a CUDA graph.  A waiter occupies
//     one warp: it can never starve the kernels it waits for.
//   * push: this rank's dx [n][row_stride] (gradient of its Concat input) -> column block of every
//     table part -> rows [first, first + n) of the part owner's output-grad buffer.  Two kernels:
//       push_bulk_kernel  one warp per CTA moves tiles of up to 8 KB: the rows of a tile are fetched with
//                         cp.async.bulk (global -> shared, mbarrier complete_tx) and leave as ONE
//                         contiguous cp.async.bulk (shared -> peer global) — large posted NVLink writes,
//                         no registers, a few dozen resident warps: it runs UNDER the gather-pool and
//                         segmented-sum kernels of the neighbouring table group;
//       exchange_bwd_kernel (ffemb_forward.cu) 16-byte LDG/STG per lane — the measured comparison and the
//                         fallback for unaligned shapes.
#include <string.h>

#include "ffemb_internal.cuh"

namespace ffemb {

// ---- flags ---------------------------------------------------------------------------------------
// pad of one rank: FFEMB_SYNC_SLOTS x (FFEMB_MAX_PEERS + 2) uint64 words
//   [slot][src]              last epoch rank `src` signalled on this slot (written by src, remotely)
//   [slot][FFEMB_MAX_PEERS]     this rank's signal count on the slot
//   [slot][FFEMB_MAX_PEERS + 1] this rank's wait count on the slot
static constexpr int kPadStride = FFEMB_MAX_PEERS + 2;

struct SyncArgs {
  uint64_t *pad[FFEMB_MAX_PEERS];
  int32_t npeers, rank;
};

__device__ __forceinline__ void st_release_sys(uint64_t *p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(32) sync_signal_kernel(const __grid_constant__ SyncArgs a, int slot) {
  uint64_t *mine = a.pad[a.rank] + (size_t)slot * kPadStride;
  uint64_t e = 0;
  if (threadIdx.x == 0) {
    e = mine[FFEMB_MAX_PEERS] + 1;
    mine[FFEMB_MAX_PEERS] = e;
  }
  e = __shfl_sync(0xffffffffu, (unsigned long long)e, 0);
  __threadfence_system();  // the producers' stores (earlier kernels on this stream) before the flag
  if ((int)threadIdx.x < a.npeers)
    st_release_sys(a.pad[threadIdx.x] + (size_t)slot * kPadStride + a.rank, e);
}

// A lost peer must not hang the box: after kWaitTimeoutNs the waiter traps (the stream then fails
// with a launch error the caller sees) instead of spinning to the scheduler's limit.
static constexpr uint64_t kWaitTimeoutNs = 20ull * 1000 * 1000 * 1000;

__global__ void __launch_bounds__(32) sync_wait_kernel(const __grid_constant__ SyncArgs a, int slot) {
  uint64_t *mine = a.pad[a.rank] + (size_t)slot * kPadStride;
  uint64_t e = 0;
  if (threadIdx.x == 0) {
    e = mine[FFEMB_MAX_PEERS + 1] + 1;
    mine[FFEMB_MAX_PEERS + 1] = e;
  }
  e = __shfl_sync(0xffffffffu, (unsigned long long)e, 0);
  if ((int)threadIdx.x < a.npeers) {
    const uint64_t *f = mine + threadIdx.x;
    const uint64_t t0 = globaltimer_ns();
    unsigned spins = 0;
    while (ld_acquire_sys(f) < e) {
      if ((++spins & 1023u) == 0 && globaltimer_ns() - t0 > kWaitTimeoutNs) __trap();
      __nanosleep(64);
    }
  }
  __syncwarp();
  __threadfence_system();
}

// ---- push with bulk copies -------------------------------------------------------------------------
static constexpr int kPushTileBytes = 8192, kPushStages = 4, kPushAhead = kPushStages - 2;

struct PushPart {
  const float *src;  // dx + column offset of the part
  float *dst;        // owner's output-grad buffer for the part (peer-mapped), row `first` already applied
  int32_t w;         // floats per row of the part
  int32_t rows_per_tile;
};
struct PushArgs {
  PushPart p[FFEMB_MAX_TABLES];
  int32_t nparts;
  int32_t max_tiles;  // tiles of the part with the most tiles
  int64_t n, stride;  // samples, floats between consecutive rows of dx
};

// Work item i -> (row block i / nparts, part i % nparts): every wave of CTAs writes to all parts, i.e.
// to all destination GPUs at once.  One warp per CTA; lane r fetches row r of the tile, lane 0 owns the
// barriers and issues the single contiguous store of the tile.
__global__ void __launch_bounds__(32) push_bulk_kernel(const __grid_constant__ PushArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *stage = smem_raw;
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + kPushStages * kPushTileBytes);
  const int lane = threadIdx.x;
  if (lane == 0) {
    for (int s = 0; s < kPushStages; s++) mbar_init(bar + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();
  const int64_t items = (int64_t)a.max_tiles * a.nparts;
  // tile number `it` of this CTA -> item
  auto item_of = [&](int64_t it) { return (int64_t)blockIdx.x + it * (int64_t)gridDim.x; };
  auto geometry = [&](int64_t item, const PushPart *&pp, int64_t &row0, int &rows) {
    const int part = (int)(item % a.nparts);
    const int64_t blk = item / a.nparts;
    pp = &a.p[part];
    row0 = blk * pp->rows_per_tile;
    const int64_t left = a.n - row0;
    rows = left <= 0 ? 0 : (int)(left < pp->rows_per_tile ? left : pp->rows_per_tile);
  };
  auto load = [&](int64_t it) {  // fetch the rows of tile `it` into stage it % S
    const int64_t item = item_of(it);
    if (item >= items) return;
    const PushPart *pp;
    int64_t row0;
    int rows;
    geometry(item, pp, row0, rows);
    if (rows == 0) return;
    const int s = (int)(it % kPushStages);
    const uint32_t rb = (uint32_t)pp->w * 4u;
    if (lane == 0) mbar_expect_tx(bar + s, rb * (uint32_t)rows);
    __syncwarp();
    for (int r = lane; r < rows; r += 32)
      bulk_g2s(stage + (size_t)s * kPushTileBytes + (size_t)r * rb, pp->src + (row0 + r) * a.stride, rb, bar + s);
  };
  uint32_t phases = 0;  // bit s = parity of the phase stage s's barrier completes next
#pragma unroll
  for (int k = 0; k < kPushAhead; k++) load(k);
  for (int64_t it = 0; item_of(it) < items; it++) {
    // stage of tile it + AHEAD was last read by the store of tile it - 2: at most one younger group
    // (tile it - 1) may still be pending
    if (lane == 0) bulk_wait_read<1>();
    __syncwarp();
    load(it + kPushAhead);
    const PushPart *pp;
    int64_t row0;
    int rows;
    geometry(item_of(it), pp, row0, rows);
    const int s = (int)(it % kPushStages);
    if (rows > 0) {
      if (lane == 0) {
        unsigned spins = 0;
        while (!mbar_try_wait(bar + s, (phases >> s) & 1u))
          if (++spins > (1u << 26)) __trap();  // a lost bulk copy must not hang the box
        bulk_s2g(pp->dst + row0 * pp->w, stage + (size_t)s * kPushTileBytes, (uint32_t)pp->w * 4u * (uint32_t)rows);
      }
      phases ^= 1u << s;
    }
    if (lane == 0) bulk_commit();  // one group per tile (empty for a skipped item)
    __syncwarp();
  }
  if (lane == 0) bulk_wait<0>();
}

int launch_push_bulk(const TableBatch &tb, const int32_t *widths, const float *dx, int64_t n, int64_t first,
                     int64_t stride, cudaStream_t s, int warps_per_sm) {
  PushArgs a;
  memset(&a, 0, sizeof(a));
  a.nparts = tb.num_tables;
  a.n = n;
  a.stride = stride;
  int64_t max_tiles = 0;
  for (int t = 0; t < tb.num_tables; t++) {
    const int w = widths[t];
    a.p[t].src = dx + tb.t[t].lookup_base;
    a.p[t].dst = tb.t[t].out + first * (int64_t)w;
    a.p[t].w = w;
    a.p[t].rows_per_tile = kPushTileBytes / (w * 4);
    const int64_t tiles = (n + a.p[t].rows_per_tile - 1) / a.p[t].rows_per_tile;
    max_tiles = tiles > max_tiles ? tiles : max_tiles;
  }
  if (max_tiles >= (1ll << 31)) {
    set_error("push: too many samples");
    return FFEMB_EINVAL;
  }
  a.max_tiles = (int32_t)max_tiles;
  const int64_t items = max_tiles * tb.num_tables;
  // resident CTAs (= warps) per SM.  Measured at N = 2 (profiles/r2_scaling.md): the push alone moves
  // 218 MB in 0.56 / 0.48 / 0.42 / 0.37 ms with 2 / 4 / 8 / 16, but UNDER the neighbouring group's
  // gather-pool / segmented-sum kernels the step is fastest with 2 (1.754 vs 1.841 / 1.926 / 1.926 ms):
  // a caller that overlaps the push asks for few warps, one that runs it alone for many.
  int ctas_per_sm = warps_per_sm > 0 ? (warps_per_sm > 16 ? 16 : warps_per_sm) : 4;
  if (tuning().xpush > 1) ctas_per_sm = tuning().xpush;  // tuning aid
  const int64_t cap = 148ll * ctas_per_sm;
  const unsigned grid = (unsigned)(items < cap ? items : cap);
  const size_t sm = (size_t)kPushStages * kPushTileBytes + kPushStages * sizeof(uint64_t);
  static bool attr_done = false;  // (idempotent; a race only repeats the call)
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(push_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e != cudaSuccess) return cuda_fail(e, "push smem opt-in");
    attr_done = true;
  }
  ProfScope prof(FFEMB_PHASE_EXCHANGE, 1, s);
  push_bulk_kernel<<<grid, 32, sm, s>>>(a);
  return cuda_fail(cudaGetLastError(), "push launch");
}

static int pack_sync(const ffemb_sync_t *sy, int slot, SyncArgs &a) {
  if (!sy || sy->num_peers < 1 || sy->num_peers > FFEMB_MAX_PEERS || sy->rank < 0 || sy->rank >= sy->num_peers ||
      slot < 0 || slot >= FFEMB_SYNC_SLOTS) {
    set_error("sync: invalid descriptor or slot %d", slot);
    return FFEMB_EINVAL;
  }
  memset(&a, 0, sizeof(a));
  for (int h = 0; h < sy->num_peers; h++) {
    if (!sy->peer_pad[h] || (uintptr_t)sy->peer_pad[h] % 8 != 0) {
      set_error("sync: peer_pad[%d] is NULL or misaligned", h);
      return FFEMB_EINVAL;
    }
    a.pad[h] = sy->peer_pad[h];
  }
  a.npeers = sy->num_peers;
  a.rank = sy->rank;
  return FFEMB_OK;
}

}  // namespace ffemb

using namespace ffemb;

extern "C" {

int ffemb_exchange_backward_parts(const ffemb_table_t *parts, int num_parts, const float *dx,
                                  int64_t n_samples, int64_t first_sample, int64_t row_stride,
                                  ffemb_stream_t stream) {
  if (!parts || num_parts < 0 || num_parts > FFEMB_MAX_TABLES || n_samples < 0 || first_sample < 0 ||
      row_stride < 1 || (n_samples > 0 && !dx)) {
    set_error("exchange_backward_parts: invalid arguments");
    return FFEMB_EINVAL;
  }
  if (num_parts == 0 || n_samples == 0) return FFEMB_OK;
  TableBatch tb;
  memset(&tb, 0, sizeof(tb));
  tb.num_tables = num_parts;
  int32_t widths[FFEMB_MAX_TABLES];
  bool bulk = row_stride % 4 == 0 && (uintptr_t)dx % 16 == 0 && tuning().xpush != 0;
  for (int p = 0; p < num_parts; p++) {
    const int32_t w = parts[p].L;
    if (!parts[p].out || w < 1 || parts[p].reserved < 0 || (int64_t)parts[p].reserved + w > row_stride) {
      set_error("exchange_backward_parts: part %d: invalid destination, width or column offset", p);
      return FFEMB_EINVAL;
    }
    tb.t[p].out = parts[p].out;
    tb.t[p].lookup_base = (uint32_t)parts[p].reserved;
    widths[p] = w;
    bulk = bulk && w % 4 == 0 && w * 4 <= 8192 && parts[p].reserved % 4 == 0 && (uintptr_t)parts[p].out % 16 == 0;
  }
  if (bulk)
    return launch_push_bulk(tb, widths, dx, n_samples, first_sample, row_stride, (cudaStream_t)stream,
                            parts[0].R > 0 && parts[0].R <= 64 ? (int)parts[0].R : 0);
  // fallback / comparison (FFEMB_XPUSH=0): the load-store kernel, one launch per run of equal widths
  for (int p0 = 0; p0 < num_parts;) {
    int p1 = p0;
    TableBatch sub;
    memset(&sub, 0, sizeof(sub));
    while (p1 < num_parts && widths[p1] == widths[p0]) sub.t[sub.num_tables++] = tb.t[p1++];
    int rc = launch_exchange_backward(sub, widths[p0], dx, n_samples, first_sample, row_stride, (cudaStream_t)stream);
    if (rc) return rc;
    p0 = p1;
  }
  return FFEMB_OK;
}

size_t ffemb_sync_pad_bytes(void) { return (size_t)FFEMB_SYNC_SLOTS * kPadStride * sizeof(uint64_t); }

int ffemb_sync_signal(const ffemb_sync_t *sync, int slot, ffemb_stream_t stream) {
  SyncArgs a;
  int rc = pack_sync(sync, slot, a);
  if (rc) return rc;
  ProfScope prof(FFEMB_PHASE_SYNC, 1, (cudaStream_t)stream);
  sync_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, slot);
  return cuda_fail(cudaGetLastError(), "sync signal launch");
}

int ffemb_sync_wait(const ffemb_sync_t *sync, int slot, ffemb_stream_t stream) {
  SyncArgs a;
  int rc = pack_sync(sync, slot, a);
  if (rc) return rc;
  ProfScope prof(FFEMB_PHASE_SYNC, 1, (cudaStream_t)stream);
  sync_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, slot);
  return cuda_fail(cudaGetLastError(), "sync wait launch");
}

}  // extern "C"
