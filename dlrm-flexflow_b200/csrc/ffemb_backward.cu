// ffemb_backward.cu — deterministic sort-then-segmented-reduce scatter-add, fused with the
// sparse SGD row update, for sm_100a; plus the atomic variant kept as the measured comparison.
//
// Replaces, per training step (reference = /root/reference):
//   Op::zero_grad's dense zero-fill of the weight-grad   src/runtime/model.cc:466-490
//   embed_backward's B*L*D float atomicAdds               src/ops/embedding.cu:192-217
//   sgd_update over all R*D elements                      src/runtime/optimizer_kernel.cu:23-41
// with work proportional to the lookups only; no dense gradient exists.
//
// Pipeline (all tables of a GPU at once):
//   1. keygen : lookup n=(table t, bag b, slot j) -> key = row_base[t] + idx  (global row id),
//               val = (t << 26) | b (table, bag).  Lookups are enumerated in (t, b, j) order.
//   2. sort   : stable LSD radix sort of (key, val) on the ceil(log2(total_rows)) key bits
//               (cub::DeviceRadixSort, onesweep) -> equal rows adjacent, bags ascending.
//   3. level 1: the sorted array is cut into fixed tiles of CH1 entries; a slot of G=D/4 lanes
//               walks one tile sequentially (entries prefetched UNR at a time: the gathered
//               output-grad row AND, for an entry that ends a run, the table row), summing
//               runs of equal keys in order.  A run that lies inside the tile is applied
//               straight to the table: W[r] = fma(-lr, sum, W[r]).  A run cut by a tile edge
//               is written out as a partial row (head partial -> slot 0, tail partial ->
//               slot 1 of the tile) with its key.
//   4. level k: the same walk over the previous level's partial rows (fixed tiles of CH2
//               entries, invalid slots are transparent) until one tile covers everything.
//               Small inputs finish all remaining levels inside one CTA (one launch).
// The shape of the summation tree depends only on (sorted order, CH1, CH2): bit-reproducible
// run to run, unlike the reference's atomics.  Rows with a single lookup get exactly the
// reference's arithmetic (g summed into 0, then one FFMA), i.e. bit-identical updates.
#include <cub/block/block_radix_sort.cuh>
#include <cub/device/device_radix_sort.cuh>

#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <type_traits>

#include "ffemb_internal.cuh"

namespace ffemb {

static constexpr int kThreads = 256;
// level-1 tile (sorted lookups per slot).  FFEMB_CH1_WIDE (rows of >= 64 floats, where a slot is half a
// warp or more) is a tuning aid: 128 halves the partial rows the upper levels walk (cfg3 reduce_up
// 0.120 -> 0.093 ms) but the longer serial walk costs level 1 as much (1.175 -> 1.188 ms; cfg5
// 11.8 -> 12.3 ms), so it stays at 64.
#ifndef FFEMB_CH1_WIDE
#define FFEMB_CH1_WIDE 64
#endif
template <int G> struct Ch1 { static constexpr int value = G >= 16 ? FFEMB_CH1_WIDE : 64; };
static inline int ch1_of(int64_t D) { return D / 4 > 8 ? FFEMB_CH1_WIDE : 64; }
static constexpr int CH2 = 32;  // level-k tile (partial rows per slot)
static constexpr int kSmallLevel = 1 << 8;   // <= this many partial slots: finish in one CTA (a CTA walks 8 tiles at a time, ~6 us a round)
static constexpr int64_t kStreamTableBytes = 64ll << 20;  // tables above this do not stay in L2
static constexpr int BAG_BITS = 26;          // sort payload = (table << 26) | bag
static constexpr uint32_t BAG_MASK = (1u << BAG_BITS) - 1;

__device__ __forceinline__ void f4_add(float4 &a, const float4 &b) {
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
}
__device__ __forceinline__ float4 f4_div(float4 a, float d) {
  return make_float4(a.x / d, a.y / d, a.z / d, a.w / d);
}
__device__ __forceinline__ float4 f4_div_rn(float4 a, float d, float r) {
  return f4_div_exact(a, d, r);
}
__device__ __forceinline__ float4 ld_f4(const float *p) { return *reinterpret_cast<const float4 *>(p); }
__device__ __forceinline__ void st_f4(float *p, float4 v) { *reinterpret_cast<float4 *>(p) = v; }

template <typename K> struct KeyTraits;
template <> struct KeyTraits<uint32_t> { static constexpr uint32_t INV = 0xffffffffu; };
template <> struct KeyTraits<uint64_t> { static constexpr uint64_t INV = ~0ull; };

template <typename K> __device__ __forceinline__ K shfl_key(K v, int src);
template <> __device__ __forceinline__ uint32_t shfl_key<uint32_t>(uint32_t v, int src) {
  return __shfl_sync(0xffffffffu, v, src);
}
template <> __device__ __forceinline__ uint64_t shfl_key<uint64_t>(uint64_t v, int src) {
  return (uint64_t)shfl_i64((int64_t)v, src);
}

// ---- 1. keygen -----------------------------------------------------------------------------
template <typename K>
__global__ void __launch_bounds__(kThreads)
keygen_kernel(const __grid_constant__ TableBatch tb, int64_t B, K *keys, uint32_t *vals) {
  const TableDesc &td = tb.t[blockIdx.y];
  const int L = td.L;
  const int64_t n = B * (int64_t)L;
  const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  if (i >= n) return;
  const int64_t b = i / L;
  const int64_t row = __ldg(td.idx + i);
  K key = (K)(tb.row_base[blockIdx.y] + (uint64_t)row);
  if (tb.padded && row < 0) key = (K)tb.row_base[tb.num_tables];  // empty slot: past every row
  keys[td.lookup_base + i] = key;
  vals[td.lookup_base + i] = ((uint32_t)blockIdx.y << BAG_BITS) | (uint32_t)b;
}


// ---- 1+2 for small batches: keygen + sort of a table's lookups inside ONE CTA ---------------------
// The device-wide radix sort is a chain of 6-7 dependent launches whatever the size; at 2K-16K lookups
// per table (BASELINE configs 1 and 2) that chain IS the step (55 of 82 us, 87 of 147 us).  The key
// order is (table, row) and a table's lookups are contiguous in the lookup numbering, so each table
// can be sorted on its own: one CTA of 1024 threads per table reads the indices straight from the
// index tensor (keygen fused), sorts (local row, (table, bag)) pairs on the ceil(log2 R_t) row bits in
// shared memory (cub::BlockRadixSort: LSD, stable — bags ascend inside a run of equal rows, like the
// device-wide sort) and writes the table's stretch of the sorted arrays.  One launch, all tables in
// parallel, 2-6 passes of 4 bits each in shared memory.
static constexpr int kSegThreads = 1024;
static constexpr int kSegMaxItems = 16;  // up to 16384 lookups per table

template <int ITEMS>
__global__ void __launch_bounds__(kSegThreads)
seg_sort_kernel(const __grid_constant__ TableBatch tb, int64_t B, uint32_t *__restrict__ keys_out,
                uint32_t *__restrict__ vals_out) {
  using Sort = cub::BlockRadixSort<uint32_t, kSegThreads, ITEMS, uint32_t>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  typename Sort::TempStorage &tmp = *reinterpret_cast<typename Sort::TempStorage *>(smem_raw);
  const int t = blockIdx.x;
  const TableDesc &td = tb.t[t];
  const int L = td.L;
  const int n = (int)(B * (int64_t)L);
  if (n == 0) return;
  uint32_t k[ITEMS], v[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const int q = threadIdx.x * ITEMS + i;
    k[i] = 0xffffffffu;  // padding: sorts behind every lookup (ties keep the input order: padding is last)
    v[i] = 0u;
    if (q < n) {
      k[i] = (uint32_t)__ldg(td.idx + q);
      v[i] = ((uint32_t)t << BAG_BITS) | (uint32_t)(q / L);
    }
  }
  int end_bit = 32 - __clz((unsigned)(td.R > 1 ? td.R - 1 : 1));
  Sort(tmp).Sort(k, v, 0, end_bit);
  const uint32_t base = (uint32_t)tb.row_base[t];
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    const int q = threadIdx.x * ITEMS + i;
    if (q < n) {
      keys_out[td.lookup_base + q] = base + k[i];
      vals_out[td.lookup_base + q] = v[i];
    }
  }
}

template <int ITEMS>
static cudaError_t launch_seg_sort(const TableBatch &tb, int64_t B, uint32_t *keys_out, uint32_t *vals_out,
                                   cudaStream_t s) {
  using Sort = cub::BlockRadixSort<uint32_t, kSegThreads, ITEMS, uint32_t>;
  const size_t sm = sizeof(typename Sort::TempStorage);
  cudaError_t e = cudaFuncSetAttribute(seg_sort_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  if (e != cudaSuccess) return e;
  seg_sort_kernel<ITEMS><<<tb.num_tables, kSegThreads, sm, s>>>(tb, B, keys_out, vals_out);
  return cudaGetLastError();
}

// table of a global row id: row_base is ascending, <= 65 entries
__device__ __forceinline__ int find_table(const TableBatch &tb, uint64_t key) {
  int lo = 0, hi = tb.num_tables - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (tb.row_base[mid] <= key) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// ---- 4. one tile of the upper levels of the segmented sum ------------------------------------
// Entries are the previous level's partial rows rows_in[p]; keys may be INV (transparent: skipped,
// and runs continue across ONE invalid slot — by construction two partials of the same run are
// never separated by more than one, see header).  (Level 1, whose entries are sorted lookups, has
// its own kernels: reduce1_kernel / reduce1_ring_kernel.)
template <typename K, int G, int C, int MODE>
__device__ __forceinline__ void reduce_tile(const TableBatch &tb, const K *keys,
                                            const float *rows_in, int64_t n, int64_t tile, int CH,
                                            bool active, int D, float lr,
                                            K *keys_out, float *rows_out, int sub, int sl,
                                            const OptArgs *oa = nullptr, const uint8_t *consumed = nullptr) {
  constexpr K INV = KeyTraits<K>::INV;
  constexpr int UNR = (C == 1) ? 4 : (C == 2 ? 2 : 1);
  const int D4 = D >> 2;
  // a slot combine_kernel has already summed and applied reads as invalid
  auto ldk = [&](int64_t q) -> K { return (consumed && consumed[q]) ? INV : keys[q]; };
  const int64_t t0 = tile * CH;
  const int64_t t1 = active ? min(n, t0 + CH) : t0;

  float4 acc[C];
#pragma unroll
  for (int c = 0; c < C; c++) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  bool seen = false;        // a valid entry was met in this tile
  bool first_open = false;  // the tile's first run continues a run of the previous tile
  bool pending = false;     // acc holds an unflushed run
  K pend_key = INV;
  bool wrote_head = false;

  for (int64_t p0 = t0; p0 < t0 + CH; p0 += G) {  // trip count uniform across the warp
    const int64_t p = p0 + sl;
    K k = INV, kn = INV, kn2 = INV;
    if (p < t1) {
      k = ldk(p);
      if (p + 1 < n) kn = ldk(p + 1);
      if (p + 2 < n) kn2 = ldk(p + 2);
    }
    // does the run of entry p stop at p?  (lookahead crosses the tile edge on purpose)
    const bool is_end = (k != INV) && !(kn == k || (kn == INV && kn2 == k));

    for (int jj = 0; jj < G; jj += UNR) {
      if (p0 + jj >= t0 + CH) break;  // uniform
      float4 g[UNR][C], w[UNR][C];
      K ek[UNR];
      bool eend[UNR];
      float *wp[UNR];
      int et[UNR];
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const int src = sub * G + jj + u;
        ek[u] = shfl_key<K>(k, src);
        eend[u] = __shfl_sync(0xffffffffu, (int)is_end, src) != 0;
        wp[u] = nullptr;
        et[u] = 0;
        if (jj + u < G && ek[u] != INV) {
          const float *gp = rows_in + (p0 + jj + u) * (int64_t)D;
          if (eend[u]) {
            const int t = find_table(tb, (uint64_t)ek[u]);
            et[u] = t;
            wp[u] = tb.t[t].table + (int64_t)((uint64_t)ek[u] - tb.row_base[t]) * D;
          }
#pragma unroll
          for (int c = 0; c < C; c++) {
            if (sl + c * G < D4) {
              g[u][c] = ld_f4(gp + 4 * (sl + c * G));
              if (wp[u]) w[u][c] = ld_f4(wp[u] + 4 * (sl + c * G));
            }
          }
        } else {
          ek[u] = INV;
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        if (ek[u] == INV) continue;
        if (!seen) {  // first valid entry of the tile: does it continue the previous tile's run?
          seen = true;
          const int64_t q = p0 + jj + u;
          K km1 = INV, km2 = INV;
          if (q >= 1) km1 = ldk(q - 1);
          if (q >= 2) km2 = ldk(q - 2);
          first_open = (km1 == ek[u]) || (km1 == INV && km2 == ek[u]);
        }
#pragma unroll
        for (int c = 0; c < C; c++) {
          if (sl + c * G < D4) f4_add(acc[c], g[u][c]);
        }
        pending = true;
        pend_key = ek[u];
        if (eend[u]) {
          if (first_open) {  // head partial: the run began in an earlier tile
            if (sl == 0) keys_out[2 * tile] = ek[u];
#pragma unroll
            for (int c = 0; c < C; c++)
              if (sl + c * G < D4) st_f4(rows_out + (2 * tile) * (int64_t)D + 4 * (sl + c * G), acc[c]);
            wrote_head = true;
          } else {  // complete run: the one read-modify-write of this table row
#pragma unroll
            for (int c = 0; c < C; c++) {
              if (sl + c * G < D4) {
                float4 o = w[u][c];
                if (MODE == APPLY_SGD) {  // W -= lr*gt, contracted like the reference's FFMA
                  o.x = fmaf(-acc[c].x, lr, o.x); o.y = fmaf(-acc[c].y, lr, o.y);
                  o.z = fmaf(-acc[c].z, lr, o.z); o.w = fmaf(-acc[c].w, lr, o.w);
                } else if (MODE == APPLY_SGDX) {
                  o = sgdx_fragment(oa, et[u], tb.t[et[u]].table, wp[u] + 4 * (sl + c * G), o, acc[c],
                                    (uint64_t)ek[u] - tb.row_base[et[u]], sl == 0 && c == 0);
                } else {
                  f4_add(o, acc[c]);
                }
                st_f4(wp[u] + 4 * (sl + c * G), o);
              }
            }
          }
          first_open = false;
          pending = false;
#pragma unroll
          for (int c = 0; c < C; c++) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
  }
  if (active) {
    if (pending) {  // tail partial: the run goes on in a later tile
#pragma unroll
      for (int c = 0; c < C; c++)
        if (sl + c * G < D4) st_f4(rows_out + (2 * tile + 1) * (int64_t)D + 4 * (sl + c * G), acc[c]);
    }
    if (sl == 0) {
      if (!wrote_head) keys_out[2 * tile] = INV;
      keys_out[2 * tile + 1] = pending ? pend_key : INV;
    }
  }
}

// ---- 3. level 1, the hot kernel ---------------------------------------------------------------
// One slot of G lanes owns a tile of CH1 sorted lookups.  Two phases per warp:
//   phase 1 (parallel over lanes): every lane turns its share of the warp's SLOTS*CH1 entries
//     into a 16-byte descriptor in shared memory {output-grad row address, table row address if
//     the entry ENDS its run of equal keys, else 0}.  All integer decoding (table id, bag id,
//     run ends) happens here, once per entry, off the serial path.
//   phase 2 (serial per slot, no warp collectives): walk the tile in order, UNR entries per
//     group, software-pipelined (group k+1's loads are issued before group k is consumed).  Per
//     entry: one broadcast LDS.128, one 16-byte gather of the output-grad row fragment per lane
//     (L2 evict_last: a table's output-grad is re-read L times while its rows are reduced) and,
//     for run ends, the table row (L2 evict_first: touched once per step), then FADDs in sorted
//     order and at a run end the FFMA + store.
// Runs cut by a tile edge leave as partial rows for the upper levels (same contract as before).
// CTA -> work item map of level 1.  The sorted array holds the lookups of the tables far larger
// than L2 first (CTAs [0, split): their row read-modify-writes are HBM-bound), then those of the
// L2-resident tables (output-grad gathers, L2-bound).  Handing the two regions out proportionally
// interleaved keeps one table of each kind in flight at any time, so the two regimes overlap;
// two output-grads (2 x B*D*4 bytes) still fit L2.  Tile contents do not change: same results.
__device__ __forceinline__ uint32_t mix_cta(uint32_t i, uint32_t grid, uint32_t split) {
  if (split == 0 || split >= grid) return i;
  const uint32_t a0 = (uint32_t)(((uint64_t)i * split) / grid);
  const uint32_t a1 = (uint32_t)(((uint64_t)(i + 1) * split) / grid);
  return a1 != a0 ? a0 : split + (i - a0);
}

// Marks of the level-1 partial slots, one byte each (two per tile), written 0 by level 1 and 1 by
// combine_kernel for the slots it has summed and applied; the 32-bit word in front of them is the
// "chains left for the tree" flag (reset by level 1, set by combine_kernel).
__host__ __device__ __forceinline__ uint32_t *up_flag(uint16_t *consumed) { return reinterpret_cast<uint32_t *>(consumed) - 1; }

struct __align__(16) EntryDesc {
  const float *gp;  // output-grad row of this lookup (nullptr: past the end of the array)
  float *wp;        // table row to update when this entry ends its run, else nullptr
};

template <int G> struct L1Threads { static constexpr int value = G >= 16 ? 256 : (G >= 4 ? 128 : (G == 2 ? 64 : 32)); };

template <typename K, int G, int C, int MODE, bool AVG, bool PIPE>
__global__ void __launch_bounds__(L1Threads<G>::value, 512 / L1Threads<G>::value)
reduce1_kernel(const __grid_constant__ TableBatch tb, const K *__restrict__ keys,
               const uint32_t *__restrict__ vals, int64_t n, int D, float lr, K *keys_out,
               float *rows_out, int flags, uint32_t split, const OptArgs *oa, uint16_t *consumed) {
  constexpr int THREADS = L1Threads<G>::value;
  constexpr int CH1 = Ch1<G>::value;
  constexpr int SLOTS = 32 / G, WARPS = THREADS / 32, EPW = SLOTS * CH1;
  constexpr int UNR = (C == 1) ? 4 : (C == 2 ? 2 : 1);
  constexpr int NG = CH1 / UNR;
  constexpr K INV = KeyTraits<K>::INV;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  EntryDesc *wd = reinterpret_cast<EntryDesc *>(smem_raw) + warp * EPW;
  float2 *wscale = reinterpret_cast<float2 *>(smem_raw + sizeof(EntryDesc) * WARPS * EPW) + warp * EPW;
  if (blockIdx.x == 0 && threadIdx.x == 0) up_flag(consumed)[0] = 0;  // "chains left for the tree": set by combine_kernel
  const int64_t wbase = ((int64_t)mix_cta(blockIdx.x, gridDim.x, split) * WARPS + warp) * EPW;
  if (wbase >= n) return;  // warp-uniform; only __syncwarp below
  const int D4 = D >> 2;

  // ---- phase 1 ----
#pragma unroll
  for (int i = 0; i < EPW / 32; i++) {
    const int e = i * 32 + lane;
    const int64_t p = wbase + e;
    EntryDesc d;
    d.gp = nullptr;
    d.wp = nullptr;
    if (p < n && (uint64_t)keys[p] < tb.row_base[tb.num_tables]) {  // (an empty slot's key is past every row)
      const K k = keys[p];
      const uint32_t v = vals[p];
      const bool is_end = (p + 1 == n) || (keys[p + 1] != k);
      const uint32_t t = v >> BAG_BITS, b = v & BAG_MASK;
      d.gp = tb.t[t].out + (int64_t)b * tb.og_stride;
      if (is_end) d.wp = tb.t[t].table + (int64_t)((uint64_t)k - tb.row_base[t]) * D;
      if ((flags & 1) && is_end) prefetch_l2_bulk(d.wp, D * 4);
      if (flags & 2) prefetch_l2_bulk(d.gp, D * 4);
      if (AVG) wscale[e] = make_float2((float)tb.t[t].L, __frcp_rn((float)tb.t[t].L));
    }
    wd[e] = d;
  }
  __syncwarp();

  // ---- phase 2 ----
  const int sub = lane / G, sl = lane % G;
  const int64_t tile = wbase / CH1 + sub;
  const int64_t t0 = tile * CH1;
  if (t0 >= n) return;  // this slot's tile is empty (its descriptors are all null)
  if (sl == 0) consumed[tile] = 0;  // (both partial slots of the tile: not yet taken by combine_kernel)
  const EntryDesc *dq = wd + sub * CH1;
  const float2 *sq = wscale + sub * CH1;
  const int cnt = (int)min((int64_t)CH1, n - t0);
  bool first_open = (t0 > 0) && (keys[t0 - 1] == keys[t0]);  // first run continues the previous tile
  bool pending = false, wrote_head = false;
  const uint64_t pol_keep = policy_evict_last(), pol_stream = policy_evict_first();
  float4 acc[C];
#pragma unroll
  for (int c = 0; c < C; c++) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 g[2][UNR][C], w[2][UNR][C];
  float *wp[2][UNR];
  bool ok[2][UNR];

  if constexpr (PIPE) {
    // Gathers that miss L2 (output-grad of one table larger than L2): latency-bound.  The loop
    // body is branch-free (predicated loads/stores + selects) so that ptxas keeps the two
    // register buffers on separate scoreboard barriers and group k+1's loads really are in
    // flight while group k is added (a branch in the body puts every load on one barrier).
    int head_e = 0;
    bool own[C];
#pragma unroll
    for (int c = 0; c < C; c++) own[c] = sl + c * G < D4;
    float *const head_row = rows_out + (2 * tile) * (int64_t)D;
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int u = 0; u < UNR; u++)
#pragma unroll
        for (int c = 0; c < C; c++) g[i][u][c] = w[i][u][c] = make_float4(0.f, 0.f, 0.f, 0.f);
    auto issue = [&](int bf, int grp) {
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const EntryDesc d = dq[min(grp * UNR + u, CH1 - 1)];
        const bool in = grp < NG;
        ok[bf][u] = in && d.gp != nullptr;
        wp[bf][u] = in ? d.wp : nullptr;
#pragma unroll
        for (int c = 0; c < C; c++) {
          ldg_row_hint_if(g[bf][u][c], d.gp + 4 * (sl + c * G), pol_keep, ok[bf][u] && own[c]);
          ld_row_hint_if(w[bf][u][c], d.wp + 4 * (sl + c * G), pol_stream, wp[bf][u] != nullptr && own[c]);
        }
      }
    };
    auto consume = [&](int bf, int grp) {
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const bool on = ok[bf][u];
        const bool end = on && wp[bf][u] != nullptr;
        const bool head = end && first_open;  // head partial: the run began in an earlier tile
        float2 sc = make_float2(1.f, 1.f);
        if (AVG) sc = sq[min(grp * UNR + u, CH1 - 1)];
        float *dst = head ? head_row : wp[bf][u];
#pragma unroll
        for (int c = 0; c < C; c++) {
          float4 x = g[bf][u][c];
          if (AVG) x = f4_div_rn(x, sc.x, sc.y);  // g / in_dim per element, embedding.cu:210
          if (!on) x = make_float4(0.f, 0.f, 0.f, 0.f);
          f4_add(acc[c], x);
          float4 o = w[bf][u][c];
          if (MODE == APPLY_SGD) {  // W -= lr*gt, contracted like the reference's FFMA
            o.x = fmaf(-acc[c].x, lr, o.x); o.y = fmaf(-acc[c].y, lr, o.y);
            o.z = fmaf(-acc[c].z, lr, o.z); o.w = fmaf(-acc[c].w, lr, o.w);
          } else {
            f4_add(o, acc[c]);
          }
          if (head) o = acc[c];
          st_row_hint_if(dst + 4 * (sl + c * G), o, pol_stream, end && own[c]);
          if (end) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        head_e = head ? grp * UNR + u : head_e;
        wrote_head = wrote_head || head;
        pending = end ? false : (pending || on);
        first_open = first_open && !end;
      }
    };
    issue(0, 0);
#pragma unroll 1
    for (int grp = 0; grp < NG; grp += 2) {
      issue(1, grp + 1);
      consume(0, grp);
      issue(0, grp + 2);  // predicated off past the tile
      consume(1, grp + 1);
      if ((grp + 2) * UNR >= cnt) break;
    }
    if (sl == 0) keys_out[2 * tile] = wrote_head ? keys[t0 + head_e] : INV;
  } else {
    // Gathers that hit L2 (the usual case: one table's output-grad is a few tens of MB): the
    // walk is instruction-bound, so the run-end handling is a branch and the common entry costs
    // four FADDs; loads are still issued a group ahead.
    auto issue = [&](auto BF, int grp) {
      constexpr int bf = decltype(BF)::value;
  #pragma unroll
      for (int u = 0; u < UNR; u++) {
        const EntryDesc d = dq[grp * UNR + u];
        ok[bf][u] = d.gp != nullptr;
        wp[bf][u] = d.wp;
        if (d.gp) {
  #pragma unroll
          for (int c = 0; c < C; c++)
            if (sl + c * G < D4) g[bf][u][c] = ldg_row_hint(d.gp + 4 * (sl + c * G), pol_keep);
        }
        if (d.wp) {
  #pragma unroll
          for (int c = 0; c < C; c++)
            if (sl + c * G < D4) w[bf][u][c] = ld_row_hint(d.wp + 4 * (sl + c * G), pol_stream);
        }
      }
    };
    auto consume = [&](auto BF, int grp) {
      constexpr int bf = decltype(BF)::value;
  #pragma unroll
      for (int u = 0; u < UNR; u++) {
        if (!ok[bf][u]) continue;
  #pragma unroll
        for (int c = 0; c < C; c++) {
          if (sl + c * G < D4) {
            float4 x = g[bf][u][c];
            if (AVG) x = f4_div_rn(x, sq[grp * UNR + u].x, sq[grp * UNR + u].y);  // g / in_dim per element, embedding.cu:210
            f4_add(acc[c], x);
          }
        }
        pending = true;
        if (wp[bf][u]) {
          if (first_open) {  // head partial: the run began in an earlier tile
            if (sl == 0) keys_out[2 * tile] = keys[t0 + grp * UNR + u];
  #pragma unroll
            for (int c = 0; c < C; c++)
              if (sl + c * G < D4) st_f4(rows_out + (2 * tile) * (int64_t)D + 4 * (sl + c * G), acc[c]);
            wrote_head = true;
          } else {  // complete run: the one read-modify-write of this table row
  #pragma unroll
            for (int c = 0; c < C; c++) {
              if (sl + c * G < D4) {
                float4 o = w[bf][u][c];
                if (MODE == APPLY_SGD) {  // W -= lr*gt, contracted like the reference's FFMA
                  o.x = fmaf(-acc[c].x, lr, o.x); o.y = fmaf(-acc[c].y, lr, o.y);
                  o.z = fmaf(-acc[c].z, lr, o.z); o.w = fmaf(-acc[c].w, lr, o.w);
                } else if (MODE == APPLY_SGDX) {
                  const uint64_t key = (uint64_t)keys[t0 + grp * UNR + u];
                  const int t = find_table(tb, key);
                  o = sgdx_fragment(oa, t, tb.t[t].table, wp[bf][u] + 4 * (sl + c * G), o, acc[c],
                                    key - tb.row_base[t], sl == 0 && c == 0);
                } else {
                  f4_add(o, acc[c]);
                }
                st_row_hint(wp[bf][u] + 4 * (sl + c * G), o, pol_stream);
              }
            }
          }
          first_open = false;
          pending = false;
  #pragma unroll
          for (int c = 0; c < C; c++) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    };
    using B0 = std::integral_constant<int, 0>;
    using B1 = std::integral_constant<int, 1>;
    issue(B0{}, 0);
  #pragma unroll 1
    for (int grp = 0; grp < NG; grp += 2) {
      issue(B1{}, grp + 1);
      consume(B0{}, grp);
      if (grp + 2 < NG && (grp + 2) * UNR < cnt) issue(B0{}, grp + 2);
      else {
  #pragma unroll
        for (int u = 0; u < UNR; u++) ok[0][u] = false;
      }
      consume(B1{}, grp + 1);
      if ((grp + 2) * UNR >= cnt) break;
    }
    if (sl == 0 && !wrote_head) keys_out[2 * tile] = INV;
  }
  if (pending) {  // tail partial: the run goes on in a later tile
#pragma unroll
    for (int c = 0; c < C; c++)
      if (sl + c * G < D4) st_f4(rows_out + (2 * tile + 1) * (int64_t)D + 4 * (sl + c * G), acc[c]);
  }
  if (sl == 0) keys_out[2 * tile + 1] = pending ? keys[t0 + cnt - 1] : INV;
}

}  // namespace ffemb
#include "ffemb_reduce1_ring.cuh"
namespace ffemb {


// ---- everything above level 1 in ONE launch, for runs cut by a few tile edges (the usual case) ----
// Level 1 leaves, per tile s, a head partial (slot 2s: the tile's first run began earlier and ends
// here) and a tail partial (slot 2s+1: its last run goes on).  A run cut by tile edges is therefore a
// CHAIN  tail(s), tail(s+1) ... tail(e-1), head(e)  — tiles s+1 .. e-1 lie inside the run — and chains
// of different runs are independent.  One slot of G lanes per tile s: if tail(s) is the FIRST piece of
// its chain (tail(s-1) has another key) the slot finds the chain's end (its G lanes look at G tiles per
// round), adds the pieces in order and does the run's one read-modify-write of the table row — all
// chains in parallel, no serial walk over tiles of partial slots, no further levels.  The pieces it
// used are marked in `consumed`.  A chain of more than kMaxChain pieces (one row with thousands of
// lookups: skewed indices) is left to the tree below (reduce_level / reduce_tail, which read marked
// slots as invalid) and the "chains left" flag is raised; when it stays 0 the tree kernels return at
// once.  The order of the adds depends on the sorted order and CH1 only: bit-reproducible.
static constexpr int kMaxChain = 128;

template <typename K, int G, int C, int MODE>
__global__ void __launch_bounds__(kThreads)
combine_kernel(const __grid_constant__ TableBatch tb, const K *__restrict__ keys,
               const float *__restrict__ rows, uint16_t *consumed16, int64_t ntiles, int D, float lr,
               const OptArgs *oa) {
  constexpr K INV = KeyTraits<K>::INV;
  constexpr int SLOTS = 32 / G;
  constexpr int PER = G >= 8 ? 1 : 8 / G;           // tiles a lane probes per round (narrow slots: 8 per round)
  constexpr int ROUNDS = kMaxChain / (G * PER) > 0 ? kMaxChain / (G * PER) : 1;
  uint8_t *consumed = reinterpret_cast<uint8_t *>(consumed16);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int D4 = D >> 2;
  const int64_t nwarps = (int64_t)gridDim.x * (kThreads / 32);
  // a warp scans 32 tiles at a time for chain starts (coalesced key reads; most tiles start none) ...
  for (int64_t base = ((int64_t)blockIdx.x * (kThreads / 32) + warp) * 32; base < ntiles; base += nwarps * 32) {
    const int64_t sc = base + lane;
    K kc = INV;
    if (sc < ntiles) kc = keys[2 * sc + 1];
    // (a tail partial the hot-row path has already used — the ragged head of a hot run — starts nothing)
    const bool is_start = kc != INV && !consumed[2 * sc + 1] && (sc == 0 || keys[2 * sc - 1] != kc);
    uint32_t pending = __ballot_sync(0xffffffffu, is_start);
    // ... and its slots of G lanes take the starts SLOTS at a time
    while (pending) {  // warp-uniform
      const int pick = (int)__fns(pending, 0, sub + 1);       // the (sub+1)-th start, 0xffffffff: none left for this slot
      const bool start = pick >= 0 && pick < 32;
      const int src = start ? pick : 0;
      const int64_t s = base + src;
      const K kt = shfl_key<K>(kc, src);
#pragma unroll
      for (int i = 0; i < SLOTS; i++) pending &= pending - 1;  // the SLOTS lowest starts are taken
      // 1. where does the chain end?  (keys only: a chain that is too long costs a few probes, no row loads)
      bool done = !start, ok = false;
      int64_t t = s + 1;      // next tile of the chain
      int conts = 0;          // tiles lying inside the run (their tail partial is a piece)
#pragma unroll 1
      for (int r = 0; r < ROUNDS; r++) {  // uniform trip count: the ballots need every lane
#pragma unroll
        for (int pp = 0; pp < PER; pp++) {
          const int64_t tt = t + pp * G + sl;
          K kh = INV, kn = INV;
          if (!done && tt < ntiles) {
            kh = keys[2 * tt];
            kn = keys[2 * tt + 1];
          }
          constexpr uint32_t GMASK = G == 32 ? 0xffffffffu : ((1u << G) - 1u);
          const uint32_t endm = (__ballot_sync(0xffffffffu, !done && kh == kt) >> (sub * G)) & GMASK;
          const uint32_t conm = (__ballot_sync(0xffffffffu, !done && kh != kt && kn == kt) >> (sub * G)) & GMASK;
          if (!done) {
            const uint32_t stop = ~conm & GMASK;             // first tile of this probe that is not inside the run
            const int nc = stop ? __ffs(stop) - 1 : G;
            conts += nc;
            if (nc < G) {
              done = true;
              ok = ((endm >> nc) & 1u) != 0;                 // the chain's last piece: the head partial of that tile
            }
          }
        }
        t += (int64_t)PER * G;
        if (__ballot_sync(0xffffffffu, !done) == 0) break;   // (warp-uniform)
      }
      if (!start) continue;
      if (!ok) {  // longer than kMaxChain pieces: the tree's
        if (sl == 0) up_flag(consumed16)[0] = 1u;
        continue;
      }
  // 2. the pieces in order: tail(s), tail(s+1 .. s+conts), head(s+conts+1); loads UB at a time, adds in order
  constexpr int UB = C == 1 ? 8 : (C == 2 ? 4 : 2);
  float4 acc[C], w[C];
  const int tbl = find_table(tb, (uint64_t)kt);
  float *wp = tb.t[tbl].table + (int64_t)((uint64_t)kt - tb.row_base[tbl]) * D;
#pragma unroll
  for (int c = 0; c < C; c++) {
    if (sl + c * G < D4) {
      w[c] = ld_f4(wp + 4 * (sl + c * G));  // (in flight while the pieces are added)
      acc[c] = ld_f4(rows + (2 * s + 1) * (int64_t)D + 4 * (sl + c * G));
    }
  }
  const int npieces = conts + 1;  // after tail(s)
#pragma unroll 1
  for (int i0 = 0; i0 < npieces; i0 += UB) {
    float4 v[UB][C];
#pragma unroll
    for (int u = 0; u < UB; u++) {
      const int i = i0 + u;
      if (i < npieces) {
        const int64_t q = i < conts ? 2 * (s + 1 + i) + 1 : 2 * (s + 1 + conts);
#pragma unroll
        for (int c = 0; c < C; c++)
          if (sl + c * G < D4) v[u][c] = ld_f4(rows + q * (int64_t)D + 4 * (sl + c * G));
      }
    }
#pragma unroll
    for (int u = 0; u < UB; u++) {
      if (i0 + u < npieces) {
#pragma unroll
        for (int c = 0; c < C; c++)
          if (sl + c * G < D4) f4_add(acc[c], v[u][c]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < C; c++) {
    if (sl + c * G < D4) {
      float4 o = w[c];
      if (MODE == APPLY_SGD) {  // W -= lr*gt, contracted like the reference's FFMA
        o.x = fmaf(-acc[c].x, lr, o.x); o.y = fmaf(-acc[c].y, lr, o.y);
        o.z = fmaf(-acc[c].z, lr, o.z); o.w = fmaf(-acc[c].w, lr, o.w);
      } else if (MODE == APPLY_SGDX) {
        o = sgdx_fragment(oa, tbl, tb.t[tbl].table, wp + 4 * (sl + c * G), o, acc[c],
                          (uint64_t)kt - tb.row_base[tbl], sl == 0 && c == 0);
      } else {
        f4_add(o, acc[c]);
      }
      st_f4(wp + 4 * (sl + c * G), o);
    }
  }
  // the pieces are spent: tail(s), tail(s+1 .. s+conts), head(s+conts+1)
  for (int i = sl; i < conts; i += G) consumed[2 * (s + 1 + i) + 1] = 1;
  if (sl == 0) {
    consumed[2 * s + 1] = 1;
    consumed[2 * (s + 1 + conts)] = 1;
  }
    }
  }
}

template <typename K, int G, int C, int MODE>
__global__ void __launch_bounds__(kThreads)
reduce_level_kernel(const __grid_constant__ TableBatch tb, const K *keys,
                    const float *rows_in, int64_t n, int CH, int D, float lr,
                    K *keys_out, float *rows_out, const OptArgs *oa, const uint32_t *flag,
                    const uint8_t *consumed) {
  if (*flag == 0) return;  // combine_kernel left no chain for the tree (the usual case)
  constexpr int SLOTS = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int64_t tile = ((int64_t)blockIdx.x * (kThreads / 32) + warp) * SLOTS + sub;
  const int64_t ntiles = (n + CH - 1) / CH;
  reduce_tile<K, G, C, MODE>(tb, keys, rows_in, n, tile, CH, tile < ntiles, D, lr, keys_out, rows_out,
                             sub, sl, oa, consumed);
}

// All remaining levels in ONE CTA: level i reads buffer (i&1), writes buffer ((i+1)&1).
template <typename K, int G, int C, int MODE>
__global__ void __launch_bounds__(kThreads)
reduce_tail_kernel(const __grid_constant__ TableBatch tb, K *keys_a, float *rows_a, K *keys_b,
                   float *rows_b, int64_t n, int CH, int D, float lr, const OptArgs *oa, const uint32_t *flag,
                   const uint8_t *consumed) {
  if (*flag == 0) return;  // combine_kernel left no chain for the tree (the usual case)
  constexpr int SLOTS = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int nslots = (kThreads / 32) * SLOTS;
  const int slot = warp * SLOTS + sub;
  K *kin = keys_a, *kout = keys_b;
  float *rin = rows_a, *rout = rows_b;
  while (true) {
    const int64_t ntiles = (n + CH - 1) / CH;
    for (int64_t base = 0; base < ntiles; base += nslots) {
      const int64_t tile = base + slot;
      reduce_tile<K, G, C, MODE>(tb, kin, rin, n, tile, CH, tile < ntiles, D, lr, kout, rout, sub, sl, oa, consumed);
    }
    if (ntiles <= 1) break;
    __syncthreads();  // partial rows written by this CTA are visible to it after the barrier
    consumed = nullptr;  // (the marks belong to the level-1 partials only)
    n = 2 * ntiles;
    K *tk = kin; kin = kout; kout = tk;
    float *tr = rin; rin = rout; rout = tr;
  }
}


// ---- AVG / MEAN: the gradient of a bag's rows is output-grad / in_dim (embedding.cu:206-211) --------
// One IEEE division per output-grad element, ONCE per bag, into scratch rows [B][D]; level 1 then adds
// the quotients like plain sums.  Same bits as dividing per lookup (the quotient does not depend on the
// lookup), 1/L of the divisions.  Descriptor reuse: t[].out = source (row stride `ogs`), t[].table =
// destination (row stride D), t[].L = divisor.
static __global__ void __launch_bounds__(256)
prescale_kernel(const __grid_constant__ TableBatch tp, int64_t B, int D, int64_t ogs) {
  const TableDesc &td = tp.t[blockIdx.y];
  const int D4 = D >> 2;
  const float Lf = (float)td.L;
  const int64_t total = B * (int64_t)D4;
  const uint64_t pol_keep = policy_evict_last(), pol_stream = policy_evict_first();
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t b = i / D4;
    const int c4 = (int)(i - b * D4);
    float4 v = ldg_row_hint(td.out + b * ogs + 4 * c4, pol_stream);
    v.x = __fdiv_rn(v.x, Lf); v.y = __fdiv_rn(v.y, Lf); v.z = __fdiv_rn(v.z, Lf); v.w = __fdiv_rn(v.w, Lf);
    st_row_hint(td.table + b * (int64_t)D + 4 * c4, v, pol_keep);  // re-read L times by level 1
  }
}

// ---- workspace carving -----------------------------------------------------------------------
static inline size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }
static inline bool wide_keys(int64_t total_rows) { return (uint64_t)total_rows >= 0xffffffffull; }

// ---- hot rows ---------------------------------------------------------------------------------------
static constexpr int64_t kHotMinLookups = 1 << 22;  // calls below this many sorted lookups never look for hot rows
static constexpr int kHotMin = 8192;                // a run of this many sorted lookups is "hot"
static constexpr int kHotStep = kHotMin / 2;        // the sorted keys are probed every kHotStep entries
static constexpr int kHotCtas = 64;                 // CTAs (bag ranges) per table in hot_sum_kernel
struct HotRun {
  uint32_t lo, hi;      // the run's WHOLE tiles: sorted positions [lo, hi), multiples of CH1
  uint32_t start, end;  // the run itself: [start, end)
  uint64_t key;         // global row id
  int32_t table;
  int32_t pad;
};


// Hot rows.  Under skew (BASELINE config 5: Zipf 1.05, bag 64) a few hundred rows per table hold half of
// all lookups; through level 1 each of those lookups is a 4*D-byte gather of an output-grad row from L2
// although the SAME few thousand output-grad rows are asked for again and again.  A run of >= kHotMin
// sorted lookups therefore leaves level 1:
//   hot_select : the sorted keys are probed every kHotStep entries; the first probe inside a run finds
//                the run's ends by binary search; runs of >= kHotMin entries go on a list.
//   hot_mark   : the run's WHOLE level-1 tiles are marked — level 1 skips them (and writes no partials);
//                the run's ragged head and tail stay ordinary level-1 entries and arrive as the tail
//                partial of the tile before / the head partial of the tile after.
//   hot_sum    : BAG-major.  A CTA owns a range of bags of one table, stages the output-grad rows of 64 KB
//                slices of it in shared memory (one TMA bulk copy per slice, double-buffered) and every
//                warp adds, for the runs it owns, the slice's entries — a run's entries are bag-ascending,
//                so a cursor per run walks them slice by slice — into REGISTER accumulators: an
//                output-grad row is read from L2 once per CTA instead of once per lookup.
//   hot_apply  : per run, head fragment + the CTAs' partial sums in bag order + tail fragment, then the
//                run's one read-modify-write of the table row; the two fragments are marked consumed.
// Every order is fixed by the sorted order and the shapes: bit-reproducible.  The set of hot runs is
// "every run of >= kHotMin entries" — never capped, so it does not depend on the list's (atomic) order.
__host__ __device__ __forceinline__ int64_t hot_base(const TableBatch &tb, int t) {
  return (int64_t)(tb.t[t].lookup_base / kHotMin) + t;
}

template <typename K>
__global__ void __launch_bounds__(256)
hot_select_kernel(const __grid_constant__ TableBatch tb, const K *__restrict__ keys, int64_t n, int CH1,
                  HotRun *list, uint32_t *cnt, int64_t cap) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const int64_t q = i * kHotStep;
  if (q >= n) return;
  const K k = keys[q];
  if ((uint64_t)k >= tb.row_base[tb.num_tables]) return;           // empty slots
  if (i > 0 && keys[q - kHotStep] == k) return;                    // not the first probe inside its run
  int64_t lo = i > 0 ? q - kHotStep + 1 : 0, hi = q;               // first position of k in (q - step, q]
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (keys[mid] < k) lo = mid + 1; else hi = mid;
  }
  const int64_t start = lo;
  lo = q + 1; hi = n;                                               // first position after k in (q, n]
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (keys[mid] <= k) lo = mid + 1; else hi = mid;
  }
  const int64_t end = lo;
  if (end - start < kHotMin) return;
  const int64_t tlo = (start + CH1 - 1) / CH1 * CH1, thi = end / CH1 * CH1;
  if (thi <= tlo) return;
  // the list is cut into one segment per table (a table of n_t lookups has at most n_t / kHotMin hot runs),
  // so that hot_sum finds a table's runs without searching; the order inside a segment is arbitrary
  const int t = find_table(tb, (uint64_t)k);
  const int64_t slot = hot_base(tb, t) + atomicAdd(cnt + t, 1u);
  if (slot >= cap) return;  // (cannot happen)
  HotRun r;
  r.lo = (uint32_t)tlo; r.hi = (uint32_t)thi; r.start = (uint32_t)start; r.end = (uint32_t)end;
  r.key = (uint64_t)k;
  r.table = t;
  r.pad = 0;
  list[slot] = r;
}

static __global__ void __launch_bounds__(256)
hot_mark_kernel(const __grid_constant__ TableBatch tb, const HotRun *__restrict__ list,
                const uint32_t *__restrict__ cnt, int CH1, uint8_t *marks) {
  const int tab = blockIdx.y;
  const uint32_t nrun = cnt[tab];
  for (uint32_t h = blockIdx.x; h < nrun; h += gridDim.x) {
    const HotRun r = list[hot_base(tb, tab) + h];
    for (int64_t t = r.lo / CH1 + threadIdx.x; t < r.hi / CH1; t += 256) marks[t] = 1;
  }
}

static constexpr int kHotWarps = 8, kHotKW = 32;         // runs per pass and CTA: kHotWarps * kHotKW (lane i of a warp keeps run i's cursor)
static constexpr int kHotSliceBytes = 64 << 10;          // output-grad rows staged per slice

static __global__ void __launch_bounds__(kHotWarps * 32, 1)
hot_sum_kernel(const __grid_constant__ TableBatch tb, const uint32_t *__restrict__ vals, int64_t B, int D,
               const HotRun *__restrict__ list, const uint32_t *__restrict__ cnt, float *__restrict__ partials) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float *slice[2] = {reinterpret_cast<float *>(smem_raw), reinterpret_cast<float *>(smem_raw + kHotSliceBytes)};
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + 2 * kHotSliceBytes);
  const int t = blockIdx.y, c = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nm = (int)cnt[t];
  if (nm == 0) return;
  const int64_t base = hot_base(tb, t);
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(bar + 1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const TableDesc &td = tb.t[t];
  const int64_t per = (B + gridDim.x - 1) / gridDim.x;
  const int64_t bag0 = (int64_t)c * per, bag1 = min(B, bag0 + per);
  const int SB = kHotSliceBytes / (D * 4);                 // bags per slice
  const int D4 = D >> 2;
  const bool own = lane < D4;
  uint32_t phases = 0u;  // bit b: parity of the phase buffer b's barrier completes next
  for (int g0 = 0; g0 < nm; g0 += kHotWarps * kHotKW) {   // passes of 256 runs (one for a few hundred hot rows)
    // this warp's runs: lane i keeps the cursor / limit / list slot of run g0 + warp*KW + i
    uint32_t cur = 0, lim = 0, slot = 0xffffffffu;
    if (lane < kHotKW && g0 + warp * kHotKW + lane < nm) {
      slot = (uint32_t)(base + g0 + warp * kHotKW + lane);
      const HotRun r = list[slot];
      auto first_bag_ge = [&](int64_t bag) {               // entries of a run are bag-ascending
        uint32_t lo = r.lo, hi = r.hi;
        while (lo < hi) {
          const uint32_t mid = (lo + hi) >> 1;
          if ((int64_t)(vals[mid] & BAG_MASK) < bag) lo = mid + 1; else hi = mid;
        }
        return lo;
      };
      cur = first_bag_ge(bag0);
      lim = first_bag_ge(bag1);
    }
    float4 acc[kHotKW];
#pragma unroll
    for (int i = 0; i < kHotKW; i++) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int nslices = bag1 > bag0 ? (int)((bag1 - bag0 + SB - 1) / SB) : 0;
    auto load = [&](int sidx) {                            // one bulk copy per slice (og_stride == D: contiguous)
      if (threadIdx.x == 0 && sidx < nslices) {
        const int64_t sb0 = bag0 + (int64_t)sidx * SB;
        const uint32_t bytes = (uint32_t)(min((int64_t)SB, bag1 - sb0) * D * 4);
        mbar_expect_tx(bar + (sidx & 1), bytes);
        bulk_g2s(slice[sidx & 1], td.out + sb0 * (int64_t)D, bytes, bar + (sidx & 1));
      }
    };
    load(0);
    for (int sidx = 0; sidx < nslices; sidx++) {
      load(sidx + 1);                                      // (its buffer was released by the barrier below)
      const int bf = sidx & 1;
      while (!mbar_try_wait(bar + bf, (phases >> bf) & 1u)) {}
      phases ^= 1u << bf;
      const int64_t sb0 = bag0 + (int64_t)sidx * SB, sb1 = min(bag1, sb0 + SB);
      const float *gs = slice[bf];
      // the first 32 entries of every run of this warp at its cursor, all loads in flight together (one
      // L2 round trip per slice instead of one per run)
      const int nsb = (int)(sb1 - sb0);                     // bags of this slice
      int first[kHotKW];                                   // bag - sb0 (>= nsb: not in this slice)
#pragma unroll
      for (int i = 0; i < kHotKW; i++) {
        const uint32_t cu = __shfl_sync(0xffffffffu, cur, i), li = __shfl_sync(0xffffffffu, lim, i);
        first[i] = cu + lane < li ? (int)((int64_t)(vals[cu + lane] & BAG_MASK) - sb0) : nsb;
      }
#pragma unroll
      for (int i = 0; i < kHotKW; i++) {
        uint32_t cu = __shfl_sync(0xffffffffu, cur, i);
        const uint32_t li = __shfl_sync(0xffffffffu, lim, i);
        // the slice's entries of the run are summed on their own and then added to the run's accumulator:
        // blocks of a few hundred addends, like the tiles of level 1 (one long serial fp32 sum of 10^4..10^5
        // addends loses two digits the tiled sum keeps)
        float4 part = make_float4(0.f, 0.f, 0.f, 0.f);
        auto add_entries = [&](int rel, int nin) {        // the first nin lanes hold bags of this slice
          for (int e0 = 0; e0 < nin; e0 += 8) {            // 8 shared-memory loads in flight, adds in entry order
            float4 x[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
              const int b = __shfl_sync(0xffffffffu, rel, (e0 + u) & 31);
              x[u] = make_float4(0.f, 0.f, 0.f, 0.f);      // (past the slice's entries: + 0 changes nothing)
              if (own && e0 + u < nin) x[u] = *reinterpret_cast<const float4 *>(gs + (size_t)b * D + 4 * lane);
            }
#pragma unroll
            for (int u = 0; u < 8; u++) f4_add(part, x[u]);
          }
        };
        auto ld_rel = [&](uint32_t p) {                    // bag - sb0 of entry p + lane (nsb: not this slice's)
          return p + lane < li ? (int)((int64_t)(vals[p + lane] & BAG_MASK) - sb0) : nsb;
        };
        int rel = first[i];
        while (cu < li) {                                  // warp-uniform
          const int nin = __popc(__ballot_sync(0xffffffffu, rel < nsb));   // a prefix: bags ascend
          add_entries(rel, nin);
          cu += nin;
          if (nin < 32) break;
          rel = ld_rel(cu);
        }
        if (own) f4_add(acc[i], part);
        if (lane == i) cur = cu;
      }
      __syncthreads();                                     // every warp is done with slice[bf]
    }
#pragma unroll
    for (int i = 0; i < kHotKW; i++) {
      const uint32_t sl_i = __shfl_sync(0xffffffffu, slot, i);
      if (sl_i != 0xffffffffu && own)
        st_f4(partials + ((size_t)sl_i * kHotCtas + c) * D + 4 * lane, acc[i]);
    }
  }
}

template <typename K, int MODE>
__global__ void __launch_bounds__(256)
hot_apply_kernel(const __grid_constant__ TableBatch tb, const HotRun *__restrict__ list,
                 const uint32_t *__restrict__ cnt, int CH1, int D, float lr,
                 const float *__restrict__ partials, const float *__restrict__ prows, uint16_t *consumed16) {
  uint8_t *consumed = reinterpret_cast<uint8_t *>(consumed16);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tab = blockIdx.y;
  const uint32_t nrun = cnt[tab];
  const int D4 = D >> 2;
  const bool own = lane < D4;
  for (uint32_t hh = blockIdx.x * 8 + warp; hh < nrun; hh += gridDim.x * 8) {
    const int64_t h = hot_base(tb, tab) + hh;
    const HotRun r = list[h];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool has_head = r.start < r.lo, has_tail = r.hi < r.end;
    const int64_t th = r.start / CH1, tt = r.hi / CH1;    // tiles of the ragged head / tail
    if (has_head && own) acc = ld_f4(prows + (2 * th + 1) * (int64_t)D + 4 * lane);    // its tail partial
    for (int cc = 0; cc < kHotCtas; cc++)
      if (own) f4_add(acc, ld_f4(partials + ((size_t)h * kHotCtas + cc) * D + 4 * lane));
    if (has_tail && own) f4_add(acc, ld_f4(prows + (2 * tt) * (int64_t)D + 4 * lane));  // its head partial
    float *wp = tb.t[r.table].table + (int64_t)(r.key - tb.row_base[r.table]) * D;
    if (own) {
      float4 o = ld_f4(wp + 4 * lane);
      if (MODE == APPLY_SGD) {  // W -= lr*gt, contracted like the reference's FFMA
        o.x = fmaf(-acc.x, lr, o.x); o.y = fmaf(-acc.y, lr, o.y);
        o.z = fmaf(-acc.z, lr, o.z); o.w = fmaf(-acc.w, lr, o.w);
      } else {
        f4_add(o, acc);
      }
      st_f4(wp + 4 * lane, o);
    }
    if (lane == 0) {
      if (has_head) consumed[2 * th + 1] = 1;
      if (has_tail) consumed[2 * tt] = 1;
    }
  }
}

struct Levels {
  int count;            // number of partial buffers
  int64_t slots[8];     // partial slots of level i output (2 * tiles)
};
static Levels plan_levels(int64_t n, int CH1) {
  Levels lv;
  lv.count = 0;
  int64_t m = 2 * ((n + CH1 - 1) / CH1);
  lv.slots[lv.count++] = m;
  while (m > 2 && lv.count < 8) {  // more than one tile at the previous level
    const int64_t tiles = (m + CH2 - 1) / CH2;
    m = 2 * tiles;
    lv.slots[lv.count++] = m;
    if (tiles <= 1) break;
  }
  return lv;
}

struct Carve {
  char *keys_a, *keys_b;
  uint32_t *vals_a, *vals_b;
  void *cub_tmp;
  size_t cub_bytes;
  char *pkeys[2];
  float *prows[2];
  uint16_t *consumed;   // marks of the level-1 partial slots (one byte each), behind the 32-bit "chains left" flag
  // hot rows (skewed indices): runs of >= kHotMin sorted lookups leave level 1, see hot_select_kernel
  uint32_t *hot_cnt;    // number of hot runs found this step
  HotRun *hot_list;
  uint8_t *hot_marks;   // one byte per level-1 tile: 1 = the tile lies inside a hot run (level 1 skips it)
  float *hot_partials;  // [run][kHotCtas][D]
  int64_t hot_cap;      // capacity of hot_list (0: no hot path at this size)
  size_t total;
};
static Carve carve(char *base, int64_t n, int64_t total_rows, int64_t D) {
  Carve c;
  const size_t ksz = wide_keys(total_rows) ? 8 : 4;
  size_t off = 0;
  auto take = [&](size_t bytes) { char *p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  c.keys_a = take((size_t)n * ksz);
  c.keys_b = take((size_t)n * ksz);
  c.vals_a = (uint32_t *)take((size_t)n * 4);
  c.vals_b = (uint32_t *)take((size_t)n * 4);
  size_t tmp = 0;
  if (ksz == 4)
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, (uint32_t *)nullptr, (uint32_t *)nullptr,
                                    (uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 32);
  else
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, (uint64_t *)nullptr, (uint64_t *)nullptr,
                                    (uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 64);
  c.cub_bytes = tmp;
  c.cub_tmp = take(tmp);
  // partial buffers ping-pong: level 1 output is the largest; level 2 output the second largest
  const Levels lv = plan_levels(n, ch1_of(D));
  const int64_t s0 = lv.slots[0], s1 = lv.count > 1 ? lv.slots[1] : 2;
  c.pkeys[0] = take((size_t)s0 * ksz);
  c.prows[0] = (float *)take((size_t)s0 * D * 4);
  c.pkeys[1] = take((size_t)s1 * ksz);
  c.prows[1] = (float *)take((size_t)s1 * D * 4);
  char *marks = take(16 + (size_t)s0);
  c.consumed = marks ? (uint16_t *)(marks + 4) : nullptr;
  c.hot_cap = tuning().hot > 0 && n >= kHotMinLookups && D <= 128 ? n / kHotMin + FFEMB_MAX_TABLES + 2 : 0;  // (opt-in)
  c.hot_cnt = (uint32_t *)take(c.hot_cap ? 256 : 0);
  c.hot_list = (HotRun *)take((size_t)c.hot_cap * sizeof(HotRun));
  c.hot_marks = (uint8_t *)take(c.hot_cap ? (size_t)(s0 / 2) : 0);
  c.hot_partials = (float *)take((size_t)c.hot_cap * kHotCtas * D * 4);
  c.total = off;
  return c;
}

#ifndef FFEMB_BACKWARD_WIDE_TU  // (the 64-bit-key instantiations are a second translation unit, see the end)
static constexpr size_t kOptBytes = (sizeof(OptArgs) + 255) & ~(size_t)255;  // device copy of OptArgs
size_t backward_workspace_bytes(int64_t num_lookups, int64_t total_rows, int64_t D) {
  if (num_lookups <= 0 || total_rows <= 0 || D <= 0 || num_lookups >= (1ll << 31)) return 0;
  return carve(nullptr, num_lookups, total_rows, D).total + small_partials_bound(D) + kOptBytes;
}

// ---- APPLY_SGDX, second half: the rows the step did NOT touch ---------------------------------
// sgd_update runs over every element of the table (optimizer_kernel.cu:23-41); with a zero gradient
// that is gt = wd*W (+ momentum).  One warp per 32 rows = one word of the touched bitmap: rows whose
// bit is set were updated by the sparse pass and are skipped; the word is cleared for the next step.
// Streaming read-modify-write of W (and V): 8 (16) bytes per element, against the reference's 4 (zero
// fill) + 12 (20) (sgd_update) + the atomics — and no dense gradient exists.
template <int G>
__global__ void __launch_bounds__(256)
sgdx_untouched_kernel(float *__restrict__ W, float *__restrict__ V, float *__restrict__ M,
                      uint32_t *__restrict__ touched, int64_t R, int D, float lr, float wd, float mom,
                      int nesterov, int adam, float alpha_t, float beta1, float beta2, float eps) {
  constexpr int SLOTS = 32 / G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int D4 = D >> 2;
  const uint64_t pol = policy_evict_first();
  const int64_t nwords = (R + 31) >> 5;
  for (int64_t w0 = (int64_t)blockIdx.x * 8 + warp; w0 < nwords; w0 += (int64_t)gridDim.x * 8) {
    const uint32_t bits = touched[w0];
    __syncwarp();
    if (lane == 0 && bits) touched[w0] = 0u;
    for (int r0 = 0; r0 < 32; r0 += SLOTS) {
      const int64_t row = w0 * 32 + r0 + sub;
      if (row >= R || ((bits >> (r0 + sub)) & 1u)) continue;
      for (int c4 = sl; c4 < D4; c4 += G) {
        float *wp = W + row * (int64_t)D + 4 * c4;
        float4 w = ld_row_hint(wp, pol), v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (adam) {  // adam_update with a zero gradient: M and V decay, W moves by alpha_t*M/(sqrt(V)+eps)
          float *mp = M + row * (int64_t)D + 4 * c4, *vq = V + row * (int64_t)D + 4 * c4;
          float4 m = ld_row_hint(mp, pol), o;
          v = ld_row_hint(vq, pol);
          o.x = adam_elem(w.x, 0.0f, m.x, v.x, alpha_t, beta1, beta2, wd, eps);
          o.y = adam_elem(w.y, 0.0f, m.y, v.y, alpha_t, beta1, beta2, wd, eps);
          o.z = adam_elem(w.z, 0.0f, m.z, v.z, alpha_t, beta1, beta2, wd, eps);
          o.w = adam_elem(w.w, 0.0f, m.w, v.w, alpha_t, beta1, beta2, wd, eps);
          st_row_hint(wp, o, pol);
          st_row_hint(mp, m, pol);
          st_row_hint(vq, v, pol);
          continue;
        }
        float *vp = nullptr;
        if (mom > 0.0f) {
          vp = V + row * (int64_t)D + 4 * c4;
          v = ld_row_hint(vp, pol);
        }
        float4 o;
        o.x = sgdx_elem(w.x, 0.0f, v.x, lr, wd, mom, nesterov != 0);
        o.y = sgdx_elem(w.y, 0.0f, v.y, lr, wd, mom, nesterov != 0);
        o.z = sgdx_elem(w.z, 0.0f, v.z, lr, wd, mom, nesterov != 0);
        o.w = sgdx_elem(w.w, 0.0f, v.w, lr, wd, mom, nesterov != 0);
        st_row_hint(wp, o, pol);
        if (vp) st_row_hint(vp, v, pol);
      }
    }
  }
}

static int launch_sgdx_untouched(const TableBatch &tb, int D, const OptArgs &o, cudaStream_t s) {
  int launches = 0;
  for (int t = 0; t < tb.num_tables; t++) launches += tb.t[t].R > 0;
  if (!launches) return FFEMB_OK;
  ProfScope prof(FFEMB_PHASE_DENSE, launches, s);
  const int D4 = D / 4;
  for (int t = 0; t < tb.num_tables; t++) {
    const int64_t R = tb.t[t].R;
    if (R <= 0) continue;
    const int64_t nwords = (R + 31) / 32;
    const unsigned grid = (unsigned)((nwords + 7) / 8 < 148 * 8 ? (nwords + 7) / 8 : 148 * 8);
    auto go = [&](auto kern) {
      kern<<<grid, 256, 0, s>>>(tb.t[t].table, o.V[t], o.M[t], o.touched[t], R, D, o.lr, o.wd, o.mom, o.nesterov,
                                o.adam, o.alpha_t, o.beta1, o.beta2, o.eps);
    };
    if (D4 <= 1) go(sgdx_untouched_kernel<1>);
    else if (D4 <= 2) go(sgdx_untouched_kernel<2>);
    else if (D4 <= 4) go(sgdx_untouched_kernel<4>);
    else if (D4 <= 8) go(sgdx_untouched_kernel<8>);
    else if (D4 <= 16) go(sgdx_untouched_kernel<16>);
    else go(sgdx_untouched_kernel<32>);
  }
  return cuda_fail(cudaGetLastError(), "sgdx dense pass launch");
}

#endif  // !FFEMB_BACKWARD_WIDE_TU

// ---- host driver -----------------------------------------------------------------------------
template <typename K, int G, int C, int MODE>
static int run_sorted(const TableBatch &tb, int D, int64_t B, int aggr, float lr, int64_t n,
                      int64_t total_rows, const Carve &c, cudaStream_t s, int stage, int64_t n_stream,
                      const OptArgs *oa) {
  K *keys_a = (K *)c.keys_a, *keys_b = (K *)c.keys_b;
  int max_l = 1;
  for (int t = 0; t < tb.num_tables; t++) max_l = tb.t[t].L > max_l ? tb.t[t].L : max_l;
  // small batches: every table's lookups sorted inside one CTA, keygen fused (FFEMB_OWN_SORT=0: the
  // device-wide sort, A/B runs); empty slots (padded) need the device-wide order
  bool seg = std::is_same<K, uint32_t>::value && !tb.padded && tuning().own_sort != 0 &&
             B * (int64_t)max_l <= (int64_t)kSegThreads * kSegMaxItems;
  for (int t = 0; t < tb.num_tables; t++) seg = seg && tb.t[t].R <= 0x7fffffffll;
  if (seg) {
    if constexpr (std::is_same<K, uint32_t>::value) if (stage != STAGE_APPLY) {
      ProfScope prof(FFEMB_PHASE_SORT, 1, s);
      const int64_t per = (B * max_l + kSegThreads - 1) / kSegThreads;
      uint32_t *ko = (uint32_t *)keys_b;
      cudaError_t e = per <= 2 ? launch_seg_sort<2>(tb, B, ko, c.vals_b, s)
                    : per <= 4 ? launch_seg_sort<4>(tb, B, ko, c.vals_b, s)
                    : per <= 8 ? launch_seg_sort<8>(tb, B, ko, c.vals_b, s)
                               : launch_seg_sort<16>(tb, B, ko, c.vals_b, s);
      if (e != cudaSuccess) return cuda_fail(e, "per-table sort launch");
    }
  } else if (stage != STAGE_APPLY) {
    ProfScope prof(FFEMB_PHASE_KEYGEN, 1, s);
    dim3 grid((unsigned)((B * max_l + kThreads - 1) / kThreads), tb.num_tables);
    keygen_kernel<K><<<grid, kThreads, 0, s>>>(tb, B, keys_a, c.vals_a);
  }
  int end_bit = 1;
  const uint64_t max_key = (uint64_t)total_rows - (tb.padded ? 0 : 1);  // (empty slots carry key = total_rows)
  while (end_bit < (int)sizeof(K) * 8 && (max_key >> end_bit) != 0) end_bit++;
  if (stage != STAGE_APPLY && !seg && tb.sort_bits > 0) {
    // aligned key space: one sort per table on that table's row bits (the table bits above them are the
    // same for the whole stretch); same temp storage, same stream
    int launches = 0;
    for (int t = 0; t < tb.num_tables; t++) {
      int bits = 1;
      while (bits < 32 && ((uint64_t)(tb.t[t].R > 1 ? tb.t[t].R - 1 : 1) >> bits) != 0) bits++;
      launches += B * tb.t[t].L > 0 ? 2 + (bits + 7) / 8 : 0;
    }
    ProfScope prof(FFEMB_PHASE_SORT, launches, s);
    for (int t = 0; t < tb.num_tables; t++) {
      const int64_t nt = B * (int64_t)tb.t[t].L;
      if (nt == 0) continue;
      int bits = 1;
      while (bits < 32 && ((uint64_t)(tb.t[t].R > 1 ? tb.t[t].R - 1 : 1) >> bits) != 0) bits++;
      const size_t lb = tb.t[t].lookup_base;
      size_t tmp = c.cub_bytes;
      cudaError_t e = cub::DeviceRadixSort::SortPairs(c.cub_tmp, tmp, keys_a + lb, keys_b + lb, c.vals_a + lb,
                                                      c.vals_b + lb, (int)nt, 0, bits, s);
      if (e != cudaSuccess) return cuda_fail(e, "radix sort (per table)");
    }
  } else if (stage != STAGE_APPLY && !seg) {
    // onesweep: histogram + scan + one pass per 8 key bits
    ProfScope prof(FFEMB_PHASE_SORT, 2 + (end_bit + 7) / 8, s);
    size_t tmp = c.cub_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(c.cub_tmp, tmp, keys_a, keys_b, c.vals_a,
                                                    c.vals_b, (int)n, 0, end_bit, s);
    if (e != cudaSuccess) return cuda_fail(e, "radix sort");
  }
  if (stage == STAGE_SORT) return cuda_fail(cudaGetLastError(), "sort launch");

  constexpr int SLOTS = 32 / G;
  constexpr int CH1 = Ch1<G>::value;
  const int64_t slots_per_cta = (int64_t)(kThreads / 32) * SLOTS;
  const int64_t ntiles1 = (n + CH1 - 1) / CH1;
  K *pk0 = (K *)c.pkeys[0], *pk1 = (K *)c.pkeys[1];
  // hot rows (skew): runs of >= kHotMin lookups leave level 1 (see hot_select_kernel).  Needs plain [B][D]
  // output-grads (the pre-scaled ones of the AVG path qualify), one fragment per lane, no empty slots.
  // OPT-IN (FFEMB_HOT=1): parity-green, but as measured in round 2 (profiles/r2_single_gpu.md) hot_sum
  // takes 8-9 ms on cfg5 where it relieves level 1 of 2.2 ms and the upper levels of 0.3 ms
  int ring_depth = 4;
  if (tuning().r1_ring >= 0) ring_depth = tuning().r1_ring;
  const bool hot = C == 1 && ring_depth > 0 && MODE != APPLY_SGDX && c.hot_cap > 0 && !tb.padded &&
                   tb.og_stride == D && aggr == FFEMB_AGGR_SUM && tuning().hot > 0 && n < (1ll << 32);
  if (hot) {
    ProfScope prof(FFEMB_PHASE_HOT, 4, s);
    cudaMemsetAsync(c.hot_cnt, 0, 256, s);
    cudaMemsetAsync(c.hot_marks, 0, (size_t)ntiles1, s);
    const int64_t probes = (n + kHotStep - 1) / kHotStep;
    hot_select_kernel<K><<<(unsigned)((probes + 255) / 256), 256, 0, s>>>(tb, keys_b, n, CH1, c.hot_list, c.hot_cnt,
                                                                         c.hot_cap);
    hot_mark_kernel<<<dim3(64, tb.num_tables), 256, 0, s>>>(tb, c.hot_list, c.hot_cnt, CH1, c.hot_marks);
    const size_t sm = 2 * kHotSliceBytes + 64;
    cudaFuncSetAttribute(hot_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    hot_sum_kernel<<<dim3(kHotCtas, tb.num_tables), kHotWarps * 32, sm, s>>>(tb, c.vals_b, B, D, c.hot_list, c.hot_cnt,
                                                                            c.hot_partials);
  }
  {
    ProfScope prof(FFEMB_PHASE_REDUCE1, 1, s);
    constexpr int T1 = L1Threads<G>::value;
    constexpr int EPC = (T1 / 32) * SLOTS * CH1;  // entries per CTA
    const unsigned grid1 = (unsigned)((n + EPC - 1) / EPC);
    // CTAs [0, split) hold the streaming tables' lookups (they come first in the key order)
    uint32_t split = (uint32_t)((n_stream + EPC - 1) / EPC);
    if (n_stream == 0 || n_stream >= n) split = 0;
    if (tuning().r1_mix == 0) split = 0;  // tuning aid
    // output-grad of one table larger than (half of) L2 -> the gathers miss: pipelined variant
    bool pipe = B * (int64_t)D * 4 > (48ll << 20);
    int flags = 0;
    if (tuning().r1_flags >= 0) flags = tuning().r1_flags;  // tuning aid
    if (tuning().r1_pipe >= 0) pipe = tuning().r1_pipe != 0;
    const size_t sm_sum = EPC * sizeof(EntryDesc), sm_avg = EPC * (sizeof(EntryDesc) + sizeof(float2));
    auto go = [&](auto kern, size_t sm) {
      kern<<<grid1, T1, sm, s>>>(tb, keys_b, c.vals_b, n, D, lr, pk0, c.prows[0], flags, split, oa, c.consumed);
    };
    // D <= 128: the cp.async ring variant, 4 entries ahead (measured: cfg3 1.36 -> 1.22 ms against the
    // register double buffer; cfg5, whose gathers miss L2, 11.4 ms against 14.5 for the register-
    // pipelined variant and 14.2 for an 8-deep ring).  D > 128 (several fragments per lane): the
    // register variants — pipelined when one table's output-grad exceeds L2 (gathers miss), lean
    // otherwise.  FFEMB_R1_RING=0|4|8 / FFEMB_R1_PIPE override (tuning aids).
    bool pipe_done = false;
    int ring = 4;
    if (tuning().r1_ring >= 0) ring = tuning().r1_ring;
    if (C == 1 && ring > 0) {
      auto gor = [&](auto kern, int S) {
        const size_t sm = EPC * sizeof(EntryDesc) + (size_t)(T1 / 32) * S * 2 * 32 * 16;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        kern<<<grid1, T1, sm, s>>>(tb, keys_b, c.vals_b, n, D, lr, pk0, c.prows[0], split, oa, c.consumed,
                                   hot ? c.hot_marks : nullptr);
      };
      if constexpr (C == 1) {
        const bool avg = aggr != FFEMB_AGGR_SUM;
        if (ring >= 8) { if (avg) gor(reduce1_ring_kernel<K, G, MODE, true, 8>, 8); else gor(reduce1_ring_kernel<K, G, MODE, false, 8>, 8); }
        else { if (avg) gor(reduce1_ring_kernel<K, G, MODE, true, 4>, 4); else gor(reduce1_ring_kernel<K, G, MODE, false, 4>, 4); }
      }
    } else if (aggr == FFEMB_AGGR_SUM) {
      if constexpr (MODE != APPLY_SGDX) { if (pipe) { go(reduce1_kernel<K, G, C, MODE, false, true>, sm_sum); pipe_done = true; } }
      if (!pipe_done) go(reduce1_kernel<K, G, C, MODE, false, false>, sm_sum);
    } else {
      if constexpr (MODE != APPLY_SGDX) { if (pipe) { go(reduce1_kernel<K, G, C, MODE, true, true>, sm_avg); pipe_done = true; } }
      if (!pipe_done) go(reduce1_kernel<K, G, C, MODE, true, false>, sm_avg);
    }
  }
  if (hot) {  // head fragment + the bag ranges' partial sums + tail fragment, then the row's one update
    ProfScope prof(FFEMB_PHASE_HOT, 1, s);
    if constexpr (MODE != APPLY_SGDX)
      hot_apply_kernel<K, MODE><<<dim3(32, tb.num_tables), 256, 0, s>>>(tb, c.hot_list, c.hot_cnt, CH1, D, lr,
                                                                        c.hot_partials, c.prows[0], c.consumed);
  }
  int64_t m = 2 * ntiles1;
  if (ntiles1 > 1) {
    int ups = 1;
    // one CTA walks (kThreads / 32) * SLOTS tiles per round: that many tiles of partial slots (256 slots at
    // D = 128, 2048 at D = 16) finish in the single-CTA tail instead of paying another grid launch
    const int64_t small_level = (int64_t)kSmallLevel * SLOTS;
    for (int64_t mm = m; mm > small_level; mm = 2 * ((mm + CH2 - 1) / CH2)) ups++;
    const bool fused = tuning().up_fused != 0;  // FFEMB_UP_FUSED=0: the tree alone (the round-1 form, A/B runs)
    ProfScope prof(FFEMB_PHASE_REDUCE_UP, ups + (fused ? 1 : 0), s);
    const uint32_t *flag = up_flag(c.consumed);
    const uint8_t *marks = (const uint8_t *)c.consumed;
    if (fused) {
      // every chain of <= kMaxChain partials in one launch; the tree below then usually has nothing left
      const int64_t want = (ntiles1 + 32 * (kThreads / 32) - 1) / (32 * (kThreads / 32));  // 32 tiles per warp and round
      combine_kernel<K, G, C, MODE><<<(unsigned)(want < 148 * 8 ? want : 148 * 8), kThreads, 0, s>>>(
          tb, pk0, c.prows[0], c.consumed, ntiles1, D, lr, oa);
    } else {
      cudaMemsetAsync((void *)flag, 1, 4, s);  // (non-zero: the tree runs unconditionally)
    }
    int cur = 0;
    K *pk[2] = {pk0, pk1};
    while (m > small_level) {  // grid-wide levels while the partial array is large
      const int64_t tiles = (m + CH2 - 1) / CH2;
      reduce_level_kernel<K, G, C, MODE>
          <<<(unsigned)((tiles + slots_per_cta - 1) / slots_per_cta), kThreads, 0, s>>>(
              tb, pk[cur], c.prows[cur], m, CH2, D, lr, pk[cur ^ 1], c.prows[cur ^ 1], oa, flag, cur == 0 && m == 2 * ntiles1 ? marks : nullptr);
      marks = nullptr;
      m = 2 * tiles;
      cur ^= 1;
    }
    reduce_tail_kernel<K, G, C, MODE><<<1, kThreads, 0, s>>>(tb, pk[cur], c.prows[cur], pk[cur ^ 1],
                                                            c.prows[cur ^ 1], m, CH2, D, lr, oa, flag, marks);
  }
  return cuda_fail(cudaGetLastError(), "sorted backward launch");
}

template <typename K, int MODE>
static int dispatch_g(const TableBatch &tb, int D, int64_t B, int aggr, float lr, int64_t n,
                      int64_t rows, const Carve &c, cudaStream_t s, int stage, int64_t n_stream,
                      const OptArgs *oa) {
  const int D4 = D / 4;
  if (D4 <= 1) return run_sorted<K, 1, 1, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  if (D4 <= 2) return run_sorted<K, 2, 1, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  if (D4 <= 4) return run_sorted<K, 4, 1, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  if (D4 <= 8) return run_sorted<K, 8, 1, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  if (D4 <= 16) return run_sorted<K, 16, 1, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  if (D4 <= 32) return run_sorted<K, 32, 1, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  if (D4 <= 64) return run_sorted<K, 32, 2, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  return run_sorted<K, 32, 4, MODE>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
}

// 64-bit keys (more than 2^32 - 1 rows on one GPU): the same templates instantiated in
// ffemb_backward_wide.cu, which includes this file with FFEMB_BACKWARD_WIDE_TU defined — the two
// halves then compile side by side.
int dispatch_wide(const TableBatch &tb, int D, int64_t B, int aggr, float lr, int64_t n, int64_t rows,
                  const Carve &c, cudaStream_t s, int stage, int64_t n_stream, const OptArgs *oa, int mode);
#ifdef FFEMB_BACKWARD_WIDE_TU
int dispatch_wide(const TableBatch &tb, int D, int64_t B, int aggr, float lr, int64_t n, int64_t rows,
                  const Carve &c, cudaStream_t s, int stage, int64_t n_stream, const OptArgs *oa, int mode) {
  if (mode == APPLY_SGD) return dispatch_g<uint64_t, APPLY_SGD>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  if (mode == APPLY_SGDX) return dispatch_g<uint64_t, APPLY_SGDX>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
  return dispatch_g<uint64_t, APPLY_GRAD>(tb, D, B, aggr, lr, n, rows, c, s, stage, n_stream, oa);
}
#else

// A side stream + fork/join events per host thread and device (created on first use, never freed:
// a handful per process).  FFEMB_SMALL_FORK=0 turns the fork off (tuning aid / A-B runs).
struct SideStream {
  cudaStream_t s;
  cudaEvent_t fork, join;
  bool ok, tried;
};
static SideStream *side_stream() {
  static thread_local SideStream ctx[32];
  const bool enabled = tuning().small_fork != 0;
  if (!enabled) return nullptr;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 32) return nullptr;
  SideStream &c = ctx[dev];
  if (!c.tried) {
    c.tried = true;
    c.ok = cudaStreamCreateWithFlags(&c.s, cudaStreamNonBlocking) == cudaSuccess &&
           cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&c.join, cudaEventDisableTiming) == cudaSuccess;
    if (!c.ok) cudaGetLastError();
  }
  return c.ok ? &c : nullptr;
}

int launch_backward_sorted(const TableBatch &tb_all, int64_t D, int64_t B, int aggr, float lr,
                           ApplyMode mode, void *ws, size_t ws_bytes, cudaStream_t s, int stage,
                           const OptArgs *opt_host) {
  int64_t n_all = 0;
  for (int t = 0; t < tb_all.num_tables; t++) n_all += B * tb_all.t[t].L;
  if (mode == APPLY_SGDX && (!opt_host || stage != STAGE_ALL)) {
    set_error("general SGD update: optimizer arguments missing, or a split (prepared) call");
    return FFEMB_EINVAL;
  }
  if (n_all == 0 && mode != APPLY_SGDX) return FFEMB_OK;
  if (D % 4 != 0 || D > 512) {
    set_error("sorted backward needs D %% 4 == 0 and D <= 512 (got %lld)", (long long)D);
    return FFEMB_EINVAL;
  }
  for (int t = 0; t < tb_all.num_tables; t++) {
    if (stage != STAGE_SORT && ((uintptr_t)tb_all.t[t].table | (uintptr_t)tb_all.t[t].out) % 16 != 0) {
      set_error("sorted backward needs 16-byte aligned table and output-grad pointers (table %d)", t);
      return FFEMB_EINVAL;
    }
  }
  if (n_all >= (1ll << 31) || B > (1ll << BAG_BITS)) {
    set_error("sorted backward: too many lookups for one call (%lld)", (long long)n_all);
    return FFEMB_EINVAL;
  }
  const size_t need = backward_workspace_bytes(n_all, (int64_t)tb_all.row_base[tb_all.num_tables], D);
  if (n_all > 0 && (!ws || ((uintptr_t)ws & 255) || ws_bytes < need)) {
    set_error("workspace: need %zu bytes, 256-byte aligned (got %zu at %p)", need, ws_bytes, ws);
    return FFEMB_EWORKSPACE;
  }
  // Split: tables whose whole gradient fits on chip go bag-major (no sort), the rest are sorted.
  TableBatch tb, ts;
  memset(&tb, 0, sizeof(tb));
  memset(&ts, 0, sizeof(ts));
  tb.og_stride = ts.og_stride = tb_all.og_stride;
  tb.padded = tb_all.padded;
  const bool no_chip = mode == APPLY_SGDX || tb_all.padded;  // (the bag-major kernels know no empty slots)
  int n_tiny = 0;
  for (int pass = 1; pass <= 2; pass++) {  // tiny tables first, then small ones
    for (int t = 0; t < tb_all.num_tables; t++) {
      if (!no_chip && tb_all.t[t].L > 0 && backward_path(tb_all.t[t].R, tb_all.t[t].L, D, B) == pass) {
        ts.t[ts.num_tables++] = tb_all.t[t];
        n_tiny += pass == 1;
      }
    }
  }
  // key order: the tables far larger than L2 first, then the L2-resident ones (see mix_cta)
  uint64_t rows_l = 0, n = 0, n_stream = 0;
  int src_table[FFEMB_MAX_TABLES];  // position in the caller's list of the k-th table in key order
  for (int pass = 0; pass < 2; pass++) {
    for (int t = 0; t < tb_all.num_tables; t++) {
      if (!no_chip && tb_all.t[t].L > 0 && backward_path(tb_all.t[t].R, tb_all.t[t].L, D, B) != 0) continue;
      const bool streams = tb_all.t[t].R * D * 4 > kStreamTableBytes;
      if (streams != (pass == 0)) continue;
      TableDesc &d = tb.t[tb.num_tables];
      d = tb_all.t[t];
      d.lookup_base = (uint32_t)n;
      src_table[tb.num_tables] = t;
      tb.row_base[tb.num_tables++] = rows_l;
      rows_l += (uint64_t)d.R;
      n += (uint64_t)B * d.L;
    }
    if (pass == 0) n_stream = n;
  }
  tb.row_base[tb.num_tables] = rows_l;
  // Few large tables: an ALIGNED key space, (t << kbits) + row, lets every table be sorted on its own row
  // bits — cfg5: eight 10 M-row tables need 24 bits = 3 radix passes each instead of the 4 passes of the
  // 27-bit global key (4.7 -> 3.6 ms for 134 M pairs).  Worth it when the passes saved outweigh the extra
  // chains of launches (measured: ~9.5 ps per pair and pass — 118 us for 12.4 M pairs —, ~45 us per chain); FFEMB_SORT_ALIGNED=0/1 forces.
  if (tb.num_tables > 1 && !tb.padded && !wide_keys((int64_t)rows_l)) {
    int kbits = 1, gbits = 1;
    auto bits_of = [](uint64_t maxv) { int b = 1; while (b < 64 && (maxv >> b) != 0) b++; return b; };
    gbits = bits_of(rows_l > 1 ? rows_l - 1 : 1);
    for (int t = 0; t < tb.num_tables; t++) kbits = std::max(kbits, bits_of(tb.t[t].R > 1 ? (uint64_t)tb.t[t].R - 1 : 1));
    double saved_ns = 0;
    for (int t = 0; t < tb.num_tables; t++) {
      const int bt = bits_of(tb.t[t].R > 1 ? (uint64_t)tb.t[t].R - 1 : 1);
      saved_ns += (double)B * tb.t[t].L * ((gbits + 7) / 8 - (bt + 7) / 8) * 0.0095;
    }
    bool aligned = saved_ns > 45e3 * (tb.num_tables - 1) && ((uint64_t)tb.num_tables << kbits) < 0xffffffffull &&
                   (int64_t)n > (int64_t)kSegThreads * kSegMaxItems * tb.num_tables;
    if (tuning().sort_aligned >= 0) aligned = tuning().sort_aligned != 0 && ((uint64_t)tb.num_tables << kbits) < 0xffffffffull;
    if (aligned) {
      tb.sort_bits = kbits;
      for (int t = 0; t <= tb.num_tables; t++) tb.row_base[t] = (uint64_t)t << kbits;
      rows_l = (uint64_t)tb.num_tables << kbits;
    }
  }
  const int64_t rows = (int64_t)rows_l;
  const Carve c = carve((char *)ws, (int64_t)n, rows > 0 ? rows : 1, D);
  int aggr_l1 = aggr;
  // AVG / MEAN with room in the workspace (ffemb_backward_workspace_avg): divide once per bag up front
  // and let the segmented sum run as a plain sum (FFEMB_PRESCALE=0: divide in the walk, A/B runs)
  if (aggr != FFEMB_AGGR_SUM && stage != STAGE_SORT && n > 0 && tuning().prescale != 0) {
    const bool all = tb.og_stride != D;  // (one row stride per call: a strided source moves every table)
    TableBatch tp;
    memset(&tp, 0, sizeof(tp));
    size_t extra = 0;
    for (int t = 0; t < tb.num_tables; t++) {
      if (tb.t[t].L >= 2 || all) {
        tp.t[tp.num_tables] = tb.t[t];
        tp.t[tp.num_tables].lookup_base = (uint32_t)t;
        tp.num_tables++;
        extra += prescale_bytes(B, D);
      }
    }
    if (tp.num_tables > 0 && ws_bytes >= need + extra) {
      char *dst = (char *)ws + need;
      for (int k = 0; k < tp.num_tables; k++) {
        tp.t[k].table = (float *)dst;
        tb.t[tp.t[k].lookup_base].out = (float *)dst;
        dst += prescale_bytes(B, D);
      }
      {
        ProfScope prof(FFEMB_PHASE_PRESCALE, 1, s);
        const int64_t want = (B * (D / 4) + 255) / 256;
        dim3 grid((unsigned)(want < 148 * 8 ? want : 148 * 8), tp.num_tables);
        prescale_kernel<<<grid, 256, 0, s>>>(tp, B, (int)D, tb.og_stride);
      }
      tb.og_stride = D;
      aggr_l1 = FFEMB_AGGR_SUM;  // (for the sorted tables only: the bag-major kernels keep the call's aggr)
    }
  }
  // The bag-major kernels of the small tables are shared-memory-bound and touch neither the sort
  // buffers nor the sorted tables: when there is sorted work too they are forked onto a side stream
  // (per host thread and device) and run UNDER the radix sort / the segmented sum instead of in front
  // of them; the caller's stream joins at the end, so stream semantics are unchanged.
  SideStream *fork = nullptr;
  if (ts.num_tables && stage != STAGE_SORT) {
    if (n > 0 && mode != APPLY_SGDX) fork = side_stream();
    cudaStream_t ss = s;
    if (fork) {
      if (cudaEventRecord(fork->fork, s) != cudaSuccess || cudaStreamWaitEvent(fork->s, fork->fork, 0) != cudaSuccess)
        fork = nullptr;
      else
        ss = fork->s;
    }
    int rc = launch_backward_small(ts, n_tiny, D, B, aggr, lr, mode, (float *)((char *)ws + c.total), ss);
    if (fork) {
      if (cudaEventRecord(fork->join, fork->s) != cudaSuccess && !rc) rc = cuda_fail(cudaGetLastError(), "side stream join");
      if (rc) cudaStreamWaitEvent(s, fork->join, 0);
    }
    if (rc) return rc;
  }
  struct Join {  // the caller's stream waits for the forked kernels on every way out
    SideStream *f;
    cudaStream_t s;
    ~Join() { if (f) cudaStreamWaitEvent(s, f->join, 0); }
  } join{fork, s};
  if (mode == APPLY_SGDX) {
    // optimizer arguments in key order of the tables, copied to the tail of the workspace (pageable
    // source: the copy is staged before cudaMemcpyAsync returns)
    OptArgs o = *opt_host;
    for (int k = 0; k < tb.num_tables; k++) {
      o.V[k] = opt_host->V[src_table[k]];
      o.M[k] = opt_host->M[src_table[k]];
      o.touched[k] = opt_host->touched[src_table[k]];
    }
    int rc = FFEMB_OK;
    if (n > 0) {  // (an empty batch has no sparse pass; every row still gets the zero-gradient update)
      OptArgs *oa = (OptArgs *)((char *)ws + c.total + small_partials_bound(D));
      cudaError_t e = cudaMemcpyAsync(oa, &o, sizeof(o), cudaMemcpyHostToDevice, s);
      if (e != cudaSuccess) return cuda_fail(e, "optimizer arguments upload");
      rc = wide_keys(rows) ? dispatch_wide(tb, (int)D, B, aggr_l1, o.lr, n, rows, c, s, stage, (int64_t)n_stream, oa, APPLY_SGDX)
                           : dispatch_g<uint32_t, APPLY_SGDX>(tb, (int)D, B, aggr_l1, o.lr, n, rows, c, s, stage, (int64_t)n_stream, oa);
    }
    if (rc) return rc;
    return launch_sgdx_untouched(tb, (int)D, o, s);
  }
  if (n == 0) return FFEMB_OK;
  if (wide_keys(rows))
    return dispatch_wide(tb, (int)D, B, aggr_l1, lr, n, rows, c, s, stage, (int64_t)n_stream, nullptr, (int)mode);
  return mode == APPLY_SGD ? dispatch_g<uint32_t, APPLY_SGD>(tb, (int)D, B, aggr_l1, lr, n, rows, c, s, stage, (int64_t)n_stream, nullptr)
                           : dispatch_g<uint32_t, APPLY_GRAD>(tb, (int)D, B, aggr_l1, lr, n, rows, c, s, stage, (int64_t)n_stream, nullptr);
}

// ---- the atomic comparison: reference-equivalent unordered scatter ---------------------------
template <int G, int C>
__global__ void __launch_bounds__(kThreads)
bwd_atomic_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, int aggr) {
  constexpr int SLOTS = 32 / G;
  const TableDesc &td = tb.t[blockIdx.y];
  const int L = td.L;
  const int D4 = D >> 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int64_t bag = ((int64_t)blockIdx.x * (kThreads / 32) + warp) * SLOTS + sub;
  const bool valid = bag < B;
  const int64_t b = valid ? bag : (B - 1);
  const int64_t *ip = td.idx + b * (int64_t)L;
  float4 g[C];
#pragma unroll
  for (int c = 0; c < C; c++) {
    g[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (sl + c * G < D4) {
      g[c] = ldg_row(td.out + b * tb.og_stride + 4 * (sl + c * G));
      if (aggr != FFEMB_AGGR_SUM) g[c] = f4_div(g[c], (float)L);
    }
  }
  for (int j0 = 0; j0 < L; j0 += G) {
    const int64_t mine = (j0 + sl < L) ? __ldg(ip + j0 + sl) : 0;
    const int cnt = min(G, L - j0);
    for (int jj = 0; jj < cnt; jj++) {
      const int64_t r = shfl_i64(mine, sub * G + jj);
      if (valid) {
        float *wp = td.table + r * (int64_t)D;
#pragma unroll
        for (int c = 0; c < C; c++)
          if (sl + c * G < D4) atomicAdd(reinterpret_cast<float4 *>(wp + 4 * (sl + c * G)), g[c]);
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads)
bwd_atomic_generic_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, int aggr) {
  const TableDesc &td = tb.t[blockIdx.y];
  const int L = td.L;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t bag = (int64_t)blockIdx.x * (kThreads / 32) + warp;
  if (bag >= B) return;
  for (int d = lane; d < D; d += 32) {
    float g = td.out[bag * tb.og_stride + d];
    if (aggr != FFEMB_AGGR_SUM) g = g / (float)L;
    for (int j = 0; j < L; j++)
      atomicAdd(td.table + __ldg(td.idx + bag * (int64_t)L + j) * (int64_t)D + d, g);
  }
}

template <int G, int C>
static void launch_atomic(const TableBatch &tb, int D, int64_t B, int aggr, cudaStream_t s) {
  const int64_t per_cta = (int64_t)(kThreads / 32) * (32 / G);
  dim3 grid((unsigned)((B + per_cta - 1) / per_cta), tb.num_tables);
  bwd_atomic_kernel<G, C><<<grid, kThreads, 0, s>>>(tb, D, B, aggr);
}

int launch_backward_atomic(const TableBatch &tb, int64_t D64, int64_t B, int aggr, cudaStream_t s) {
  const int D = (int)D64;
  if (B == 0 || tb.num_tables == 0) return FFEMB_OK;
  ProfScope prof(FFEMB_PHASE_ATOMIC, 1, s);
  bool vec = (D % 4 == 0) && D <= 512;
  for (int t = 0; t < tb.num_tables; t++)
    vec = vec && (((uintptr_t)tb.t[t].table | (uintptr_t)tb.t[t].out) % 16 == 0);
  if (!vec) {
    dim3 grid((unsigned)((B + kThreads / 32 - 1) / (kThreads / 32)), tb.num_tables);
    bwd_atomic_generic_kernel<<<grid, kThreads, 0, s>>>(tb, D, B, aggr);
  } else {
    const int D4 = D / 4;
    if (D4 <= 1) launch_atomic<1, 1>(tb, D, B, aggr, s);
    else if (D4 <= 2) launch_atomic<2, 1>(tb, D, B, aggr, s);
    else if (D4 <= 4) launch_atomic<4, 1>(tb, D, B, aggr, s);
    else if (D4 <= 8) launch_atomic<8, 1>(tb, D, B, aggr, s);
    else if (D4 <= 16) launch_atomic<16, 1>(tb, D, B, aggr, s);
    else if (D4 <= 32) launch_atomic<32, 1>(tb, D, B, aggr, s);
    else if (D4 <= 64) launch_atomic<32, 2>(tb, D, B, aggr, s);
    else launch_atomic<32, 4>(tb, D, B, aggr, s);
  }
  return cuda_fail(cudaGetLastError(), "atomic backward launch");
}

#endif  // !FFEMB_BACKWARD_WIDE_TU
}  // namespace ffemb
