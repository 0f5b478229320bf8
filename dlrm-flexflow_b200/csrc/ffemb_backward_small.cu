// ffemb_backward_small.cu — backward (+ SGD row update) for tables whose WHOLE gradient fits on
// chip: no sort, no gather of output-grad rows through L2.
//
// Why (profiles/r1b_summary.md): on the MLPerf shape 38 % of all lookups hit tables with <= 155
// rows (3, 4, 10, 14, 36, 63, 108, 155).  Through the sorted path every one of those lookups is a
// 512-byte gather of an output-grad row from L2 (the same 33.5 MB re-read L times) plus a sort
// pair; reduce level 1 is bound by exactly that L2->SM traffic (~8 TB/s measured).  Bag-major,
// the output-grad of such a table is streamed from HBM ONCE and the table gradient is accumulated
// on chip:
//   tiny  (R <= 16)            : per-lane REGISTER accumulators, one per row.  Per bag a slot counts
//                                the multiplicity c_r of every row among the bag's indices with
//                                warp votes and does acc[r] = fma(c_r, g, acc[r]).
//   small (R*D*4 <= 80 KB)     : one [R][D] accumulator per CTA in SHARED memory.  Output-grad rows
//                                and indices are staged chunk-wise with cp.async (double buffered);
//                                every accumulator row is OWNED by one slot (row % #slots), which
//                                adds that row's lookups in (bag, j) order — no atomics, no races.
// Both leave one partial [R][D] per CTA; small_apply_kernel adds the partials of a row in CTA
// order and applies W = fma(-lr, sum, W) (or accumulates into the dense grad).  Every order is
// fixed by (B, L, R, D) alone: bit-reproducible run to run, like the sorted path.  A row looked
// up once gets exactly the reference's arithmetic (0 + g, then one FFMA).
//
// Replaces, for these tables, the same reference code as ffemb_backward.cu:
//   src/runtime/model.cc:466-490 (zero-fill), src/ops/embedding.cu:192-217 (embed_backward),
//   src/runtime/optimizer_kernel.cu:23-41 (sgd_update).
#include <stdlib.h>

#include "ffemb_internal.cuh"

namespace ffemb {

static constexpr int kTinyRows = 16;
static constexpr int64_t kSmallAccBytes = 80ll << 10;
static constexpr int kSmallMaxBag = 64;
static constexpr int kTinyThreads = 256, kTinyCtas = 148 * 4;
static constexpr int kSmallThreads = 512, kSmallCtas = 148 * 2, kChunkBags = 16;  // bags per CTA are a multiple of kChunkBags

// 0 = sorted path, 1 = tiny (register accumulators), 2 = small (shared-memory accumulator)
int backward_path(int64_t R, int64_t L, int64_t D, int64_t B) {
  if (tuning().small == 0) return 0;  // tuning aid: FFEMB_SMALL=0 forces the sorted path
  // (D = 16 was tried for the tiny path in round 2: its kernels cost cfg2 0.11 ms where the sorted path's
  // share was 0.03 — too little work per bag for a chunk-staged kernel)
  if (D % 4 != 0 || D < 32 || D > 128 || L < 1 || L > kSmallMaxBag || R < 1) return 0;
  const int64_t n = B * L;
  if (n < 16384 || n < 16 * R) return 0;  // little duplication / little work: sorting is fine
  if (R <= kTinyRows) return 1;
  if (R * D * 4 <= kSmallAccBytes && R <= 256 && L >= 8) return 2;  // g re-read L times otherwise
  return 0;
}

size_t small_partials_bound(int64_t D) {
  return (size_t)kTinyCtas * kTinyRows * D * 4 + (size_t)kSmallCtas * kSmallAccBytes + 4096;
}

__device__ __forceinline__ int find_by_cta(const TableBatch &tb, uint32_t cta) {
  int lo = 0, hi = tb.num_tables - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (tb.cta_base[mid] <= cta) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__device__ __forceinline__ float4 f4_scale_div(float4 a, float d, float r) {
  return f4_div_exact(a, d, r);
}

// ---- chunk staging shared by both kernels ---------------------------------------------------------
// A CTA owns a contiguous range of bags of ONE table and walks it in chunks: the chunk's
// output-grad rows ([nb][D] floats, contiguous) and indices ([nb][L] int64, contiguous) are copied
// into shared memory with cp.async, one chunk ahead of the one being consumed (two buffers), so a
// resident CTA always has a whole chunk (tens of KB) in flight and no registers are tied up.
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)),
               "l"(gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)),
               "l"(gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

struct ChunkStage {
  const float *g;      // output-grad of the table
  const int64_t *idx;  // indices of the table
  float *gs;           // [2][chb][D]
  int64_t *is;         // [2][chb][L]
  int64_t b0, b1;      // bag range of this CTA
  int chb, D, L, nthreads;
  int64_t ogs;         // floats between consecutive output-grad rows (D, or the Concat row width)
  __device__ __forceinline__ int bags(int k) const {
    return (int)min((int64_t)chb, b1 - (b0 + (int64_t)k * chb));
  }
  __device__ __forceinline__ void issue(int k, int buf) const {
    const int64_t bb = b0 + (int64_t)k * chb;
    const int nb = bags(k);
    const float *gsrc = g + bb * ogs;
    float *gdst = gs + (size_t)buf * chb * D;
    if (ogs == D) {  // plain [B][D] buffer: one contiguous stretch
      for (int i = threadIdx.x; i < nb * (D >> 2); i += nthreads) cp_async16(gdst + 4 * i, gsrc + 4 * i);
    } else {         // column block of a wider matrix: D contiguous floats per row
      const int D4 = D >> 2;
      for (int i = threadIdx.x; i < nb * D4; i += nthreads) {
        const int r = i / D4, c4 = i - r * D4;
        cp_async16(gdst + 4 * i, gsrc + r * ogs + 4 * c4);
      }
    }
    const int64_t *isrc = idx + bb * (int64_t)L;
    int64_t *idst = is + (size_t)buf * chb * L;
    for (int i = threadIdx.x; i < nb * L; i += nthreads) cp_async8(idst + i, isrc + i);
    cp_async_commit();
  }
};

// ---- tiny: register accumulators --------------------------------------------------------------
// TableDesc fields reused: lookup_base = bags per CTA; row_base[t] = first float of the table's
// partials; cta_base[t] = first CTA of the table.
static constexpr int kTinyChunk = 64;  // bags per chunk
template <int G, int RMAX, bool AVG>
__global__ void __launch_bounds__(kTinyThreads, 2)
bwd_tiny_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, float *partials) {
  constexpr int SLOTS = 32 / G, NSLOT = (kTinyThreads / 32) * SLOTS;
  constexpr int NB = (kSmallMaxBag + G - 1) / G;  // index registers per lane
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int t = find_by_cta(tb, blockIdx.x);
  const TableDesc &td = tb.t[t];
  const int L = td.L, R = (int)td.R;
  const int64_t bpc = td.lookup_base;
  const int64_t c = blockIdx.x - tb.cta_base[t];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int slot = warp * SLOTS + sub;
  const unsigned slotmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  const bool own = sl < (D >> 2);
  const float Lf = (float)L, Linv = __frcp_rn(Lf);
  ChunkStage st;
  st.g = td.out; st.idx = td.idx; st.D = D; st.L = L; st.ogs = tb.og_stride; st.chb = kTinyChunk; st.nthreads = kTinyThreads;
  st.b0 = c * bpc; st.b1 = min(B, st.b0 + bpc);
  st.gs = reinterpret_cast<float *>(smem_raw);
  st.is = reinterpret_cast<int64_t *>(st.gs + 2 * kTinyChunk * D);

  float4 acc[RMAX];
#pragma unroll
  for (int r = 0; r < RMAX; r++) acc[r] = make_float4(0.f, 0.f, 0.f, 0.f);

  const int nchunks = (int)((st.b1 - st.b0 + kTinyChunk - 1) / kTinyChunk);
  if (nchunks > 0) st.issue(0, 0);
  for (int k = 0; k < nchunks; k++) {
    const int buf = k & 1;
    if (k + 1 < nchunks) { st.issue(k + 1, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const int nb = st.bags(k);
    const float *gb = st.gs + (size_t)buf * kTinyChunk * D;
    const int64_t *ib = st.is + (size_t)buf * kTinyChunk * L;
    for (int bb0 = 0; bb0 < nb; bb0 += NSLOT) {  // CTA-uniform (votes need whole warps)
      const int bb = bb0 + slot;
      const bool valid = bb < nb;
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
      if (valid && own) g = *reinterpret_cast<const float4 *>(gb + bb * D + 4 * sl);
      if (AVG) g = f4_scale_div(g, Lf, Linv);  // gradient = output-grad / in_dim, embedding.cu:210
      int v[NB];
#pragma unroll
      for (int kk = 0; kk < NB; kk++) {
        const int j = kk * G + sl;
        v[kk] = (valid && j < L) ? (int)ib[bb * L + j] : -1;
      }
#pragma unroll
      for (int r = 0; r < RMAX; r++) {
        if (r < R) {  // CTA-uniform
          int cnt = 0;
#pragma unroll
          for (int kk = 0; kk < NB; kk++)
            if (kk * G < L) cnt += __popc(__ballot_sync(0xffffffffu, v[kk] == r) & slotmask);
          if (cnt) {  // (0 * inf would poison the row)
            const float cf = (float)cnt;
            acc[r].x = fmaf(cf, g.x, acc[r].x); acc[r].y = fmaf(cf, g.y, acc[r].y);
            acc[r].z = fmaf(cf, g.z, acc[r].z); acc[r].w = fmaf(cf, g.w, acc[r].w);
          }
        }
      }
    }
    __syncthreads();  // everyone is done with buffer `buf` before chunk k+2 overwrites it
  }
  // CTA partial = slots added in slot order (the staging buffers are free now)
  float *sacc = reinterpret_cast<float *>(smem_raw);
  for (int s2 = 0; s2 < NSLOT; s2++) {
    if (slot == s2 && own) {
#pragma unroll
      for (int r = 0; r < RMAX; r++) {
        if (r < R) {
          float4 *p = reinterpret_cast<float4 *>(sacc + r * D + 4 * sl);
          float4 a = acc[r];
          if (s2) { const float4 o = *p; a = make_float4(o.x + a.x, o.y + a.y, o.z + a.z, o.w + a.w); }
          *p = a;
        }
      }
    }
    __syncthreads();
  }
  float4 *dst = reinterpret_cast<float4 *>(partials + tb.row_base[t] + (uint64_t)c * R * D);
  for (int i = threadIdx.x; i < R * (D >> 2); i += kTinyThreads)
    dst[i] = reinterpret_cast<const float4 *>(sacc)[i];
}

// ---- small: shared-memory accumulator, rows owned by slots ---------------------------------------
// Per chunk (<= 512 lookups) the lookups are first put in ROW order — owner-major, so that the rows
// a slot owns (row % #slots == slot) are one contiguous stretch — with a stable counting sort in
// shared memory (warp match + ranks, per-batch counts, one prefix): positions depend on the data
// only, so the order is reproducible.  Then every slot walks its stretch row by row with the row's
// accumulator fragment in REGISTERS: per lookup one broadcast LDS (bag number), one LDS.128 of the
// staged output-grad fragment and four FADDs; the accumulator row is read and written once per
// chunk.  (The first version sorted by owner only and did LDS.128 + STS.128 on the accumulator per
// lookup: bound by the shared-memory pipe, l1tex 84 %, profiles/r1d_summary.md.)
static constexpr int kSmallMaxLookups = 512, kSmallMaxChunk = 32, kSmallMaxRows = 256;
__host__ __device__ __forceinline__ int small_chunk_bags(int L) {
  const int c = kSmallMaxLookups / L;
  return c > kSmallMaxChunk ? kSmallMaxChunk : c;
}
__host__ __device__ __forceinline__ int small_bins(int R, int NU) { return ((R + NU - 1) / NU) * NU; }

template <int G, bool AVG>
__global__ void __launch_bounds__(kSmallThreads)
bwd_small_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, float *partials, uint32_t cta_off) {
  constexpr int SLOTS = 32 / G, NW = kSmallThreads / 32, NU = NW * SLOTS;  // NU: power of two
  constexpr int NBATCH = kSmallMaxLookups / 32;
  static_assert(kSmallMaxLookups == kSmallThreads, "one lookup per thread and chunk");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t cta = blockIdx.x + cta_off;
  const int t = find_by_cta(tb, cta);
  const TableDesc &td = tb.t[t];
  const int L = td.L, R = (int)td.R;
  const int rpo = (R + NU - 1) / NU, NBIN = rpo * NU;  // rows per owner; bin = owner * rpo + row / NU
  const int64_t bpc = td.lookup_base;
  const int64_t c = cta - tb.cta_base[t];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const bool own = sl < (D >> 2);
  const float Lf = (float)L, Linv = __frcp_rn(Lf);
  const uint32_t magic = (uint32_t)((0x100000000ull + (uint32_t)L - 1) / (uint32_t)L);  // e / L, e < 2^16
  ChunkStage st;
  st.g = td.out; st.idx = td.idx; st.D = D; st.L = L; st.ogs = tb.og_stride; st.chb = small_chunk_bags(L); st.nthreads = kSmallThreads;
  st.b0 = c * bpc; st.b1 = min(B, st.b0 + bpc);
  float *acc = reinterpret_cast<float *>(smem_raw);                         // [R][D]
  st.gs = acc + (size_t)R * D;                                              // [2][chb][D]
  st.is = reinterpret_cast<int64_t *>(st.gs + 2 * st.chb * D);              // [2][chb][L]
  uint32_t *sorted = reinterpret_cast<uint32_t *>(st.is + 2 * st.chb * L);  // [512] bag number
  uint32_t *wcnt = sorted + kSmallMaxLookups;                               // [NBATCH][NBIN]
  uint32_t *start = wcnt + NBATCH * NBIN;                                   // [NBIN + 1]
  uint32_t *total = start + NBIN + 1;                                       // [NBIN]

  for (int i = threadIdx.x; i < R * (D >> 2); i += kSmallThreads)
    reinterpret_cast<float4 *>(acc)[i] = make_float4(0.f, 0.f, 0.f, 0.f);

  const int nchunks = (int)((st.b1 - st.b0 + st.chb - 1) / st.chb);
  if (nchunks > 0) st.issue(0, 0);
  for (int k = 0; k < nchunks; k++) {
    const int buf = k & 1;
    if (k + 1 < nchunks) { st.issue(k + 1, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    for (int i = threadIdx.x; i < NBATCH * NBIN; i += kSmallThreads) wcnt[i] = 0;
    __syncthreads();  // (a) chunk k landed; counts cleared; accumulator zero-fill visible (k == 0)
    const int nb = st.bags(k);
    float *gb = st.gs + (size_t)buf * st.chb * D;
    const int64_t *ib = st.is + (size_t)buf * st.chb * L;
    const int n = nb * L;
    // 1. rank of every lookup among the lookups of its warp-batch that hit the same row
    const int e = threadIdx.x;  // batch = warp
    const int row = e < n ? (int)ib[e] : -1;
    const int bin = row >= 0 ? (row & (NU - 1)) * rpo + row / NU : NBIN;
    const unsigned m = __match_any_sync(0xffffffffu, bin);
    const int rank = __popc(m & ((1u << lane) - 1u));
    if (rank == 0 && row >= 0) wcnt[warp * NBIN + bin] = __popc(m);
    if (AVG)  // gradient = output-grad / in_dim per element, once per staged row (embedding.cu:210)
      for (int i = threadIdx.x; i < nb * (D >> 2); i += kSmallThreads)
        reinterpret_cast<float4 *>(gb)[i] = f4_div_exact(reinterpret_cast<const float4 *>(gb)[i], Lf, Linv);
    __syncthreads();  // (b)
    // 2. per bin: exclusive prefix of the per-batch counts (lane = batch)
    for (int o = warp; o < NBIN; o += NW) {
      const uint32_t x = lane < NBATCH ? wcnt[lane * NBIN + o] : 0u;
      uint32_t incl = x;
#pragma unroll
      for (int d = 1; d < NBATCH; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
      }
      if (lane < NBATCH) wcnt[lane * NBIN + o] = incl - x;
      if (lane == NBATCH - 1) total[o] = incl;
    }
    __syncthreads();  // (c)
    if (warp == 0) {  // exclusive prefix over the bins: a lane takes a run of consecutive bins
      const int per = (NBIN + 31) / 32;
      uint32_t mine = 0;
      for (int i = 0; i < per; i++) {
        const int o = lane * per + i;
        if (o < NBIN) mine += total[o];
      }
      uint32_t incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
      }
      uint32_t run = incl - mine;
      for (int i = 0; i < per; i++) {
        const int o = lane * per + i;
        if (o < NBIN) { start[o] = run; run += total[o]; }
      }
      if (lane == 31) start[NBIN] = incl;
    }
    __syncthreads();  // (d)
    // 3. stable scatter into row order
    if (row >= 0) {
      const uint32_t bag = L == 1 ? (uint32_t)e : __umulhi((uint32_t)e, magic);
      sorted[start[bin] + wcnt[warp * NBIN + bin] + rank] = bag;
    }
    __syncthreads();  // (e)
    // 4. every slot adds the lookups of the rows it owns, row by row, in (bag, j) order
    if (own) {
      const int unit = warp * SLOTS + sub;
      for (int i = 0; i < rpo; i++) {
        const int r = i * NU + unit;
        const uint32_t lo = start[unit * rpo + i], hi = start[unit * rpo + i + 1];
        if (r < R && hi > lo) {
          float4 *ap = reinterpret_cast<float4 *>(acc + r * D + 4 * sl);
          float4 a = *ap;
          for (uint32_t q = lo; q < hi; q++) {
            const float4 x = *reinterpret_cast<const float4 *>(gb + sorted[q] * D + 4 * sl);
            a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
          }
          *ap = a;
        }
      }
    }
    __syncthreads();  // (f) everyone is done with buffer `buf` and `sorted` before they are rewritten
  }
  float4 *dst = reinterpret_cast<float4 *>(partials + tb.row_base[t] + (uint64_t)c * R * D);
  for (int i = threadIdx.x; i < R * (D >> 2); i += kSmallThreads)
    dst[i] = reinterpret_cast<const float4 *>(acc)[i];
}

// ---- cross-CTA sum in CTA order + the row update ---------------------------------------------------
template <int G, int MODE>
__global__ void __launch_bounds__(256)
small_apply_kernel(const __grid_constant__ TableBatch tb, int D, const float *partials, float lr) {
  constexpr int SLOTS = 32 / G;
  const TableDesc &td = tb.t[blockIdx.y];
  const int R = (int)td.R;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int row = ((int)blockIdx.x * 8 + warp) * SLOTS + sub;
  if (row >= R || sl >= (D >> 2)) return;
  const int nct = (int)(tb.cta_base[blockIdx.y + 1] - tb.cta_base[blockIdx.y]);
  const float *p = partials + tb.row_base[blockIdx.y] + (size_t)row * D + 4 * sl;
  float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
  const size_t stride = (size_t)R * D;
  int c = 0;
  for (; c + 8 <= nct; c += 8) {  // 8 independent loads in flight, added in CTA order
    float4 x[8];
#pragma unroll
    for (int u = 0; u < 8; u++) x[u] = *reinterpret_cast<const float4 *>(p + (size_t)(c + u) * stride);
#pragma unroll
    for (int u = 0; u < 8; u++) { sum.x += x[u].x; sum.y += x[u].y; sum.z += x[u].z; sum.w += x[u].w; }
  }
  for (; c < nct; c++) {
    const float4 x = *reinterpret_cast<const float4 *>(p + (size_t)c * stride);
    sum.x += x.x; sum.y += x.y; sum.z += x.z; sum.w += x.w;
  }
  float4 *wp = reinterpret_cast<float4 *>(td.table + (size_t)row * D + 4 * sl);
  float4 o = *wp;
  if (MODE == APPLY_SGD) {  // W -= lr*gt, contracted like the reference's FFMA
    o.x = fmaf(-sum.x, lr, o.x); o.y = fmaf(-sum.y, lr, o.y);
    o.z = fmaf(-sum.z, lr, o.z); o.w = fmaf(-sum.w, lr, o.w);
  } else {
    o.x += sum.x; o.y += sum.y; o.z += sum.z; o.w += sum.w;
  }
  *wp = o;
}

// ---- host side ---------------------------------------------------------------------------------
// tb holds the eligible tables only, tiny ones first (n_tiny of them).
template <int G>
static int run_small(TableBatch &tb, int n_tiny, int D, int64_t B, int aggr, float lr, ApplyMode mode,
                     float *partials, cudaStream_t s) {
  const int T = tb.num_tables, n_small = T - n_tiny;
  const bool avg = aggr != FFEMB_AGGR_SUM;
  // CTAs per table in proportion to its work; bags per CTA even (8-byte index copies stay aligned)
  auto plan = [&](int t0, int t1, int total_ctas, int64_t min_bags, auto cost, uint32_t cta0) {
    double sum = 0;
    for (int t = t0; t < t1; t++) sum += cost(tb.t[t]);
    uint32_t cta = cta0;
    for (int t = t0; t < t1; t++) {
      int64_t want = (int64_t)(total_ctas * (cost(tb.t[t]) / sum) + 0.5);
      if (want < 1) want = 1;
      int64_t bpc = (B + want - 1) / want;
      if (bpc < min_bags) bpc = min_bags;
      bpc = (bpc + kChunkBags - 1) / kChunkBags * kChunkBags;
      tb.t[t].lookup_base = (uint32_t)bpc;
      tb.cta_base[t] = cta;
      cta += (uint32_t)((B + bpc - 1) / bpc);
    }
    return cta;
  };
  const uint32_t c_tiny = plan(0, n_tiny, kTinyCtas, 64, [](const TableDesc &d) { return (double)(d.R + 4); }, 0);
  const uint32_t c_all = plan(n_tiny, T, kSmallCtas, 64, [](const TableDesc &d) { return (double)d.L; }, c_tiny);
  tb.cta_base[T] = c_all;
  uint64_t off = 0;
  int max_r = 1;
  for (int t = 0; t < T; t++) {
    tb.row_base[t] = off;
    off += (uint64_t)(tb.cta_base[t + 1] - tb.cta_base[t]) * tb.t[t].R * D;
    max_r = tb.t[t].R > max_r ? (int)tb.t[t].R : max_r;
  }
  if (off * 4 > small_partials_bound(D)) {
    set_error("small backward: partials exceed the reserved workspace");
    return FFEMB_EWORKSPACE;
  }
  ProfScope prof(FFEMB_PHASE_SMALL, (n_tiny ? 1 : 0) + (n_small ? 1 : 0) + 1, s);
  if (n_tiny) {
    int rmax = 0, lmax = 1;
    for (int t = 0; t < n_tiny; t++) {
      rmax = tb.t[t].R > rmax ? (int)tb.t[t].R : rmax;
      lmax = tb.t[t].L > lmax ? tb.t[t].L : lmax;
    }
    const size_t sm = 2 * (size_t)kTinyChunk * D * 4 + 2 * (size_t)kTinyChunk * lmax * 8;  // >= rmax*D*4
    auto go = [&](auto kern) -> int {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)(2 * kTinyChunk * 128 * 4 + 2 * kTinyChunk * kSmallMaxBag * 8));
      if (e != cudaSuccess) return cuda_fail(e, "tiny backward smem opt-in");
      kern<<<c_tiny, kTinyThreads, sm, s>>>(tb, D, B, partials);
      return FFEMB_OK;
    };
    int rc;
    if (rmax <= 4) rc = avg ? go(bwd_tiny_kernel<G, 4, true>) : go(bwd_tiny_kernel<G, 4, false>);
    else rc = avg ? go(bwd_tiny_kernel<G, kTinyRows, true>) : go(bwd_tiny_kernel<G, kTinyRows, false>);
    if (rc) return rc;
  }
  if (n_small) {
    constexpr int NU = (kSmallThreads / 32) * (32 / G);
    auto fixed_of = [&](int R) {
      const size_t nbin = (size_t)small_bins(R, NU);
      return (size_t)kSmallMaxLookups * 4 + (size_t)(kSmallMaxLookups / 32) * nbin * 4 + (2 * nbin + 2) * 4;
    };
    const size_t fixed = fixed_of(kSmallMaxRows);
    size_t sm = 0;
    for (int t = n_tiny; t < T; t++) {
      const int chb = small_chunk_bags(tb.t[t].L);
      const size_t x = (size_t)tb.t[t].R * D * 4 + 2 * (size_t)chb * D * 4 + 2 * (size_t)chb * tb.t[t].L * 8 +
                       fixed_of((int)tb.t[t].R);
      sm = x > sm ? x : sm;
    }
    auto go = [&](auto kern) -> int {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)(kSmallAccBytes + 2 * kSmallMaxChunk * 128 * 4 + 2 * kSmallMaxLookups * 8 + fixed));
      if (e != cudaSuccess) return cuda_fail(e, "small backward smem opt-in");
      kern<<<c_all - c_tiny, kSmallThreads, sm, s>>>(tb, D, B, partials, c_tiny);
      return FFEMB_OK;
    };
    const int rc = avg ? go(bwd_small_kernel<G, true>) : go(bwd_small_kernel<G, false>);
    if (rc) return rc;
  }
  {
    constexpr int SLOTS = 32 / G;
    dim3 grid((unsigned)((max_r + 8 * SLOTS - 1) / (8 * SLOTS)), T);
    if (mode == APPLY_SGD) small_apply_kernel<G, APPLY_SGD><<<grid, 256, 0, s>>>(tb, D, partials, lr);
    else small_apply_kernel<G, APPLY_GRAD><<<grid, 256, 0, s>>>(tb, D, partials, lr);
  }
  return cuda_fail(cudaGetLastError(), "small backward launch");
}

int launch_backward_small(TableBatch &tb, int n_tiny, int64_t D64, int64_t B, int aggr, float lr,
                          ApplyMode mode, float *partials, cudaStream_t s) {
  const int D = (int)D64, D4 = D / 4;
  if (D4 <= 8) return run_small<8>(tb, n_tiny, D, B, aggr, lr, mode, partials, s);
  if (D4 <= 16) return run_small<16>(tb, n_tiny, D, B, aggr, lr, mode, partials, s);
  return run_small<32>(tb, n_tiny, D, B, aggr, lr, mode, partials, s);
}

}  // namespace ffemb
