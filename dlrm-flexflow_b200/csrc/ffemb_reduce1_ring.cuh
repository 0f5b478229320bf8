// ffemb_reduce1_ring.cuh — level 1 of the segmented sum with the gathers staged through a
// per-lane shared-memory ring (cp.async), included by ffemb_backward.cu.
//
// Same contract and same summation order as reduce1_kernel (tiles of CH1 sorted lookups per slot,
// runs inside a tile applied to the table, runs cut by a tile edge left as head/tail partial rows),
// so results are bit-identical.  What changes is where the in-flight data lives: reduce1_kernel keeps
// two register buffers of UNR gathered fragments (128 registers, 16 warps per SM, and ptxas puts
// both buffers on one scoreboard barrier), which leaves ~50 KB in flight per SM.  Here every lane
// copies ITS 16-byte fragment of the next S entries' output-grad rows (and, for entries that end a
// run, of the table row) with cp.async into a ring it alone reads back — no registers are tied
// up, no cross-lane synchronisation is needed (cp.async.wait_group is per thread), the walk is
// S entries ahead of its adds, and ~40 registers allow 3-5 CTAs per SM.
#pragma once

namespace ffemb {

__device__ __forceinline__ void cp_async16_hint(void *smem, const void *gmem, uint64_t pol) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(smem)),
               "l"(gmem), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// C == 1 only (D <= 128): one 16-byte fragment per lane and row.
template <typename K, int G, int MODE, bool AVG, int S>
__global__ void __launch_bounds__(L1Threads<G>::value)
reduce1_ring_kernel(const __grid_constant__ TableBatch tb, const K *__restrict__ keys,
                    const uint32_t *__restrict__ vals, int64_t n, int D, float lr, K *keys_out,
                    float *rows_out, uint32_t split, const OptArgs *oa, uint16_t *consumed,
                    const uint8_t *__restrict__ hot_marks) {
  constexpr int THREADS = L1Threads<G>::value;
  constexpr int CH1 = Ch1<G>::value;
  constexpr int SLOTS = 32 / G, WARPS = THREADS / 32, EPW = SLOTS * CH1;
  constexpr K INV = KeyTraits<K>::INV;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  EntryDesc *wd = reinterpret_cast<EntryDesc *>(smem_raw) + warp * EPW;
  // ring: [warp][stage][g | w][lane] float4 — the lanes of a warp touch consecutive 16-byte words
  // (conflict-free); a lane only ever reads back what it copied itself
  float4 *ring = reinterpret_cast<float4 *>(smem_raw + sizeof(EntryDesc) * WARPS * EPW) +
                 (size_t)warp * S * 2 * 32 + lane;
  if (blockIdx.x == 0 && threadIdx.x == 0) up_flag(consumed)[0] = 0;  // "chains left for the tree": set by combine_kernel
  const int64_t wbase = ((int64_t)mix_cta(blockIdx.x, gridDim.x, split) * WARPS + warp) * EPW;
  if (wbase >= n) return;  // warp-uniform
  const int D4 = D >> 2;
  if (SLOTS == 1 && hot_marks && hot_marks[wbase / CH1]) {  // the warp's one tile lies inside a hot run: nothing to do
    if (lane == 0) {
      consumed[wbase / CH1] = 0;
      keys_out[2 * (wbase / CH1)] = keys_out[2 * (wbase / CH1) + 1] = INV;
    }
    return;
  }

  // ---- phase 1: descriptors (parallel over lanes) ----
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
    }
    wd[e] = d;
  }
  __syncwarp();

  // ---- phase 2: serial walk per slot, gathers S entries ahead through the ring ----
  const int sub = lane / G, sl = lane % G;
  const int64_t tile = wbase / CH1 + sub;
  const int64_t t0 = tile * CH1;
  if (t0 >= n) return;
  if (sl == 0) consumed[tile] = 0;  // (both partial slots of the tile: not yet taken by combine_kernel)
  if (hot_marks && hot_marks[tile]) {  // the tile lies inside a hot run: hot_sum_kernel adds its entries
    if (sl == 0) keys_out[2 * tile] = keys_out[2 * tile + 1] = INV;
    return;
  }
  const bool own = sl < D4;
  const EntryDesc *dq = wd + sub * CH1;
  int cnt = (int)min((int64_t)CH1, n - t0);
  if (tb.padded) {  // the empty slots are the END of the sorted array: cut the tile at its first one
    int lo = 0, hi = cnt;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (dq[mid].gp) lo = mid + 1; else hi = mid;
    }
    cnt = lo;
    if (cnt == 0) {  // nothing but empty slots: no partial rows
      if (sl == 0) keys_out[2 * tile] = keys_out[2 * tile + 1] = INV;
      return;
    }
  }
  bool first_open = (t0 > 0) && (keys[t0 - 1] == keys[t0]);
  bool pending = false, wrote_head = false;
  const uint64_t pol_keep = policy_evict_last(), pol_stream = policy_evict_first();
  float scale = 1.f, rscale = 1.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);

  auto issue = [&](int e, int stage) {  // one commit group per entry, empty past the end
    if (e < cnt && own) {
      const EntryDesc d = dq[e];
      cp_async16_hint(&ring[(stage * 2) * 32], d.gp + 4 * sl, pol_keep);
      if (d.wp) cp_async16_hint(&ring[(stage * 2 + 1) * 32], d.wp + 4 * sl, pol_stream);
    }
    cp_commit();
  };
#pragma unroll
  for (int e = 0; e < S - 1; e++) issue(e, e);

#pragma unroll 1
  for (int e0 = 0; e0 < cnt; e0 += S) {
#pragma unroll
    for (int u = 0; u < S; u++) {  // stage index is a compile-time constant
      const int e = e0 + u;
      issue(e + S - 1, (u + S - 1) % S);
      cp_wait<S - 1>();
      if (e < cnt) {
        const EntryDesc d = dq[e];
        if (own) {
          float4 x = ring[(u * 2) * 32];
          if (AVG) {  // g / in_dim per element, embedding.cu:210
            const uint32_t t = vals[t0 + e] >> BAG_BITS;
            scale = (float)tb.t[t].L;
            rscale = __frcp_rn(scale);
            x = f4_div_rn(x, scale, rscale);
          }
          f4_add(acc, x);
        }
        pending = true;
        if (d.wp) {
          if (first_open) {  // head partial: the run began in an earlier tile
            if (sl == 0) keys_out[2 * tile] = keys[t0 + e];
            if (own) st_f4(rows_out + (2 * tile) * (int64_t)D + 4 * sl, acc);
            wrote_head = true;
          } else if (own) {  // complete run: the one read-modify-write of this table row
            float4 o = ring[(u * 2 + 1) * 32];
            if (MODE == APPLY_SGD) {  // W -= lr*gt, contracted like the reference's FFMA
              o.x = fmaf(-acc.x, lr, o.x); o.y = fmaf(-acc.y, lr, o.y);
              o.z = fmaf(-acc.z, lr, o.z); o.w = fmaf(-acc.w, lr, o.w);
            } else if (MODE == APPLY_SGDX) {
              const uint64_t key = (uint64_t)keys[t0 + e];
              const int t = find_table(tb, key);
              o = sgdx_fragment(oa, t, tb.t[t].table, d.wp + 4 * sl, o, acc, key - tb.row_base[t], sl == 0);
            } else {
              f4_add(o, acc);
            }
            st_row_hint(d.wp + 4 * sl, o, pol_stream);
          }
          first_open = false;
          pending = false;
          acc = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
  }
  cp_wait<0>();
  if (sl == 0 && !wrote_head) keys_out[2 * tile] = INV;
  if (pending && own) st_f4(rows_out + (2 * tile + 1) * (int64_t)D + 4 * sl, acc);
  if (sl == 0) keys_out[2 * tile + 1] = pending ? keys[t0 + cnt - 1] : INV;
}

}  // namespace ffemb
