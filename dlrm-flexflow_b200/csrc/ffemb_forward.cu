// ffemb_forward.cu — multi-hot gather-and-pool for sm_100a.
//
// Replaces embed_forward (/root/reference/src/ops/embedding.cu:166-190), which runs one thread
// per output float, re-reads the 8-byte index D times per bag and read-modify-writes the output
// in memory once per bag element.  Here:
//   * a "slot" of G = next_pow2(D/4) lanes owns a run of consecutive bags; every lane keeps its
//     16-byte fragment of the pooled row in registers and accumulates j = 0..L-1 SEQUENTIALLY,
//     so every output element sees exactly the reference's order of fp32 additions
//     (bit-identical results);
//   * the slot's indices are one contiguous stretch of the [B][L] index array: read once,
//     coalesced, a chunk ahead, and handed out with width-G warp shuffles; row reads are 16-byte
//     ld.global.nc issued one group ahead of the adds, so loads stay in flight across bag and
//     chunk boundaries (bag width 1, the DLRM default, is simply the case "store after every add");
//   * L2 eviction hints: rows of tables far larger than L2, the index stream and the pooled
//     outputs are evict_first; rows of small tables keep normal priority and stay resident;
//   * all tables of a GPU go in one launch: a flat grid, each CTA finds its table in a prefix
//     table passed in parameter space.
// HBM-bound integer/float streaming work: no tensor cores, no smem staging needed (the index
// stream is perfectly coalesced and the rows are random 64..512-byte gathers).
#include <stdlib.h>

#include "ffemb_internal.cuh"

namespace ffemb {

#ifndef FFEMB_FWD_KBIG
#define FFEMB_FWD_KBIG 2  // row numbers per lane and chunk for G >= 16 (1 halves the unrolled walk: measured 0.830 vs 0.822 ms on cfg3, 6.27 vs 5.89 on cfg5)
#endif
enum { AG_SUM = 0, AG_AVG = 1, AG_MEAN = 2, AG_AVG_P2 = 3 };  // AG_AVG_P2: AVG, every bag width a power of two
static constexpr int kThreads = 256;

__device__ __forceinline__ void f4_add(float4 &a, const float4 &b) {
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
}
__device__ __forceinline__ void f4_div(float4 &a, float d) {
  a.x = a.x / d; a.y = a.y / d; a.z = a.z / d; a.w = a.w / d;  // IEEE division, as the reference
}
__device__ __forceinline__ void f4_mul(float4 &a, float m) {
  a.x *= m; a.y *= m; a.z *= m; a.w *= m;
}
// a / d exactly as IEEE division rounds it, without the division subroutine on the common path
// (the reference divides the accumulator inside the bag loop, embedding.cu:186 — one division per
// lookup and element).  d = bag width, r = RN(1/d).
//   d a power of two : r is exact and a*r is the correctly rounded quotient, denormal results included;
//   else             : Markstein's exact-residual correction (div_rn), valid while the residual
//                      a - d*q does not underflow, i.e. |a| >= 2^-100; smaller non-zero values take
//                      the IEEE division (a rare, well-predicted branch).
__device__ __forceinline__ float avg_div(float a, float d, float r, bool pow2) {
  if (pow2) return a * r;
  return div_exact(a, d, r);  // tiny / zero / inf / nan accumulators take the IEEE division
}
// The general form is NOT inlined: the pipelined walk is unrolled 64 lookups deep, and the residual
// correction + fallback per element made the AVG kernel 268 KB of code — it then ran out of the
// instruction cache (ncu: icc hit rate 54 %, "no instruction" the top stall, profiles/r1_notes.md).
// One call per 16-byte fragment keeps the walk small; power-of-two bag widths (AG_AVG_P2, decided
// per launch) need no call at all.
__device__ __noinline__ void f4_avg_div(float4 &a, float d, float r, bool pow2) {
  a.x = avg_div(a.x, d, r, pow2); a.y = avg_div(a.y, d, r, pow2);
  a.z = avg_div(a.z, d, r, pow2); a.w = avg_div(a.w, d, r, pow2);
}

// ---- the vector path ---------------------------------------------------------------------
// A slot of G lanes owns `bps` consecutive bags of one table = a contiguous run of bps*L
// indices.  The run is cut into chunks of G*K lookups: every lane holds K row numbers of the
// chunk in registers (loaded coalesced, the NEXT chunk's already in flight), and the row loads
// are issued UNR at a time, one group ahead of the adds (two register buffers), so a slot keeps
// UNR..2*UNR independent 16-byte loads per lane in flight without a bubble at bag or chunk
// boundaries.  The adds stay strictly sequential in j per output element (embedding.cu:180-188),
// and a bag's pooled row is stored the moment its last lookup has been added.
template <int G> struct FwdCfg { static constexpr int K = G >= 16 ? FFEMB_FWD_KBIG : (G >= 4 ? 4 : (G == 2 ? 8 : 16)); };
static constexpr int kFwdTargetMax = 128;            // lookups per slot when there is plenty of work
static constexpr int64_t kStreamBytes = 64ll << 20;  // tables above this do not try to stay in L2
static constexpr int64_t kL1Bytes = 128ll << 10;     // tables up to this are read through L1

__host__ __device__ __forceinline__ int fwd_bags_per_slot(int L, int target) {
  return L <= 0 ? target : (L >= target ? 1 : target / L);
}

// The pipelined walk of one slot.  TINY: the table fits one SM's L1, rows are read with L1
// allocation (a handful of hot rows otherwise all queue on the same L2 slices: measured 1.1 ms
// for a 3-row table against 0.06 ms for a 1000-row one at the same lookup count).
// FULL: D4 == G*C, every lane of a slot owns a fragment (no per-lane width check).
// Output pointer of one bag (lane fragment excluded).  Plain: td.out + b*D.  Fused exchange: the
// Concat input of the rank that owns sample b (include/ffemb.h, ffemb_forward_exchange).
__device__ __forceinline__ float *out_row(const TableDesc &td, const OutMap &om, int64_t b, int D) {
  if (om.npeers == 0) return td.out + b * (int64_t)D;
  const int64_t h = b / om.ext;
  return om.peer[h] + (b - h * om.ext) * om.stride + td.lookup_base;
}

template <int G, int C, int AG, int UNR, bool FULL, bool TINY, bool EXCH>
__device__ __forceinline__ void fwd_slot(const int64_t *__restrict__ ip, const float *__restrict__ W,
                                         float *op, int L, int D, int nl, int nl_max, int sl,
                                         uint64_t pol_rows, uint64_t pol_stream, const OutMap &om,
                                         int64_t bag0, uint32_t col, bool check_hi) {
  constexpr int K = FwdCfg<G>::K, CH = G * K, NG = CH / UNR;
  static_assert(CH % UNR == 0 && NG % 2 == 0, "group pipeline needs an even number of groups");
  const int D4 = D >> 2;
  const float Lf = (float)L;
  const float Lr = __frcp_rn(Lf);
  const bool Lpow2 = (L & (L - 1)) == 0;
  bool own[C];  // does this lane hold fragment c of the row?
#pragma unroll
  for (int c = 0; c < C; c++) own[c] = FULL || (sl + c * G < D4);

  uint32_t cur[K], nxt[K];
  uint32_t hi_or = 0u;  // OR of the upper halves of every index read (row numbers travel as 32 bits)
#pragma unroll
  for (int k = 0; k < K; k++) nxt[k] = 0u;
  auto load_idx = [&](uint32_t (&r)[K], int base) {
#pragma unroll
    for (int k = 0; k < K; k++) {
      const int e = base + k * G + sl;
      const int64_t v = e < nl ? ldg_idx_stream(ip + e, pol_stream) : 0ll;
      r[k] = (uint32_t)v;
      hi_or |= (uint32_t)((uint64_t)v >> 32);
    }
  };
  float4 v[2][UNR][C];
  float4 acc[C];
#pragma unroll
  for (int c = 0; c < C; c++) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int i = 0; i < 2; i++)
#pragma unroll
    for (int u = 0; u < UNR; u++)
#pragma unroll
      for (int c = 0; c < C; c++) v[i][u][c] = make_float4(0.f, 0.f, 0.f, 0.f);
  int rem = L;
  // fused exchange: (rank, local sample) of the bag being pooled; op walks that rank's buffer
  int xh = 0;
  int64_t xb = 0;
  if (EXCH) {
    xh = (int)(bag0 / om.ext);
    xb = bag0 - xh * om.ext;
  }

  // branch-free: UNR predicated LDG.128 in a row
  auto issue = [&](int bf, const uint32_t (&r)[K], int grp, int base) {
#pragma unroll
    for (int u = 0; u < UNR; u++) {
      const int e0 = grp * UNR;
      const uint32_t row = __shfl_sync(0xffffffffu, r[(e0 + u) / G], (e0 + u) % G, G);
      const float *rp = W + (int64_t)row * D;
      const bool on = base + e0 + u < nl;
#pragma unroll
      for (int c = 0; c < C; c++) {
        if (TINY) ldg_row_l1_if(v[bf][u][c], rp + 4 * c * G, on && own[c]);
        else ldg_row_hint_if(v[bf][u][c], rp + 4 * c * G, pol_rows, on && own[c]);
      }
    }
  };
  // Measured (profiles/r1_notes.md): making this part branch-free too lets ptxas give the two
  // buffers separate scoreboard barriers, but the extra selects cost more than the finer
  // overlap wins (big-table forward 0.125 ms -> 0.19 ms), so the adds and the store branch.
  auto consume = [&](int bf, int grp, int base) {
#pragma unroll
    for (int u = 0; u < UNR; u++) {
      if (base + grp * UNR + u < nl) {
#pragma unroll
        for (int c = 0; c < C; c++) {
          f4_add(acc[c], v[bf][u][c]);
          if (AG == AG_AVG) f4_avg_div(acc[c], Lf, Lr, Lpow2);  // the in-loop divide of embedding.cu:186
          if (AG == AG_AVG_P2) f4_mul(acc[c], Lr);              // ... exact for a power-of-two width
        }
        if (--rem == 0) {  // last lookup of the bag: store the pooled row
#pragma unroll
          for (int c = 0; c < C; c++) {
            if (AG == AG_MEAN) f4_mul(acc[c], 1.0f / Lf);  // one multiply by 1/len, embedding.cc:302-313
            if (own[c]) st_row_hint(op + 4 * c * G, acc[c], pol_stream);
            acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
          if (EXCH) {
            if (++xb == om.ext) {  // next sample belongs to the next rank
              xb = 0;
              ++xh;
              op = om.peer[xh < om.npeers ? xh : om.npeers - 1] + col + 4 * sl;
            } else {
              op += om.stride;
            }
          } else {
            op += D;
          }
          rem = L;
        }
      }
    }
  };

  load_idx(cur, 0);
  issue(0, cur, 0, 0);
#pragma unroll 1
  for (int base = 0; base < nl_max; base += CH) {
    const bool more = base + CH < nl_max;
    if (more) load_idx(nxt, base + CH);
#pragma unroll
    for (int g = 0; g < NG; g++) {  // fully unrolled: buffer, group and register numbers are constants
      if (g + 1 < NG) issue((g + 1) & 1, cur, g + 1, base);
      else issue(0, nxt, 0, base + CH);  // all predicated off when there is no next chunk
      consume(g & 1, g, base);
    }
#pragma unroll
    for (int k = 0; k < K; k++) cur[k] = nxt[k];
  }
  // The row count is unknown to the single-table entry point (Embedding::forward_kernel has no such
  // argument), so the host cannot rule out rows beyond 2^32 there: an index that does not fit 32 bits
  // (or a negative one) stops the kernel instead of gathering a wrong row silently.
  if (check_hi && hi_or) __trap();
}

template <int G, int C, int AG, int UNR, bool FULL, bool EXCH>
__global__ void __launch_bounds__(kThreads, C >= 2 ? 2 : 4)
fwd_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, int target, int64_t stream_bytes,
           const __grid_constant__ OutMap om, uint32_t pmul, int64_t l1_bytes) {
  constexpr int SLOTS = 32 / G;
  // Work item of this CTA.  The flat grid lists the tables one after the other; CTAs are handed
  // their item through a multiplicative permutation (pmul coprime with the grid size, about
  // 0.618 x grid) so that every wave of resident CTAs is a near-uniform mix of ALL tables: the
  // HBM-bound gathers (tables >> L2), the L2-bound ones (tables resident in L2) and the L1-bound
  // ones (tables of a few rows) then overlap instead of running one regime after the other.
  const uint32_t vb = (uint32_t)(((uint64_t)blockIdx.x * pmul) % gridDim.x);
  // which table does this CTA work on?  cta_base is ascending, <= 65 entries
  int t = 0;
  {
    int lo = 0, hi = tb.num_tables - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (tb.cta_base[mid] <= vb) lo = mid; else hi = mid - 1;
    }
    t = lo;
  }
  const TableDesc &td = tb.t[t];
  const int L = td.L;
  const int bps = fwd_bags_per_slot(L, target);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int64_t slot = ((int64_t)(vb - tb.cta_base[t]) * (kThreads / 32) + warp) * SLOTS + sub;
  const int64_t bag0 = slot * bps;
  const int nb = bag0 < B ? (int)min((int64_t)bps, B - bag0) : 0;
  float *op = (nb > 0 ? out_row(td, om, bag0, D) : td.out) + 4 * sl;
  const uint64_t pol_stream = policy_evict_first();
  if (L == 0) {  // empty bags pool to 0 (out = 0, no lookups)
    for (int i = 0; i < nb; i++) {
      float *zp = out_row(td, om, bag0 + i, D) + 4 * sl;
#pragma unroll
      for (int c = 0; c < C; c++)
        if (sl + c * G < (D >> 2)) st_row_hint(zp + 4 * c * G, make_float4(0.f, 0.f, 0.f, 0.f), pol_stream);
    }
    return;
  }
  const int nl = nb * L;        // valid lookups of this slot
  const int nl_max = bps * L;   // CTA-uniform trip count (the shuffles need every lane)
  const int64_t *ip = td.idx + bag0 * (int64_t)L;
  const float *W = td.table + 4 * sl;
  const int64_t bytes = td.R * (int64_t)D * 4;
  const uint64_t pol_rows = bytes > stream_bytes ? pol_stream : policy_evict_normal();
  if (td.R > 0 && bytes <= l1_bytes)
    fwd_slot<G, C, AG, UNR, FULL, true, EXCH>(ip, W, op, L, D, nl, nl_max, sl, pol_rows, pol_stream, om, bag0, td.lookup_base, false);
  else
    fwd_slot<G, C, AG, UNR, FULL, false, EXCH>(ip, W, op, L, D, nl, nl_max, sl, pol_rows, pol_stream, om, bag0, td.lookup_base, td.R == 0);
}

// ---- AGGR_MODE_AVG as the reference computes it (divide inside the bag loop) ----------------
// One bag per slot, loads batched UNR at a time.  Kept apart from the pipelined kernel: the IEEE
// division is a subroutine call on this target, which does not mix with the branch-free body.
template <int G, int C, int AG>
__global__ void __launch_bounds__(kThreads)
fwd_avg_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, const __grid_constant__ OutMap om) {
  constexpr int SLOTS = 32 / G;
  constexpr int UNR = (C == 1) ? 8 : (C == 2 ? 4 : 2);
  const TableDesc &td = tb.t[blockIdx.y];
  const int L = td.L;
  const int D4 = D >> 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int64_t bag = ((int64_t)blockIdx.x * (kThreads / 32) + warp) * SLOTS + sub;
  const bool valid = bag < B;
  const int64_t b = valid ? bag : (B - 1);  // clamp: shuffles below need every lane
  const int64_t *ip = td.idx + b * (int64_t)L;
  const float *__restrict__ W = td.table;
  const float Lf = (float)L;

  float4 acc[C];
#pragma unroll
  for (int c = 0; c < C; c++) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);

  for (int j0 = 0; j0 < L; j0 += G) {
    const int64_t mine = (j0 + sl < L) ? __ldg(ip + j0 + sl) : 0;
    const int cnt = min(G, L - j0);  // warp-uniform (L is per table = per CTA)
    for (int jj = 0; jj < cnt; jj += UNR) {
      float4 v[UNR][C];
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        if (jj + u < cnt) {
          const int64_t r = shfl_i64(mine, sub * G + jj + u);
          const float *rp = W + r * (int64_t)D;
#pragma unroll
          for (int c = 0; c < C; c++)
            if (sl + c * G < D4) v[u][c] = ldg_row(rp + 4 * (sl + c * G));
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        if (jj + u < cnt) {
#pragma unroll
          for (int c = 0; c < C; c++) {
            if (sl + c * G < D4) {
              f4_add(acc[c], v[u][c]);
              if (AG == AG_AVG) f4_div(acc[c], Lf);  // the in-loop divide of embedding.cu:186
            }
          }
        }
      }
    }
  }
  if (AG == AG_MEAN && L > 0) {
    const float inv = 1.0f / Lf;  // one multiply by 1/len, as src/ops/embedding.cc:302-313
#pragma unroll
    for (int c = 0; c < C; c++) f4_mul(acc[c], inv);
  }
  if (valid) {
    float *op = out_row(td, om, b, D);
#pragma unroll
    for (int c = 0; c < C; c++)
      if (sl + c * G < D4) st_stream(op + 4 * (sl + c * G), acc[c]);
  }
}

// ---- any D / any alignment: one warp per bag, lane per float --------------------------------
__global__ void __launch_bounds__(kThreads)
fwd_generic_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, int ag, const __grid_constant__ OutMap om) {
  const TableDesc &td = tb.t[blockIdx.y];
  const int L = td.L;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t bag = (int64_t)blockIdx.x * (kThreads / 32) + warp;
  if (bag >= B) return;
  const int64_t *ip = td.idx + bag * (int64_t)L;
  const float Lf = (float)L;
  for (int d = lane; d < D; d += 32) {
    float acc = 0.f;
    for (int j = 0; j < L; j++) {
      acc += td.table[__ldg(ip + j) * (int64_t)D + d];
      if (ag == AG_AVG) acc = acc / Lf;
    }
    if (ag == AG_MEAN && L > 0) acc *= (1.0f / Lf);
    out_row(td, om, bag, D)[d] = acc;
  }
}

template <int G, int C, bool FULL>
static void launch_vec2(const TableBatch &tb, int D, int64_t B, int ag, int target, unsigned ctas,
                        const OutMap &om, cudaStream_t s) {
  constexpr int UNR = (C == 1) ? 4 : (C == 2 ? 2 : 1);
  int64_t stream_bytes = kStreamBytes;
  if (tuning().stream_mb >= 0) stream_bytes = tuning().stream_mb << 20;  // tuning aid
  bool p2 = true;  // AVG: every bag width a power of two -> the in-loop divide is one exact multiply
  for (int t = 0; t < tb.num_tables; t++) p2 = p2 && (tb.t[t].L & (tb.t[t].L - 1)) == 0;
  int64_t l1_bytes = kL1Bytes;
  if (tuning().l1_kb >= 0) l1_bytes = tuning().l1_kb << 10;  // tuning aid
  // CTA -> work item permutation: odd-ish multiplier near the golden ratio of the grid, coprime with it
  uint32_t pmul = 1;
  bool perm = tb.num_tables > 1;
  if (tuning().fwd_perm >= 0) perm = tuning().fwd_perm != 0;  // tuning aid
  if (perm && ctas > 2) {
    auto gcd = [](uint32_t a, uint32_t b) { while (b) { const uint32_t r = a % b; a = b; b = r; } return a; };
    pmul = (uint32_t)((double)ctas * 0.6180339887) | 1u;
    while (gcd(pmul, ctas) != 1) pmul += 2;
    pmul %= ctas;
  }
  if (om.npeers == 0) {
    if (ag == AG_SUM) fwd_kernel<G, C, AG_SUM, UNR, FULL, false><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
    else if (ag == AG_AVG && p2) fwd_kernel<G, C, AG_AVG_P2, UNR, FULL, false><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
    else if (ag == AG_AVG) fwd_kernel<G, C, AG_AVG, UNR, FULL, false><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
    else fwd_kernel<G, C, AG_MEAN, UNR, FULL, false><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
  } else {
    if (ag == AG_SUM) fwd_kernel<G, C, AG_SUM, UNR, FULL, true><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
    else if (ag == AG_AVG && p2) fwd_kernel<G, C, AG_AVG_P2, UNR, FULL, true><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
    else if (ag == AG_AVG) fwd_kernel<G, C, AG_AVG, UNR, FULL, true><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
    else fwd_kernel<G, C, AG_MEAN, UNR, FULL, true><<<ctas, kThreads, 0, s>>>(tb, D, B, target, stream_bytes, om, pmul, l1_bytes);
  }
}

template <int G, int C>
static void launch_vec(const TableBatch &tb_in, int D, int64_t B, int ag, const OutMap &om, cudaStream_t s) {
  constexpr int SLOTS = 32 / G;
  // AVG-literal: the pipelined kernel with the in-loop divide done by avg_div; the one-bag-per-slot
  // kernel with the IEEE division subroutine is kept as the measured comparison (FFEMB_FWD_AVG_DIV=1)
  bool avg_old = false;
  if (tuning().fwd_avg_div >= 0) avg_old = tuning().fwd_avg_div != 0;
  if (ag == AG_AVG && avg_old) {
    const int64_t per_cta = (int64_t)(kThreads / 32) * SLOTS;
    dim3 grid((unsigned)((B + per_cta - 1) / per_cta), tb_in.num_tables);
    fwd_avg_kernel<G, C, AG_AVG><<<grid, kThreads, 0, s>>>(tb_in, D, B, om);
    return;
  }
  TableBatch tb = tb_in;
  // lookups per slot: enough slots to fill 148 SMs x 64 warps first, then amortise per-slot cost
  int64_t n = 0;
  for (int t = 0; t < tb.num_tables; t++) n += B * (tb.t[t].L > 0 ? tb.t[t].L : 1);
  int target = 8;
  while (target < kFwdTargetMax && n / (2 * target) >= (int64_t)148 * 64 * SLOTS) target *= 2;
  if (tuning().fwd_target > 0) target = tuning().fwd_target;  // tuning aid
  uint32_t ctas = 0;
  for (int t = 0; t < tb.num_tables; t++) {
    tb.cta_base[t] = ctas;
    const int64_t per_cta = (int64_t)fwd_bags_per_slot(tb.t[t].L, target) * SLOTS * (kThreads / 32);
    ctas += (uint32_t)((B + per_cta - 1) / per_cta);
  }
  tb.cta_base[tb.num_tables] = ctas;
  if (D / 4 == G * C) launch_vec2<G, C, true>(tb, D, B, ag, target, ctas, om, s);
  else launch_vec2<G, C, false>(tb, D, B, ag, target, ctas, om, s);
}


// ---- the measured comparison: rows gathered with TMA bulk copies (FFEMB_FWD_BULK=S) ----------------
// north_star: "TMA tile-gather where ncu shows it wins".  A gathered row is one contiguous piece of
// 4*D bytes at a data-dependent address — no tile, nothing to multicast — so the TMA form of this
// gather is ONE cp.async.bulk per row (instead of D/4 16-byte LDGs), issued by lane 0 of the warp
// that owns the bag into a ring of S stages per warp in shared memory, completion counted on one
// mbarrier per stage; every lane then adds its 16-byte fragment from shared memory.  Adds stay
// sequential in j: bit-identical to fwd_kernel.  SUM, D <= 128, plain output only — it exists to be
// measured against fwd_kernel (profiles/r2_tma_ab.md), which stays the product path.
static constexpr int kBulkWarps = 16;
template <int S>
__global__ void __launch_bounds__(kBulkWarps * 32)
fwd_bulk_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, int target) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float *ring = reinterpret_cast<float *>(smem_raw) + (size_t)warp * S * D;
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + (size_t)kBulkWarps * S * D * 4) + warp * S;
  if (lane == 0) {
    for (int s = 0; s < S; s++) mbar_init(bar + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();
  int t = 0;
  while (t + 1 < tb.num_tables && blockIdx.x >= tb.cta_base[t + 1]) t++;
  const TableDesc &td = tb.t[t];
  const int L = td.L;
  const int bpw = fwd_bags_per_slot(L, target);
  const int64_t b0 = ((int64_t)(blockIdx.x - tb.cta_base[t]) * kBulkWarps + warp) * bpw;
  const int nb = (int)min((int64_t)bpw, B - b0);
  if (nb <= 0) return;
  const int64_t *ip = td.idx + b0 * (int64_t)L;
  const int n = nb * L;
  const uint32_t rowbytes = (uint32_t)D * 4u;
  const bool own = lane < (D >> 2);
  auto issue = [&](int e) {
    if (e < n && lane == 0) {
      const int st = e % S;
      const int64_t r = __ldg(ip + e);
      mbar_expect_tx(bar + st, rowbytes);
      bulk_g2s(ring + (size_t)st * D, td.table + r * (int64_t)D, rowbytes, bar + st);
    }
  };
#pragma unroll
  for (int e = 0; e < S - 1; e++) issue(e);
  uint32_t phases = 0;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  int j = 0;
  int64_t bag = b0;
  if (L == 0) {
    for (int i = 0; i < nb; i++)
      if (own) st_stream(td.out + (b0 + i) * (int64_t)D + 4 * lane, acc);
    return;
  }
  for (int e = 0; e < n; e++) {
    issue(e + S - 1);  // its stage was read by every lane before the __syncwarp that ended iteration e - 1
    const int st = e % S;
    while (!mbar_try_wait(bar + st, (phases >> st) & 1u)) {}
    phases ^= 1u << st;
    if (own) f4_add(acc, *reinterpret_cast<const float4 *>(ring + (size_t)st * D + 4 * lane));
    if (++j == L) {
      if (own) st_stream(td.out + bag * (int64_t)D + 4 * lane, acc);
      acc = make_float4(0.f, 0.f, 0.f, 0.f);
      j = 0;
      bag++;
    }
    __syncwarp();
  }
}

template <int S>
static int launch_forward_bulk(const TableBatch &tb_in, int D, int64_t B, cudaStream_t s) {
  TableBatch tb = tb_in;
  const int target = 128;
  uint32_t ctas = 0;
  for (int t = 0; t < tb.num_tables; t++) {
    tb.cta_base[t] = ctas;
    const int64_t per_cta = (int64_t)fwd_bags_per_slot(tb.t[t].L, target) * kBulkWarps;
    ctas += (uint32_t)((B + per_cta - 1) / per_cta);
  }
  tb.cta_base[tb.num_tables] = ctas;
  const size_t sm = (size_t)kBulkWarps * S * D * 4 + (size_t)kBulkWarps * S * sizeof(uint64_t);
  cudaError_t e = cudaFuncSetAttribute(fwd_bulk_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  if (e != cudaSuccess) return cuda_fail(e, "bulk forward smem opt-in");
  fwd_bulk_kernel<S><<<ctas, kBulkWarps * 32, sm, s>>>(tb, D, B, target);
  return cuda_fail(cudaGetLastError(), "bulk forward launch");
}

int launch_forward(const TableBatch &tb, int64_t D64, int64_t B, int aggr, const OutMap &om, cudaStream_t s) {
  const int D = (int)D64;
  const int ag = aggr == FFEMB_AGGR_SUM ? AG_SUM : (aggr == FFEMB_AGGR_AVG ? AG_AVG : AG_MEAN);
  if (B == 0 || tb.num_tables == 0) return FFEMB_OK;
  ProfScope prof(FFEMB_PHASE_FORWARD, 1, s);
  bool vec = (D % 4 == 0) && D <= 512;
  for (int t = 0; t < tb.num_tables; t++)
    vec = vec && (((uintptr_t)tb.t[t].table | (uintptr_t)(om.npeers ? nullptr : tb.t[t].out)) % 16 == 0) &&
          tb.t[t].R <= (1ll << 32);  // row numbers travel as 32 bits inside the kernel
  if (om.npeers) {  // every destination of the fused exchange must keep 16-byte fragments aligned
    vec = vec && om.stride % 4 == 0;
    for (int h = 0; h < om.npeers; h++) vec = vec && (uintptr_t)om.peer[h] % 16 == 0;
    for (int t = 0; t < tb.num_tables; t++) vec = vec && tb.t[t].lookup_base % 4 == 0;
  }
  if (vec && tuning().fwd_bulk > 0 && om.npeers == 0 && ag == AG_SUM && D <= 128) {
    // FFEMB_FWD_BULK=4|8|16: the TMA bulk-copy row gather, measured against fwd_kernel (A/B runs only)
    if (tuning().fwd_bulk >= 16) return launch_forward_bulk<16>(tb, D, B, s);
    if (tuning().fwd_bulk >= 8) return launch_forward_bulk<8>(tb, D, B, s);
    return launch_forward_bulk<4>(tb, D, B, s);
  }
  if (!vec) {
    dim3 grid((unsigned)((B + kThreads / 32 - 1) / (kThreads / 32)), tb.num_tables);
    fwd_generic_kernel<<<grid, kThreads, 0, s>>>(tb, D, B, ag, om);
  } else {
    const int D4 = D / 4;
    if (D4 <= 1) launch_vec<1, 1>(tb, D, B, ag, om, s);
    else if (D4 <= 2) launch_vec<2, 1>(tb, D, B, ag, om, s);
    else if (D4 <= 4) launch_vec<4, 1>(tb, D, B, ag, om, s);
    else if (D4 <= 8) launch_vec<8, 1>(tb, D, B, ag, om, s);
    else if (D4 <= 16) launch_vec<16, 1>(tb, D, B, ag, om, s);
    else if (D4 <= 32) launch_vec<32, 1>(tb, D, B, ag, om, s);
    else if (D4 <= 64) launch_vec<32, 2>(tb, D, B, ag, om, s);
    else launch_vec<32, 4>(tb, D, B, ag, om, s);
  }
  return cuda_fail(cudaGetLastError(), "forward launch");
}

// ---- output-grad push (the transposed exchange) -----------------------------------------------
// This rank's dx [n][stride] -> column block of table t of every sample goes to the owner's
// output-grad rows [first, first + n): contiguous D-float pieces, read locally (coalesced), written
// over NVLink peer memory with 16-byte stores.  One slot of G lanes per (sample, table).
template <int G, int C>
__global__ void __launch_bounds__(kThreads)
exchange_bwd_kernel(const __grid_constant__ TableBatch tb, int D, const float *__restrict__ dx,
                    int64_t n, int64_t first, int64_t stride) {
  constexpr int SLOTS = 32 / G;
  const TableDesc &td = tb.t[blockIdx.y];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int D4 = D >> 2;
  const uint64_t pol = policy_evict_first();
  constexpr int U = 4;  // samples per slot and iteration: U loads in flight, then U remote stores
  const int64_t step = (int64_t)gridDim.x * (kThreads / 32) * SLOTS;
  for (int64_t s0 = ((int64_t)blockIdx.x * (kThreads / 32) + warp) * SLOTS + sub; s0 < n; s0 += U * step) {
    float4 v[U][C];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t sidx = s0 + u * step;
      if (sidx < n) {
        const float *src = dx + sidx * stride + td.lookup_base;
#pragma unroll
        for (int c = 0; c < C; c++)
          if (sl + c * G < D4) v[u][c] = ldg_row_hint(src + 4 * (sl + c * G), pol);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t sidx = s0 + u * step;
      if (sidx < n) {
        float *dst = td.out + (first + sidx) * (int64_t)D;
#pragma unroll
        for (int c = 0; c < C; c++)
          if (sl + c * G < D4) st_row_hint(dst + 4 * (sl + c * G), v[u][c], pol);
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads)
exchange_bwd_generic_kernel(const __grid_constant__ TableBatch tb, int D, const float *__restrict__ dx,
                            int64_t n, int64_t first, int64_t stride) {
  const TableDesc &td = tb.t[blockIdx.y];
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n * D; i += (int64_t)gridDim.x * kThreads) {
    const int64_t sidx = i / D, d = i - sidx * D;
    td.out[(first + sidx) * D + d] = dx[sidx * stride + td.lookup_base + d];
  }
}

template <int G, int C>
static void launch_xbwd(const TableBatch &tb, int D, const float *dx, int64_t n, int64_t first, int64_t stride,
                        cudaStream_t s) {
  const int64_t per_cta = (int64_t)(kThreads / 32) * (32 / G);
  int64_t bx = (n + per_cta - 1) / per_cta;
  const int64_t cap = (148ll * 8 + tb.num_tables - 1) / tb.num_tables;  // ~8 CTAs per SM in total
  if (bx > cap) bx = cap;
  dim3 grid((unsigned)bx, tb.num_tables);
  exchange_bwd_kernel<G, C><<<grid, kThreads, 0, s>>>(tb, D, dx, n, first, stride);
}

int launch_exchange_backward(const TableBatch &tb, int64_t D64, const float *dx, int64_t n, int64_t first,
                             int64_t stride, cudaStream_t s) {
  const int D = (int)D64;
  if (n == 0 || tb.num_tables == 0) return FFEMB_OK;
  ProfScope prof(FFEMB_PHASE_EXCHANGE, 1, s);
  bool vec = (D % 4 == 0) && D <= 512 && stride % 4 == 0 && (uintptr_t)dx % 16 == 0;
  for (int t = 0; t < tb.num_tables; t++)
    vec = vec && (uintptr_t)tb.t[t].out % 16 == 0 && tb.t[t].lookup_base % 4 == 0;
  if (!vec) {
    dim3 grid(148 * 2, tb.num_tables);
    exchange_bwd_generic_kernel<<<grid, kThreads, 0, s>>>(tb, D, dx, n, first, stride);
  } else {
    const int D4 = D / 4;
    if (D4 <= 1) launch_xbwd<1, 1>(tb, D, dx, n, first, stride, s);
    else if (D4 <= 2) launch_xbwd<2, 1>(tb, D, dx, n, first, stride, s);
    else if (D4 <= 4) launch_xbwd<4, 1>(tb, D, dx, n, first, stride, s);
    else if (D4 <= 8) launch_xbwd<8, 1>(tb, D, dx, n, first, stride, s);
    else if (D4 <= 16) launch_xbwd<16, 1>(tb, D, dx, n, first, stride, s);
    else if (D4 <= 32) launch_xbwd<32, 1>(tb, D, dx, n, first, stride, s);
    else if (D4 <= 64) launch_xbwd<32, 2>(tb, D, dx, n, first, stride, s);
    else launch_xbwd<32, 4>(tb, D, dx, n, first, stride, s);
  }
  return cuda_fail(cudaGetLastError(), "exchange backward launch");
}

}  // namespace ffemb
