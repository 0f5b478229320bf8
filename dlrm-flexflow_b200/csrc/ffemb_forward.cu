// ffemb_forward.cu — multi-hot gather-and-pool for sm_100a.
//
// Replaces embed_forward (/root/reference/src/ops/embedding.cu:166-190), which runs one thread
// per output float, re-reads the 8-byte index D times per bag and read-modify-writes the output
// in memory once per bag element.  Here:
//   * a "slot" of G = next_pow2(D/4) lanes owns one bag; every lane keeps its 16-byte fragment of
//     the pooled row in registers and accumulates j = 0..L-1 SEQUENTIALLY, so every output
//     element sees exactly the reference's order of fp32 additions (bit-identical results);
//   * the bag's indices are read once, coalesced (8*G bytes per slot), and handed out with
//     warp shuffles; row reads are 16-byte ld.global.nc, batched UNR at a time so each lane
//     has UNR independent loads in flight;
//   * bag width 1 (the DLRM default) takes a dedicated path with U bags per slot in flight;
//   * all tables of a GPU go in one launch (blockIdx.y = table, descriptors in param space).
// HBM-bound integer/float streaming work: no tensor cores, no smem staging needed (the index
// stream is perfectly coalesced and the rows are random 64..512-byte gathers).
#include "ffemb_internal.cuh"

namespace ffemb {

enum { AG_SUM = 0, AG_AVG = 1, AG_MEAN = 2 };
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

// ---- general bag width -------------------------------------------------------------------
template <int G, int C, int AG>
__global__ void __launch_bounds__(kThreads)
fwd_pool_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B) {
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
    float *op = td.out + b * (int64_t)D;
#pragma unroll
    for (int c = 0; c < C; c++)
      if (sl + c * G < D4) st_stream(op + 4 * (sl + c * G), acc[c]);
  }
}

// ---- bag width 1: U bags per slot in flight ------------------------------------------------
template <int G, int C, int U>
__global__ void __launch_bounds__(kThreads)
fwd_onehot_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B) {
  constexpr int SLOTS = 32 / G;
  constexpr int NB = SLOTS * U;  // bags per warp, <= 32
  static_assert(NB <= 32, "one index per lane");
  const TableDesc &td = tb.t[blockIdx.y];
  const int D4 = D >> 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int64_t bag0 = ((int64_t)blockIdx.x * (kThreads / 32) + warp) * NB;
  if (bag0 >= B) return;  // warp-uniform
  const int64_t mine = (lane < NB && bag0 + lane < B) ? __ldg(td.idx + bag0 + lane) : 0;
  const float *__restrict__ W = td.table;
  float4 v[U][C];
#pragma unroll
  for (int u = 0; u < U; u++) {
    const int64_t r = shfl_i64(mine, u * SLOTS + sub);
    const float *rp = W + r * (int64_t)D;
#pragma unroll
    for (int c = 0; c < C; c++)
      if (sl + c * G < D4) v[u][c] = ldg_row(rp + 4 * (sl + c * G));
  }
#pragma unroll
  for (int u = 0; u < U; u++) {
    const int64_t bag = bag0 + u * SLOTS + sub;
    if (bag < B) {
      float *op = td.out + bag * (int64_t)D;
#pragma unroll
      for (int c = 0; c < C; c++) {
        if (sl + c * G < D4) {
          float4 a = make_float4(0.f, 0.f, 0.f, 0.f);  // out = 0; out += w  (-0 -> +0 like the ref)
          f4_add(a, v[u][c]);
          st_stream(op + 4 * (sl + c * G), a);
        }
      }
    }
  }
}

// ---- any D / any alignment: one warp per bag, lane per float --------------------------------
__global__ void __launch_bounds__(kThreads)
fwd_generic_kernel(const __grid_constant__ TableBatch tb, int D, int64_t B, int ag) {
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
    td.out[bag * (int64_t)D + d] = acc;
  }
}

template <int G, int C>
static void launch_vec(const TableBatch &tb, int D, int64_t B, int ag, bool onehot, cudaStream_t s) {
  constexpr int SLOTS = 32 / G;
  if (onehot) {
    constexpr int U = (C == 1) ? (G < 8 ? G : 8) : (C == 2 ? 4 : 2);
    const int64_t per_cta = (int64_t)(kThreads / 32) * SLOTS * U;
    dim3 grid((unsigned)((B + per_cta - 1) / per_cta), tb.num_tables);
    fwd_onehot_kernel<G, C, U><<<grid, kThreads, 0, s>>>(tb, D, B);
    return;
  }
  const int64_t per_cta = (int64_t)(kThreads / 32) * SLOTS;
  dim3 grid((unsigned)((B + per_cta - 1) / per_cta), tb.num_tables);
  switch (ag) {
    case AG_SUM: fwd_pool_kernel<G, C, AG_SUM><<<grid, kThreads, 0, s>>>(tb, D, B); break;
    case AG_AVG: fwd_pool_kernel<G, C, AG_AVG><<<grid, kThreads, 0, s>>>(tb, D, B); break;
    default: fwd_pool_kernel<G, C, AG_MEAN><<<grid, kThreads, 0, s>>>(tb, D, B); break;
  }
}

int launch_forward(const TableBatch &tb, int64_t D64, int64_t B, int aggr, cudaStream_t s) {
  const int D = (int)D64;
  const int ag = aggr == FFEMB_AGGR_SUM ? AG_SUM : (aggr == FFEMB_AGGR_AVG ? AG_AVG : AG_MEAN);
  if (B == 0 || tb.num_tables == 0) return FFEMB_OK;
  ProfScope prof(FFEMB_PHASE_FORWARD, 1, s);
  bool vec = (D % 4 == 0) && D <= 512;
  bool onehot = true;
  for (int t = 0; t < tb.num_tables; t++) {
    vec = vec && (((uintptr_t)tb.t[t].table | (uintptr_t)tb.t[t].out) % 16 == 0);
    onehot = onehot && tb.t[t].L == 1;
  }
  if (!vec) {
    dim3 grid((unsigned)((B + kThreads / 32 - 1) / (kThreads / 32)), tb.num_tables);
    fwd_generic_kernel<<<grid, kThreads, 0, s>>>(tb, D, B, ag);
  } else {
    const int D4 = D / 4;
    if (D4 <= 1) launch_vec<1, 1>(tb, D, B, ag, onehot, s);
    else if (D4 <= 2) launch_vec<2, 1>(tb, D, B, ag, onehot, s);
    else if (D4 <= 4) launch_vec<4, 1>(tb, D, B, ag, onehot, s);
    else if (D4 <= 8) launch_vec<8, 1>(tb, D, B, ag, onehot, s);
    else if (D4 <= 16) launch_vec<16, 1>(tb, D, B, ag, onehot, s);
    else if (D4 <= 32) launch_vec<32, 1>(tb, D, B, ag, onehot, s);
    else if (D4 <= 64) launch_vec<32, 2>(tb, D, B, ag, onehot, s);
    else launch_vec<32, 4>(tb, D, B, ag, onehot, s);
  }
  return cuda_fail(cudaGetLastError(), "forward launch");
}

}  // namespace ffemb
