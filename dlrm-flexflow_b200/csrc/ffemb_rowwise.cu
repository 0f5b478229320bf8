// ffemb_rowwise.cu — row-wise sharded tables (a table too large for one GPU's HBM).
//
// Not in the reference: Embedding::create_output_and_partition asserts that only the sample
// dimension is partitioned (/root/reference/src/ops/embedding.cu:84-85).  The shard map adopted
// mirrors Legion's blocked restriction (src/runtime/model.cc:880-892): rows_per_shard = ceil(R/G),
// owner(row) = row / rows_per_shard, local row = row % rows_per_shard (sharding.py, oracle).
//
// Forward: every rank pools ITS samples and reads each row straight from the shard of the GPU
// that owns it, over NVLink peer memory (16-byte loads, several rows in flight per lane).  No
// index exchange, no partial sums: the adds happen on the sample's rank in j order, exactly as
// on one GPU, so the pooled rows are bit-identical to the unsharded result.
// Backward: the owner of a shard runs the ordinary deterministic backward over the whole batch
// with indices made local by rowwise_localize_kernel: a lookup of a foreign row is pointed at a
// scratch row appended to the shard (row rows_per_shard), which nobody reads.
#include <string.h>

#include "ffemb_internal.cuh"

namespace ffemb {

struct RowShards {
  const float *shard[FFEMB_MAX_PEERS];  // shard g = rows [g*rps, min(R, (g+1)*rps)), peer-mapped
  int64_t rps;
  int32_t npeers;
};

static constexpr int kRwThreads = 256;

template <int G, int C, int AG>
__global__ void __launch_bounds__(kRwThreads)
fwd_rowwise_kernel(const int64_t *__restrict__ idx, float *__restrict__ out,
                   const __grid_constant__ RowShards rs, int L, int D, int64_t B) {
  constexpr int SLOTS = 32 / G;
  constexpr int UNR = (C == 1) ? 8 : (C == 2 ? 4 : 2);
  const int D4 = D >> 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, sl = lane % G;
  const int64_t bag = ((int64_t)blockIdx.x * (kRwThreads / 32) + warp) * SLOTS + sub;
  const bool valid = bag < B;
  const int64_t b = valid ? bag : (B - 1);  // clamp: the shuffles below need every lane
  const int64_t *ip = idx + b * (int64_t)L;
  const float Lf = (float)L;
  float4 acc[C];
#pragma unroll
  for (int c = 0; c < C; c++) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int j0 = 0; j0 < L; j0 += G) {
    const int64_t mine = (j0 + sl < L) ? __ldg(ip + j0 + sl) : 0;
    const int cnt = min(G, L - j0);
    for (int jj = 0; jj < cnt; jj += UNR) {
      float4 v[UNR][C];
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        if (jj + u < cnt) {
          const int64_t r = shfl_i64(mine, sub * G + jj + u);
          const int64_t g = r / rs.rps;
          const float *rp = rs.shard[g] + (r - g * rs.rps) * (int64_t)D;
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
              acc[c].x += v[u][c].x; acc[c].y += v[u][c].y; acc[c].z += v[u][c].z; acc[c].w += v[u][c].w;
              if (AG == 1) {  // AGGR_MODE_AVG as the reference computes it, embedding.cu:186
                acc[c].x = acc[c].x / Lf; acc[c].y = acc[c].y / Lf;
                acc[c].z = acc[c].z / Lf; acc[c].w = acc[c].w / Lf;
              }
            }
          }
        }
      }
    }
  }
  if (AG == 2 && L > 0) {
    const float inv = 1.0f / Lf;
#pragma unroll
    for (int c = 0; c < C; c++) { acc[c].x *= inv; acc[c].y *= inv; acc[c].z *= inv; acc[c].w *= inv; }
  }
  if (valid) {
    float *op = out + b * (int64_t)D;
#pragma unroll
    for (int c = 0; c < C; c++)
      if (sl + c * G < D4) st_stream(op + 4 * (sl + c * G), acc[c]);
  }
}

template <int G, int C>
static void launch_rw(const int64_t *idx, float *out, const RowShards &rs, int L, int D, int64_t B, int ag,
                      cudaStream_t s) {
  const int64_t per_cta = (int64_t)(kRwThreads / 32) * (32 / G);
  const unsigned grid = (unsigned)((B + per_cta - 1) / per_cta);
  if (ag == 0) fwd_rowwise_kernel<G, C, 0><<<grid, kRwThreads, 0, s>>>(idx, out, rs, L, D, B);
  else if (ag == 1) fwd_rowwise_kernel<G, C, 1><<<grid, kRwThreads, 0, s>>>(idx, out, rs, L, D, B);
  else fwd_rowwise_kernel<G, C, 2><<<grid, kRwThreads, 0, s>>>(idx, out, rs, L, D, B);
}

// idx [n] global row ids -> local [n]: row % rps on the owner, the scratch row `rps` elsewhere.
// Also clears the scratch row of the shard (it collected last step's foreign gradients).
__global__ void __launch_bounds__(256)
rowwise_localize_kernel(const int64_t *__restrict__ idx, int64_t *__restrict__ local, int64_t n,
                        int64_t rps, int64_t me, float *scratch_row, int D) {
  if (blockIdx.x == 0 && scratch_row)
    for (int d = threadIdx.x; d < D; d += 256) scratch_row[d] = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const int64_t r = idx[i], g = r / rps;
    local[i] = g == me ? r - g * rps : rps;
  }
}


// ---- backward: route every lookup of THIS rank's samples to the owner of its row ------
// Owner g keeps, per table, a slot array laid out like the GLOBAL index tensor [G*ext][L] and an
// output-grad buffer [G*ext][D]; rank h = `rank` fills the slots and rows of ITS samples
// (global bags [h*ext, h*ext + n)) on every owner, over NVLink peer memory:
//     slots_g[t][(h*ext + b)*L + j] = idx[b][j] - g*rps   if g owns the row, else -1 (empty slot)
//     rows_g [t][(h*ext + b)][:]    = dx[b][col_t : col_t + D]   on the owners of >= 1 lookup of the bag
// Positions are a function of (rank, b, j) only — no counting, no count exchange — and the owner's
// sorted backward meets equal rows in global (b, j) order, the single-GPU order.  One CTA = a run of
// whole bags (about 256 lookups): phase A, one thread per lookup, writes the slot on every owner
// (coalesced 8-byte stores) and ORs the owner into the bag's mask in shared memory; phase B, one warp
// per bag, loads the bag's output-grad row once and stores it to every owner in the mask.
struct RouteTable {
  const int64_t *idx;   // [n][L]
  int64_t rps;
  int64_t slot_base;    // first slot of the table in the owners' slot arrays (entries)
  int64_t row_base;     // first row of the table in the owners' routed output-grad buffers
  int32_t L, col;
  uint32_t cta_base;    // first CTA of the table in the flat grid
  int32_t bags_per_cta;
};
struct RouteArgs {
  RouteTable t[FFEMB_MAX_TABLES];
  int64_t *slots[FFEMB_MAX_PEERS];
  float *rows[FFEMB_MAX_PEERS];
  const float *dx;
  int64_t dx_stride, n, first;  // samples of this rank, global number of its first sample
  int32_t num_tables, npeers, D;
};

__global__ void __launch_bounds__(256) rowwise_route_kernel(const __grid_constant__ RouteArgs a) {
  __shared__ uint32_t mask[256];
  int t = 0;
  while (t + 1 < a.num_tables && blockIdx.x >= a.t[t + 1].cta_base) t++;  // <= 64 tables
  const RouteTable &rt = a.t[t];
  const int L = rt.L;
  const int64_t b0 = (int64_t)(blockIdx.x - rt.cta_base) * rt.bags_per_cta;
  const int nb = (int)min((int64_t)rt.bags_per_cta, a.n - b0);
  if (nb <= 0) return;
  for (int i = threadIdx.x; i < nb; i += 256) mask[i] = 0u;
  __syncthreads();
  // phase A
  const int64_t nl = (int64_t)nb * L;
  for (int64_t i = threadIdx.x; i < nl; i += 256) {
    const int64_t r = __ldg(rt.idx + b0 * L + i);
    const int64_t g = r / rt.rps;
    const int64_t slot = rt.slot_base + (a.first + b0) * L + i;
#pragma unroll
    for (int h = 0; h < FFEMB_MAX_PEERS; h++)
      if (h < a.npeers) a.slots[h][slot] = h == g ? r - g * rt.rps : -1;
    atomicOr(&mask[(int)(i / L)], 1u << (int)g);
  }
  __syncthreads();
  // phase B
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int D4 = a.D >> 2;
  for (int bb = warp; bb < nb; bb += 8) {
    const uint32_t m = mask[bb];
    const float *src = a.dx + (b0 + bb) * a.dx_stride + rt.col;
    const int64_t drow = rt.row_base + a.first + b0 + bb;
    for (int c = lane; c < D4; c += 32) {
      const float4 v = ldg_row(src + 4 * c);
#pragma unroll
      for (int h = 0; h < FFEMB_MAX_PEERS; h++)
        if ((m >> h) & 1u) st_stream(a.rows[h] + drow * a.D + 4 * c, v);
    }
  }
}

}  // namespace ffemb

using namespace ffemb;

extern "C" {

int ffemb_forward_rowwise(const int64_t *idx, float *out, const float *const *peer_shards,
                          int num_peers, int64_t rows_per_shard, int64_t L, int64_t D, int64_t B,
                          int aggr, ffemb_stream_t stream) {
  if (!aggr_valid(aggr)) {
    set_error("invalid aggr %d", aggr);
    return FFEMB_EINVAL;
  }
  if (num_peers < 1 || num_peers > FFEMB_MAX_PEERS || !peer_shards || rows_per_shard < 1 || L < 0 ||
      L >= (1ll << 31) || D <= 0 || D % 4 != 0 || D > 512 || B < 0 || (B > 0 && (!out || (L > 0 && !idx)))) {
    set_error("forward_rowwise: invalid arguments (D must be a multiple of 4, <= 512)");
    return FFEMB_EINVAL;
  }
  RowShards rs;
  rs.rps = rows_per_shard;
  rs.npeers = num_peers;
  for (int g = 0; g < FFEMB_MAX_PEERS; g++) rs.shard[g] = g < num_peers ? peer_shards[g] : nullptr;
  for (int g = 0; g < num_peers; g++) {
    if (!rs.shard[g] || (uintptr_t)rs.shard[g] % 16 || (uintptr_t)out % 16) {
      set_error("forward_rowwise: shard pointers and the output must be 16-byte aligned, non-NULL");
      return FFEMB_EINVAL;
    }
  }
  if (B == 0) return FFEMB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int ag = aggr == FFEMB_AGGR_SUM ? 0 : (aggr == FFEMB_AGGR_AVG ? 1 : 2);
  ProfScope prof(FFEMB_PHASE_FORWARD, 1, s);
  const int D4 = (int)D / 4;
  if (D4 <= 1) launch_rw<1, 1>(idx, out, rs, (int)L, (int)D, B, ag, s);
  else if (D4 <= 2) launch_rw<2, 1>(idx, out, rs, (int)L, (int)D, B, ag, s);
  else if (D4 <= 4) launch_rw<4, 1>(idx, out, rs, (int)L, (int)D, B, ag, s);
  else if (D4 <= 8) launch_rw<8, 1>(idx, out, rs, (int)L, (int)D, B, ag, s);
  else if (D4 <= 16) launch_rw<16, 1>(idx, out, rs, (int)L, (int)D, B, ag, s);
  else if (D4 <= 32) launch_rw<32, 1>(idx, out, rs, (int)L, (int)D, B, ag, s);
  else if (D4 <= 64) launch_rw<32, 2>(idx, out, rs, (int)L, (int)D, B, ag, s);
  else launch_rw<32, 4>(idx, out, rs, (int)L, (int)D, B, ag, s);
  return cuda_fail(cudaGetLastError(), "forward_rowwise launch");
}

int ffemb_rowwise_localize(const int64_t *idx, int64_t *local_idx, int64_t n, int64_t rows_per_shard,
                           int shard, float *scratch_row, int64_t D, ffemb_stream_t stream) {
  if (n < 0 || rows_per_shard < 1 || shard < 0 || (n > 0 && (!idx || !local_idx)) || D < 0) {
    set_error("rowwise_localize: invalid arguments");
    return FFEMB_EINVAL;
  }
  if (n == 0 && !scratch_row) return FFEMB_OK;
  int64_t blocks = (n + 255) / 256;
  blocks = blocks < 1 ? 1 : (blocks > 148 * 8 ? 148 * 8 : blocks);
  ProfScope prof(FFEMB_PHASE_KEYGEN, 1, (cudaStream_t)stream);
  rowwise_localize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(idx, local_idx, n, rows_per_shard,
                                                                               shard, scratch_row, (int)D);
  return cuda_fail(cudaGetLastError(), "rowwise_localize launch");
}


int ffemb_rowwise_route(const ffemb_table_t *tables, int num_tables, int64_t D, int64_t n_samples,
                        const float *dx, int64_t row_stride, const ffemb_route_t *rt, ffemb_stream_t stream) {
  if (!tables || num_tables < 0 || num_tables > FFEMB_MAX_TABLES || D <= 0 || D % 4 != 0 || n_samples < 0 ||
      !rt || rt->num_peers < 1 || rt->num_peers > FFEMB_MAX_PEERS || rt->rank < 0 || rt->rank >= rt->num_peers ||
      rt->samples_per_peer < n_samples || (n_samples > 0 && !dx) || row_stride < D || row_stride % 4 != 0 ||
      (uintptr_t)dx % 16 != 0) {
    set_error("rowwise_route: invalid arguments (D a multiple of 4, dx 16-byte aligned, n_samples <= samples_per_peer)");
    return FFEMB_EINVAL;
  }
  RouteArgs a;
  memset(&a, 0, sizeof(a));
  for (int h = 0; h < rt->num_peers; h++) {
    if (!rt->peer_slots[h] || !rt->peer_rows[h] || (uintptr_t)rt->peer_rows[h] % 16 != 0) {
      set_error("rowwise_route: peer_slots[%d] / peer_rows[%d] NULL or misaligned", h, h);
      return FFEMB_EINVAL;
    }
    a.slots[h] = rt->peer_slots[h];
    a.rows[h] = rt->peer_rows[h];
  }
  const int64_t ext = rt->samples_per_peer, G = rt->num_peers;
  int64_t slot_base = 0, row_base = 0, ctas = 0;
  for (int t = 0; t < num_tables; t++) {
    const ffemb_table_t &x = tables[t];
    if (x.L < 0 || x.R < 1 || x.reserved < 0 || x.reserved % 4 != 0 || (int64_t)x.reserved + D > row_stride ||
        (n_samples > 0 && x.L > 0 && !x.idx)) {
      set_error("rowwise_route: table %d: invalid descriptor (R = rows of the WHOLE table, reserved = column of "
                "its block in dx)", t);
      return FFEMB_EINVAL;
    }
    RouteTable &r = a.t[t];
    r.idx = x.idx;
    r.rps = (x.R + G - 1) / G;
    r.slot_base = slot_base;
    r.row_base = row_base;
    r.L = x.L;
    r.col = x.reserved;
    r.cta_base = (uint32_t)ctas;
    r.bags_per_cta = x.L <= 0 ? 256 : (x.L >= 256 ? 1 : 256 / x.L);
    slot_base += G * ext * x.L;
    row_base += G * ext;
    ctas += x.L > 0 ? (n_samples + r.bags_per_cta - 1) / r.bags_per_cta : 0;
    if (ctas >= (1ll << 31)) {
      set_error("rowwise_route: too many lookups for one call");
      return FFEMB_EINVAL;
    }
  }
  a.dx = dx;
  a.dx_stride = row_stride;
  a.n = n_samples;
  a.first = (int64_t)rt->rank * ext;
  a.num_tables = num_tables;
  a.npeers = rt->num_peers;
  a.D = (int)D;
  if (ctas == 0) return FFEMB_OK;
  ProfScope prof(FFEMB_PHASE_EXCHANGE, 1, (cudaStream_t)stream);
  rowwise_route_kernel<<<(unsigned)ctas, 256, 0, (cudaStream_t)stream>>>(a);
  return cuda_fail(cudaGetLastError(), "rowwise_route launch");
}

}  // extern "C"
