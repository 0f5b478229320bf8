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

}  // extern "C"
