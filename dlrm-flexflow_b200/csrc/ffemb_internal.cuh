// ffemb_internal.cuh — shared declarations of libffemb.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "ffemb.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libffemb is written for sm_100a (B200) only"
#endif

#include <type_traits>

namespace ffemb {

// compile-time loop: f(std::integral_constant<int, I>{}) for I in [0, N)
template <int I, int N, typename F>
__device__ __forceinline__ void static_for(F &&f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for<I + 1, N>(f);
  }
}

// Descriptor array passed BY VALUE as a kernel parameter (64 x 40 B = 2.5 KB of the 4 KB
// parameter space): no per-step H2D copy, no device-side descriptor allocation.
struct TableDesc {
  const int64_t *idx;
  float *out;     // pooled output (fwd) / output-grad (bwd)
  float *table;   // weights or dense weight-grad
  int64_t R;
  int32_t L;
  uint32_t lookup_base;  // first lookup number of this table: sum_{s<t} B*L_s (bwd only)
};
struct TableBatch {
  TableDesc t[FFEMB_MAX_TABLES];
  uint64_t row_base[FFEMB_MAX_TABLES + 1];  // first global row id of each table (bwd keys)
  uint32_t cta_base[FFEMB_MAX_TABLES + 1];  // first CTA of each table in the flat forward grid
  int32_t num_tables;
  // bwd: indices < 0 are EMPTY SLOTS (the routed lookups of a row-wise shard, ffemb_backward_sgd_padded):
  // keygen gives them the key `row_base[num_tables]` (one past the last row), they sort to the end
  // and level 1 never loads anything for them
  int32_t padded;
  // bwd: > 0 = the key space is ALIGNED — table t's keys are (t << sort_bits) + row — and every table's
  // stretch of the lookup array is sorted on its own, on its own row bits (fewer radix passes than the
  // global sort when the tables are large and few: launch_backward_sorted decides)
  int32_t sort_bits;
  // floats between consecutive rows of the output-grad buffers (bwd).  D: plain [B][D] buffers;
  // larger: td.out points at the table's column block inside a wider row-major matrix — the
  // gradient of the Concat output (ffemb_backward_sgd_concat)
  int64_t og_stride;
};

// Where pooled rows go.  npeers == 0: td.out + b*D (plain).  Else the fused exchange: sample b
// belongs to rank b / ext and lands in that rank's Concat input (include/ffemb.h).
struct OutMap {
  float *peer[FFEMB_MAX_PEERS];
  int64_t ext, stride;
  int32_t npeers;
};

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);

// Tuning aids (FFEMB_* environment variables, A/B runs only).  Read ONCE per process on first use
// (ffemb_api.cu), never on the launch path; -1 = not set, the built-in default applies.
struct Tuning {
  long long stream_mb, l1_kb;
  int fwd_perm, fwd_avg_div, fwd_target, r1_mix, r1_flags, r1_pipe, r1_ring, small, small_fork,
      up_fused, hot, slab_mb, own_sort, xpush, prescale, fwd_bulk, sort_aligned;
};
const Tuning &tuning();

// profiling hooks (ffemb_api.cu): no-ops unless ffemb_profile_enable(1)
void prof_begin(int phase, cudaStream_t s);
void prof_end(int phase, int launches, cudaStream_t s);
struct ProfScope {
  int phase, launches;
  cudaStream_t s;
  ProfScope(int p, int n, cudaStream_t st) : phase(p), launches(n), s(st) { prof_begin(p, st); }
  ~ProfScope() { prof_end(phase, launches, s); }
};

static inline bool aggr_valid(int aggr) {
  return aggr == FFEMB_AGGR_SUM || aggr == FFEMB_AGGR_AVG || aggr == FFEMB_AGGR_MEAN;
}

// launchers (defined in the .cu files)
int launch_forward(const TableBatch &tb, int64_t D, int64_t B, int aggr, const OutMap &om, cudaStream_t s);
int launch_exchange_backward(const TableBatch &tb, int64_t D, const float *dx, int64_t n, int64_t first,
                             int64_t stride, cudaStream_t s);
// push of parts of mixed width with bulk copies (ffemb_exchange.cu); tb.t[p].lookup_base = column offset
int launch_push_bulk(const TableBatch &tb, const int32_t *widths, const float *dx, int64_t n, int64_t first,
                     int64_t stride, cudaStream_t s, int warps_per_sm = 0);
int launch_backward_atomic(const TableBatch &tb, int64_t D, int64_t B, int aggr, cudaStream_t s);
enum ApplyMode { APPLY_SGD = 0, APPLY_GRAD = 1, APPLY_SGDX = 2 };
// APPLY_SGDX: the reference's full sgd_update (/root/reference/src/runtime/optimizer_kernel.cu:23-41:
// weight decay, momentum, nesterov) on the rows a step touches; the rows it does not touch get the
// same formula with a zero gradient in sgdx_untouched (ffemb_backward.cu).  Lives in device memory
// (tail of the caller's workspace) so the kernels' parameter space stays as it is.
struct OptArgs {
  float lr, wd, mom;
  int32_t nesterov;
  int32_t adam;                         // 0: sgd_update; 1: adam_update (optimizer_kernel.cu:206-226)
  float alpha_t, beta1, beta2, eps;     // Adam: alpha_t is the bias-corrected step of AdamOptimizer::next
  float *V[FFEMB_MAX_TABLES];           // SGD: momentum buffer ([R][D], unused when mom == 0); Adam: V
  float *M[FFEMB_MAX_TABLES];           // Adam: first moment
  uint32_t *touched[FFEMB_MAX_TABLES];  // bit r of table t: row r was updated by the sparse pass
};
// one element of adam_update, operation for operation as nvcc compiles the reference
// (SASS of oracle/_ref/optimizer_kernel_cu.o): gt = FFMA(W, wd, g); t1 = gt*(1-b1); t2 = gt*(1-b2);
// gg = gt*t2; M = FFMA(M, b1, t1); V = FFMA(V, b2, gg); W = W - (M*alpha_t) / (sqrtf(V) + eps)
// with IEEE sqrt and division.  Intrinsics: nothing here may be re-contracted.
__device__ __forceinline__ float adam_elem(float w, float g, float &m, float &v, float alpha_t, float b1,
                                           float b2, float wd, float eps) {
  const float gt = __fmaf_rn(w, wd, g);
  const float t1 = __fmul_rn(gt, __fadd_rn(-b1, 1.0f));
  const float t2 = __fmul_rn(gt, __fadd_rn(-b2, 1.0f));
  const float gg = __fmul_rn(gt, t2);
  m = __fmaf_rn(m, b1, t1);
  v = __fmaf_rn(v, b2, gg);
  return __fadd_rn(w, -__fdiv_rn(__fmul_rn(m, alpha_t), __fadd_rn(__fsqrt_rn(v), eps)));
}
// one element of sgd_update, contracted as nvcc contracts the reference (FFMA each):
//   gt = WGrad + wd*W ; V = V*mom + gt ; gt = nesterov ? gt + mom*V : V ; W -= lr*gt
__device__ __forceinline__ float sgdx_elem(float w, float g, float &v, float lr, float wd, float mom, bool nesterov) {
  float gt = fmaf(wd, w, g);
  if (mom > 0.0f) {
    v = fmaf(v, mom, gt);
    gt = nesterov ? fmaf(mom, v, gt) : v;
  }
  return fmaf(-lr, gt, w);
}
// the 16-byte fragment at wp (table row of table t, key = global row id): general update + mark the row
__device__ __forceinline__ float4 sgdx_fragment(const OptArgs *oa, int t, const float *table, float *wp,
                                                float4 w, float4 g, uint64_t local_row, bool mark) {
  if (oa->adam) {
    const float a = oa->alpha_t, b1 = oa->beta1, b2 = oa->beta2, wd = oa->wd, eps = oa->eps;
    float *mp = oa->M[t] + (wp - table), *vp = oa->V[t] + (wp - table);
    float4 m = *reinterpret_cast<const float4 *>(mp), v = *reinterpret_cast<const float4 *>(vp), o;
    o.x = adam_elem(w.x, g.x, m.x, v.x, a, b1, b2, wd, eps);
    o.y = adam_elem(w.y, g.y, m.y, v.y, a, b1, b2, wd, eps);
    o.z = adam_elem(w.z, g.z, m.z, v.z, a, b1, b2, wd, eps);
    o.w = adam_elem(w.w, g.w, m.w, v.w, a, b1, b2, wd, eps);
    *reinterpret_cast<float4 *>(mp) = m;
    *reinterpret_cast<float4 *>(vp) = v;
    if (mark) atomicOr(oa->touched[t] + (local_row >> 5), 1u << (local_row & 31));
    return o;
  }
  const float lr = oa->lr, wd = oa->wd, mom = oa->mom;
  const bool nes = oa->nesterov != 0;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  float *vp = nullptr;
  if (mom > 0.0f) {
    vp = oa->V[t] + (wp - table);
    v = *reinterpret_cast<const float4 *>(vp);
  }
  float4 o;
  o.x = sgdx_elem(w.x, g.x, v.x, lr, wd, mom, nes);
  o.y = sgdx_elem(w.y, g.y, v.y, lr, wd, mom, nes);
  o.z = sgdx_elem(w.z, g.z, v.z, lr, wd, mom, nes);
  o.w = sgdx_elem(w.w, g.w, v.w, lr, wd, mom, nes);
  if (vp) *reinterpret_cast<float4 *>(vp) = v;
  if (mark) atomicOr(oa->touched[t] + (local_row >> 5), 1u << (local_row & 31));
  return o;
}
// stage: the sort depends on the indices only, so it can be enqueued ahead of the output-grads
enum { STAGE_ALL = 0, STAGE_SORT = 1, STAGE_APPLY = 2 };
int launch_backward_sorted(const TableBatch &tb, int64_t D, int64_t B, int aggr, float lr,
                           ApplyMode mode, void *ws, size_t ws_bytes, cudaStream_t s,
                           int stage = STAGE_ALL, const OptArgs *opt_host = nullptr);
int launch_load_sparse(const ffemb_table_t *tables, int num_tables, const int64_t *full, int64_t row_len,
                       const int32_t *sample_ids, int64_t first_sample, int64_t B, cudaStream_t s);
size_t backward_workspace_bytes(int64_t num_lookups, int64_t total_rows, int64_t D);
// scratch of one table's pre-scaled output-grad (AVG / MEAN backward), 256-byte granules
static inline size_t prescale_bytes(int64_t B, int64_t D) { return ((size_t)B * (size_t)D * 4 + 255) & ~(size_t)255; }
// tables whose whole gradient fits on chip (ffemb_backward_small.cu)
int backward_path(int64_t R, int64_t L, int64_t D, int64_t B);  // 0 sorted, 1 tiny, 2 small
size_t small_partials_bound(int64_t D);
int launch_backward_small(TableBatch &tb, int n_tiny, int64_t D, int64_t B, int aggr, float lr,
                          ApplyMode mode, float *partials, cudaStream_t s);

// ---- device helpers --------------------------------------------------------------------
// 16-byte read-only load of a table row fragment.  Rows are touched once per lookup, so keep
// them out of L1 (L1::no_allocate) and let L2 hold the hot ones.
__device__ __forceinline__ float4 ldg_row(const float *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
// streaming 16-byte store (pooled outputs are consumed by another kernel / another GPU)
__device__ __forceinline__ void st_stream(float *p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
// ---- L2 eviction-priority hints (createpolicy + .L2::cache_hint) ---------------------------
// evict_last : gathered data that is re-read soon (output-grad rows of the table being reduced,
//              rows of tables small enough to live in the 126 MB L2);
// evict_first: data touched once per step (rows of big tables, index streams, pooled outputs),
//              so that it does not push the former out.
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// read-only 16-byte gather with an L2 policy (no L1 allocation)
__device__ __forceinline__ float4 ldg_row_hint(const float *p, uint64_t pol) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p), "l"(pol));
  return v;
}
// coherent 16-byte load / store with an L2 policy (table rows that this kernel also writes)
__device__ __forceinline__ float4 ld_row_hint(const float *p, uint64_t pol) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p), "l"(pol)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_row_hint(float *p, float4 v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p),
               "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol)
               : "memory");
}
// Predicated forms: ONE predicated LDG and no branch, so that a batch of them stays in a
// straight line and ptxas can give consecutive batches their own scoreboard barriers (a load
// wrapped in `if` becomes its own BSSY/BRA region and every such load ends up on one barrier,
// which turns a register double buffer into a serial chain: measured, see profiles/).
__device__ __forceinline__ void ldg_row_hint_if(float4 &v, const float *p, uint64_t pol, bool on) {
  asm volatile(
      "{\n .reg .pred q;\n setp.ne.b32 q, %6, 0;\n"
      " @q ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;\n}"
      : "+f"(v.x), "+f"(v.y), "+f"(v.z), "+f"(v.w)
      : "l"(p), "l"(pol), "r"((int)on));
}
// L1-allocating variant for tables small enough to live in one SM's L1 (a handful of hot rows
// otherwise all hit the same L2 slices: measured 1.1 ms for a 3-row table vs 0.06 ms for a
// 1000-row one at the same lookup count)
__device__ __forceinline__ void ldg_row_l1_if(float4 &v, const float *p, bool on) {
  asm volatile(
      "{\n .reg .pred q;\n setp.ne.b32 q, %5, 0;\n"
      " @q ld.global.nc.L1::evict_last.v4.f32 {%0,%1,%2,%3}, [%4];\n}"
      : "+f"(v.x), "+f"(v.y), "+f"(v.z), "+f"(v.w)
      : "l"(p), "r"((int)on));
}
__device__ __forceinline__ void ld_row_hint_if(float4 &v, const float *p, uint64_t pol, bool on) {
  asm volatile(
      "{\n .reg .pred q;\n setp.ne.b32 q, %6, 0;\n"
      " @q ld.global.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;\n}"
      : "+f"(v.x), "+f"(v.y), "+f"(v.z), "+f"(v.w)
      : "l"(p), "l"(pol), "r"((int)on)
      : "memory");
}
__device__ __forceinline__ void st_row_hint_if(float *p, float4 v, uint64_t pol, bool on) {
  asm volatile(
      "{\n .reg .pred q;\n setp.ne.b32 q, %6, 0;\n"
      " @q st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;\n}" ::"l"(p),
      "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol), "r"((int)on)
      : "memory");
}
// one-instruction L2 prefetch of a whole row (TMA bulk prefetch: no registers, no smem);
// p 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void prefetch_l2_bulk(const void *p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ int64_t ldg_idx_stream(const int64_t *p, uint64_t pol) {
  int64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.b64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
  return v;
}
// a / d, correctly rounded, without the division subroutine: r = RN(1/d), q = RN(a*r), then two
// exact-residual corrections (Markstein).  Valid while no intermediate under/overflows, i.e. for
// every gradient that is not denormal-small; d is a bag width (1 .. 2^31).
__device__ __forceinline__ float div_rn(float a, float d, float r) {
  float q = a * r;
  q = fmaf(fmaf(-d, q, a), r, q);
  q = fmaf(fmaf(-d, q, a), r, q);
  return q;
}
// The same quotient for EVERY input: outside the range where the residual trick is exact (tiny
// values whose residual a - d*q would underflow, zero, inf, nan) the IEEE division takes over — a
// rare, well-predicted branch.  Bit-identical to a / d.
__device__ __forceinline__ float div_exact(float a, float d, float r) {
  float q = div_rn(a, d, r);
  const float m = fabsf(a);
  if (!(m >= 7.9e-31f && m <= 3.0e38f)) q = a / d;
  return q;
}
// The 16-byte fragment form for the backward's gradient scaling (g / in_dim per element,
// /root/reference/src/ops/embedding.cu:210): four Markstein quotients, one combined range test, and
// the IEEE divisions behind a NON-inlined call — inlined per element they bloat the serial walk of the
// segmented sum (measured on cfg5: level 1 12.5 -> 19.2 ms).  A zero addend counts as safe: its
// quotient is a zero whose sign cannot be observed in a sum that starts from +0.
static __device__ __noinline__ float4 f4_div_ieee(float4 a, float d) {
  return make_float4(a.x / d, a.y / d, a.z / d, a.w / d);
}
__device__ __forceinline__ bool div_rn_safe(float a) {
  const float m = fabsf(a);
  return (m >= 7.9e-31f && m <= 3.0e38f) || m == 0.0f;
}
__device__ __forceinline__ float4 f4_div_exact(float4 a, float d, float r) {
  float4 q = make_float4(div_rn(a.x, d, r), div_rn(a.y, d, r), div_rn(a.z, d, r), div_rn(a.w, d, r));
  if (!(div_rn_safe(a.x) && div_rn_safe(a.y) && div_rn_safe(a.z) && div_rn_safe(a.w))) q = f4_div_ieee(a, d);
  return q;
}
// ---- TMA bulk copies (cp.async.bulk) + mbarrier: the push over NVLink (ffemb_exchange.cu) and the bulk
// row-gather experiment of the forward (ffemb_forward.cu) ---------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n .reg .pred P;\n mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n selp.b32 %0, 1, 0, P;\n}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void *smem, const void *gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem)),
               "l"(gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *gmem, const void *smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem), "r"(smem_u32(smem)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

__device__ __forceinline__ int64_t shfl_i64(int64_t v, int src) {
  int lo = __shfl_sync(0xffffffffu, (int)(v & 0xffffffffll), src);
  int hi = __shfl_sync(0xffffffffu, (int)(v >> 32), src);
  return ((int64_t)hi << 32) | (uint32_t)lo;
}

}  // namespace ffemb
