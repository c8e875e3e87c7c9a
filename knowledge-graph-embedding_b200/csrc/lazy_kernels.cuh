// lazy_kernels.cuh — row-lazy replay of the reference's DENSE per-batch table passes.
//
// The reference touches every row of every table on every batch: the start-of-batch constraint
// assigns the normalised / clipped table (TransE.py:171-172, TransD.py constraint)
// and Keras Adam's sparse apply still decays m, v and moves w on ALL rows.
// For a table far larger than a batch (C5: 5 M x 128, 2.56 GB, 15 GB of traffic per step) that
// dense work is the whole cost.  A row that receives no gradient evolves by a closed recurrence,
//      w <- constrain(w);  m <- b1 m;  v <- b2 v;  w <- w - lr_j m / (sqrt(v) + eps)       (step j)
// so it can be REPLAYED later, in registers, when the row is next needed.  `last[row]` is the
// Adam step the stored w / m / v of the row belong to.  Per training step t:
//   k_heads        unique keys of the sorted gradient records -> work list (lazy tables) and
//                  segment table (the small dense tables)
//   k_lazy_pre     before the step kernel: rows on the list are replayed to step t-1 and given the
//                  start-of-batch constraint of step t; only w is written
//   k_lazy_update  after it: m, v replayed to t-1 (pure decays), record sum, clip, Adam step t,
//                  last[row] = t
//   k_lazy_flush   (kge_flush_rows) every row to step t: the table then equals the dense result
// Every operation is the same explicitly rounded arithmetic as the dense kernels
// (update_kernels.cuh), applied in the same order, so lazy and dense agree bit for bit
// (tests/test_gpu_lazy.py).
#pragma once
#include "update_kernels.cuh"

namespace kge {

// lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t) (Keras Adam._prepare_local).  Integer powers by binary
// exponentiation in double: only IEEE multiplies, one sqrt and one divide, so host and device
// produce the same float for every t — the replay needs lr_j of PAST steps on the device.
__host__ __device__ inline float adam_lr(float lr, float b1, float b2, uint64_t t) {
  double p1 = 1.0, p2 = 1.0, a = (double) b1, b = (double) b2;
  for (uint64_t e = t; e; e >>= 1) {
    if (e & 1) { p1 *= a; p2 *= b; }
    a *= a; b *= b;
  }
  const double num = (double) lr * sqrt(1.0 - p2);
  return (float) (num / (1.0 - p1));
}

enum { LAZY_CON_NONE = -1 };
struct LazyTable {
  float *w, *m, *v;
  uint32_t* last;        // [rows] Adam step of the stored row
  const float* rec;      // gradient records of this table (k_step)
  int64_t rows;
  int cols, ld, rec_ld;
  int slot;              // index into scal (clip scale)
  int con_op;            // CON_NORMALIZE / CON_CLIP / LAZY_CON_NONE: start-of-batch constraint of the table
};
struct LazyParams {
  LazyTable tab[2];      // entity tables (ent_emb, and ent_proj for TransD): both keyed by entity id
  int n_tabs;
  const uint32_t* uniq;  // [0] = number of items; item i = uniq[4 + 3 i .. ] = key, lo, hi (sorted positions)
  const uint32_t* vals;  // record ids in sorted-key order (stable)
  const float* scal;
  float lr, beta1, beta2, eps;
  int opt;               // kge_opt_kind.  SGD / Adagrad: a row without a gradient is not moved by the optimizer at all
                         // (w - lr * 0, acc + 0), so its skipped steps replay the constraint only, to its fixed point
  uint32_t t;            // the Adam step being applied (pre: rows go to t-1 + constraint; flush: rows go to t)
};

// ---- work list ---------------------------------------------------------------------------
// One thread per sorted record: the first record of a key finds the end of its segment (galloping
// search: segments are short except for hot rows) and files the key either in the work list of the
// lazy tables (keys < lazy_end, slot by warp-aggregated atomic — the order of the list does not
// matter, every item is an independent row) or in the segment table of the dense tables.
__global__ void __launch_bounds__(256) k_heads(const uint32_t* __restrict__ keys, int64_t n, uint32_t lazy_end,
                                               uint32_t* __restrict__ uniq, uint32_t* __restrict__ seg_lo,
                                               uint32_t* __restrict__ seg_hi) {
  const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  bool head = false;
  uint32_t k = 0;
  if (i < n) { k = keys[i]; head = (i == 0) || (keys[i - 1] != k); }
  int64_t hi = i + 1;
  if (head) {
    int64_t step = 1, lo2 = i;                       // keys[lo2] == k; find first position with a different key
    while (lo2 + step < n && __ldg(keys + lo2 + step) == k) { lo2 += step; step <<= 1; }
    int64_t hi2 = (lo2 + step < n) ? lo2 + step : n; // keys[hi2] != k or hi2 == n
    while (lo2 + 1 < hi2) { const int64_t mid = (lo2 + hi2) >> 1; if (__ldg(keys + mid) == k) lo2 = mid; else hi2 = mid; }
    hi = hi2;
  }
  const bool lz = head && k < lazy_end;
  const unsigned bal = __ballot_sync(FULL, lz);
  if (bal) {
    uint32_t base = 0;
    const int leader = __ffs(bal) - 1;
    if (lane == leader) base = atomicAdd(uniq, (uint32_t) __popc(bal));
    base = __shfl_sync(FULL, base, leader);
    if (lz) {
      uint32_t* it = uniq + 4 + 3 * (size_t) (base + __popc(bal & ((1u << lane) - 1)));
      it[0] = k; it[1] = (uint32_t) i; it[2] = (uint32_t) hi;
    }
  }
  if (head && !lz) { seg_lo[k - lazy_end] = (uint32_t) i; seg_hi[k - lazy_end] = (uint32_t) hi; }
}

// ---- one row in registers ------------------------------------------------------------------
template <int CPL>
__device__ __forceinline__ void load_row(const float* base, int nch, int lane, float4 (&x)[CPL]) {
  const float4* p = reinterpret_cast<const float4*>(base);
#pragma unroll
  for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; x[u] = (c < nch) ? p[c] : f4zero(); }
}
template <int CPL>
__device__ __forceinline__ void store_row(float* base, int nch, int lane, const float4 (&x)[CPL]) {
  float4* p = reinterpret_cast<float4*>(base);
#pragma unroll
  for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; if (c < nch) p[c] = x[u]; }
}
template <int CPL>
__device__ __forceinline__ bool all_zero(const float4 (&m)[CPL], const float4 (&v)[CPL]) {
  bool z = true;
#pragma unroll
  for (int u = 0; u < CPL; ++u)
    z = z && m[u].x == 0.f && m[u].y == 0.f && m[u].z == 0.f && m[u].w == 0.f && v[u].x == 0.f && v[u].y == 0.f &&
        v[u].z == 0.f && v[u].w == 0.f;
  return __all_sync(FULL, z);
}
// the constraint of k_constrain on a row in registers; false = the row was left as it is
template <int CPL>
__device__ __forceinline__ bool con_row(float4 (&w)[CPL], int op, int nch, int lane) {
  if (op == LAZY_CON_NONE) return false;
  const float nrm = sqrtf(row_sumsq<CPL>(w));
  if (op == CON_CLIP && nrm < 1.0f) return false;
  const float den = (op == CON_CLIP) ? fmaxf(nrm, 1e-9f) : nrm;
#pragma unroll
  for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; if (c < nch) w[u] = div4(w[u], den); }
  return den != 1.0f;
}
// Adam step j with a zero gradient: adam1() with g = +0 reduced to the operations that can change
// a bit.  m <- b1 m + (1 - b1) 0 keeps its "+ 0" (it turns a -0 product into +0, as the dense kernel
// does); v >= 0, so b2 v + 0 == b2 v.  Padding elements stay zero (0 / (0 + eps)).
__device__ __forceinline__ void idle1(float& w, float& m, float& v, float b1, float b2, float lr_j, float eps) {
  m = __fadd_rn(__fmul_rn(b1, m), 0.f);
  v = __fmul_rn(b2, v);
  w = __fsub_rn(w, __fdiv_rn(__fmul_rn(lr_j, m), __fadd_rn(sqrtf(v), eps)));
}
template <int CPL>
__device__ __forceinline__ void adam_idle(float4 (&w)[CPL], float4 (&m)[CPL], float4 (&v)[CPL], float b1, float b2,
                                          float lr_j, float eps) {
#pragma unroll
  for (int u = 0; u < CPL; ++u) {
    idle1(w[u].x, m[u].x, v[u].x, b1, b2, lr_j, eps);
    idle1(w[u].y, m[u].y, v[u].y, b1, b2, lr_j, eps);
    idle1(w[u].z, m[u].z, v[u].z, b1, b2, lr_j, eps);
    idle1(w[u].w, m[u].w, v[u].w, b1, b2, lr_j, eps);
  }
}
// m, v of `n` zero-gradient steps (w is replayed elsewhere)
template <int CPL>
__device__ __forceinline__ void decay_mv(float4 (&m)[CPL], float4 (&v)[CPL], uint32_t n, float b1, float b2) {
  for (uint32_t j = 0; j < n; ++j) {
#pragma unroll
    for (int u = 0; u < CPL; ++u) {
      m[u] = f4(__fadd_rn(__fmul_rn(b1, m[u].x), 0.f), __fadd_rn(__fmul_rn(b1, m[u].y), 0.f),
                __fadd_rn(__fmul_rn(b1, m[u].z), 0.f), __fadd_rn(__fmul_rn(b1, m[u].w), 0.f));
      v[u] = f4(__fmul_rn(b2, v[u].x), __fmul_rn(b2, v[u].y), __fmul_rn(b2, v[u].z), __fmul_rn(b2, v[u].w));
    }
  }
}
// lr_j of the most recent steps, one table per CTA (adam_lr is ~100 double-precision instructions)
constexpr int LR_TAB = 256;
__device__ __forceinline__ void fill_lr_tab(float* s_lr, const LazyParams& p) {
  for (int i = threadIdx.x; i < LR_TAB; i += blockDim.x) s_lr[i] = (p.t > (uint32_t) i) ? adam_lr(p.lr, p.beta1, p.beta2, p.t - i) : 0.f;
  __syncthreads();
}
__device__ __forceinline__ float lr_of(const float* s_lr, const LazyParams& p, uint32_t j) {
  const uint32_t back = p.t - j;                       // j <= p.t
  return back < (uint32_t) LR_TAB ? s_lr[back] : adam_lr(p.lr, p.beta1, p.beta2, j);
}
// steps (s, t_end] of an idle row: constraint then zero-gradient Adam, each step.  A row whose m and
// v are all zero (never touched) is not moved by Adam, and once the constraint leaves it unchanged
// it is a fixed point: the loop stops early.
template <int CPL>
__device__ __forceinline__ void replay_row(const LazyParams& p, const LazyTable& T, const float* s_lr, float4 (&w)[CPL],
                                           float4 (&m)[CPL], float4 (&v)[CPL], bool fresh, uint32_t s, uint32_t t_end,
                                           int nch, int lane) {
  for (uint32_t j = s + 1; j <= t_end; ++j) {
    const bool moved = con_row<CPL>(w, T.con_op, nch, lane);
    if (fresh) { if (!moved) break; continue; }
    adam_idle<CPL>(w, m, v, p.beta1, p.beta2, lr_of(s_lr, p, j), p.eps);
  }
  if (!fresh) {
#pragma unroll
    for (int u = 0; u < CPL; ++u) mask_tail(w[u], m[u], v[u], lane + 32 * u, T.cols);
  }
}

// ---- before the step kernel: listed rows -> state of step t-1, constrained for step t -------
template <int CPL>
__global__ void __launch_bounds__(256) k_lazy_pre(const LazyParams p) {
  __shared__ float s_lr[LR_TAB];
  fill_lr_tab(s_lr, p);
  const int lane = threadIdx.x & 31;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  const uint32_t n = __ldg(p.uniq);
  for (int64_t it = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; it < n; it += nw) {
    const int64_t row = __ldg(p.uniq + 4 + 3 * it);
    for (int ti = 0; ti < p.n_tabs; ++ti) {
      const LazyTable& T = p.tab[ti];
      const uint32_t s = __ldg(T.last + row);
      const bool idle = s + 1 < p.t;                       // steps s+1 .. t-1 were skipped
      if (!idle && T.con_op == LAZY_CON_NONE) continue;
      const int nch = (T.cols + 3) >> 2;
      float4 w[CPL], m[CPL], v[CPL];
      load_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);
      if (idle) {
        load_row<CPL>(T.m + row * (int64_t) T.ld, nch, lane, m);
        load_row<CPL>(T.v + row * (int64_t) T.ld, nch, lane, v);
        replay_row<CPL>(p, T, s_lr, w, m, v, p.opt != KGE_OPTIM_ADAM || all_zero<CPL>(m, v), s, p.t - 1, nch, lane);
      }
      con_row<CPL>(w, T.con_op, nch, lane);                // start-of-batch constraint of step t
      store_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);
    }
  }
}

// ---- after it: record sum (dense kernels' order), clip, m / v decays of the skipped steps, Adam ----
template <int CPL>
__global__ void __launch_bounds__(256) k_lazy_update(const LazyParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  const uint32_t n = __ldg(p.uniq);
  const float lr_t = p.opt == KGE_OPTIM_ADAM ? adam_lr(p.lr, p.beta1, p.beta2, p.t) : p.lr;
  for (int64_t it = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; it < n; it += nw) {
    const uint32_t* item = p.uniq + 4 + 3 * it;
    const int64_t row = __ldg(item), lo = __ldg(item + 1), L = (int64_t) __ldg(item + 2) - lo;
    for (int ti = 0; ti < p.n_tabs; ++ti) {
      const LazyTable& T = p.tab[ti];
      const int nch = (T.cols + 3) >> 2;
      float4 w[CPL], m[CPL], v[CPL], g[CPL];
      load_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);     // issued before the dependent record chain
      load_row<CPL>(T.m + row * (int64_t) T.ld, nch, lane, m);
      load_row<CPL>(T.v + row * (int64_t) T.ld, nch, lane, v);
      const uint32_t s = __ldg(T.last + row);
#pragma unroll
      for (int u = 0; u < CPL; ++u) g[u] = f4zero();
      if (L <= UPD_HEAVY) sum_records<CPL>(p.vals, T.rec, T.rec_ld, 0u, lo, lo + L, nch, lane, g);
      else
        for (int w3 = 0; w3 < 8; ++w3) {          // k_update's eight contiguous slices, combined in slice order
          float4 part[CPL];
#pragma unroll
          for (int u = 0; u < CPL; ++u) part[u] = f4zero();
          sum_records<CPL, RecDepth<CPL>::HEAVY>(p.vals, T.rec, T.rec_ld, 0u, lo + L * w3 / 8, lo + L * (w3 + 1) / 8, nch, lane, part);
#pragma unroll
          for (int u = 0; u < CPL; ++u) g[u] += part[u];
        }
      // k_lazy_pre already moved w; m and v still belong to step s: their zero-gradient steps are
      // independent of w
      if (p.opt == KGE_OPTIM_ADAM && s + 1 < p.t && !all_zero<CPL>(m, v)) decay_mv<CPL>(m, v, p.t - 1 - s, p.beta1, p.beta2);
      const float scale = p.scal[T.slot];
#pragma unroll
      for (int u = 0; u < CPL; ++u) {
        const int c = lane + 32 * u;
        if (c >= nch) continue;
        adam_chunk(w[u], m[u], v[u], scale4(scale, g[u]), p.beta1, p.beta2, lr_t, p.eps, p.opt);
        mask_tail(w[u], m[u], v[u], c, T.cols);
      }
      store_row<CPL>(T.m + row * (int64_t) T.ld, nch, lane, m);
      store_row<CPL>(T.v + row * (int64_t) T.ld, nch, lane, v);
      store_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);
      if (lane == 0) T.last[row] = p.t;
    }
  }
}

// ---- kge_flush_rows: every row of the lazy tables to step t ---------------------------------
template <int CPL>
__global__ void __launch_bounds__(256) k_lazy_flush(const LazyParams p) {
  __shared__ float s_lr[LR_TAB];
  fill_lr_tab(s_lr, p);
  const int lane = threadIdx.x & 31;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (int ti = 0; ti < p.n_tabs; ++ti) {
    const LazyTable& T = p.tab[ti];
    const int nch = (T.cols + 3) >> 2;
    for (int64_t row = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < T.rows; row += nw) {
      const uint32_t s = __ldg(T.last + row);
      if (s >= p.t) continue;
      float4 w[CPL], m[CPL], v[CPL];
      load_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);
      load_row<CPL>(T.m + row * (int64_t) T.ld, nch, lane, m);
      load_row<CPL>(T.v + row * (int64_t) T.ld, nch, lane, v);
      const bool fresh = p.opt != KGE_OPTIM_ADAM || all_zero<CPL>(m, v);
      replay_row<CPL>(p, T, s_lr, w, m, v, fresh, s, p.t, nch, lane);
      store_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);
      if (!fresh) {
        store_row<CPL>(T.m + row * (int64_t) T.ld, nch, lane, m);
        store_row<CPL>(T.v + row * (int64_t) T.ld, nch, lane, v);
      }
      if (lane == 0) T.last[row] = p.t;
    }
  }
}

template <int WHICH>   // 0 pre, 1 update, 2 flush
inline int launch_lazy(const LazyParams& lp, int sm_count, cudaStream_t st) {
  int max_ch = 1;
  for (int j = 0; j < lp.n_tabs; ++j) max_ch = ((lp.tab[j].cols + 3) / 4 > max_ch) ? (lp.tab[j].cols + 3) / 4 : max_ch;
  const int cpl = max_ch <= 32 ? 1 : (max_ch <= 64 ? 2 : (max_ch <= 128 ? 4 : 8));
  const unsigned grid = (unsigned) sm_count * (cpl <= 2 ? 8 : 4);
#define KGE_LAZY_LAUNCH(K)                                         \
  switch (cpl) {                                                   \
    case 1: K<1><<<grid, 256, 0, st>>>(lp); break;                 \
    case 2: K<2><<<grid, 256, 0, st>>>(lp); break;                 \
    case 4: K<4><<<grid, 256, 0, st>>>(lp); break;                 \
    default: K<8><<<grid, 256, 0, st>>>(lp); break;                \
  }
  if (WHICH == 0) { KGE_LAZY_LAUNCH(k_lazy_pre) }
  else if (WHICH == 1) { KGE_LAZY_LAUNCH(k_lazy_update) }
  else { KGE_LAZY_LAUNCH(k_lazy_flush) }
#undef KGE_LAZY_LAUNCH
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("lazy-row kernel launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
  return KGE_OK;
}

}  // namespace kge
