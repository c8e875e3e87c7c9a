// update_kernels.cuh — constraint, partial reduction and row update kernels.
// All are model independent: they see tables, records and sorted keys only.
#pragma once
#include <stdlib.h>
#include "common.cuh"
#include "reduce_tail.cuh"

namespace kge {

// ---- start-of-batch constraint (constraint.py:26-31 normalise, :93-99 clip) --------
// One warp per row; the row is read twice (norm, then scale) — second read hits L1.
enum { CON_NORMALIZE = 0, CON_CLIP = 1 };
struct ConstrainJob {
  float* w;
  int64_t rows;
  int cols, ld, op;
};
struct ConstrainParams {
  ConstrainJob job[KGE_MAX_TABLES];
  int n_jobs;
  int64_t total_rows;
};

// The constraint arithmetic, one row held by a warp as CPL chunks per lane: sum of squares in
// chunk order with explicitly rounded operations (no context-dependent FMA contraction), so the
// stand-alone constraint pass and the update kernels' shadow output agree bit for bit.
template <int CPL>
__device__ __forceinline__ float row_sumsq(const float4 (&v)[CPL]) {
  float ss = 0.f;
#pragma unroll
  for (int u = 0; u < CPL; ++u) {
    ss = __fmaf_rn(v[u].x, v[u].x, ss); ss = __fmaf_rn(v[u].y, v[u].y, ss);
    ss = __fmaf_rn(v[u].z, v[u].z, ss); ss = __fmaf_rn(v[u].w, v[u].w, ss);
  }
  return warp_sum(ss);
}
__device__ __forceinline__ float4 div4(float4 a, float den) {
  return f4(__fdiv_rn(a.x, den), __fdiv_rn(a.y, den), __fdiv_rn(a.z, den), __fdiv_rn(a.w, den));   // X / norm * 1
}

// CPL = float4 chunks per lane (row width <= 128 * CPL floats); a warp takes RB consecutive rows and
// issues all their loads before the first reduction, so the DRAM latency is paid once per batch.
template <int CPL>
__global__ void __launch_bounds__(256) k_constrain(const ConstrainParams p) {
  constexpr int RB = CPL <= 2 ? 4 : (CPL == 4 ? 2 : 1);
  const int lane = threadIdx.x & 31;
  int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (int64_t g0 = gw * RB; g0 < p.total_rows; g0 += nw * RB) {
    float4 v[RB][CPL];
    float4* base[RB];
    int nch[RB], op[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      int64_t row = g0 + r;
      base[r] = nullptr; nch[r] = 0; op[r] = CON_NORMALIZE;
      if (row < p.total_rows) {
        int j = 0;
        while (j < p.n_jobs - 1 && row >= p.job[j].rows) { row -= p.job[j].rows; ++j; }
        const ConstrainJob& J = p.job[j];
        base[r] = reinterpret_cast<float4*>(J.w + row * (int64_t) J.ld);
        nch[r] = (J.cols + 3) >> 2;
        op[r] = J.op;
      }
#pragma unroll
      for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; v[r][u] = (c < nch[r]) ? base[r][c] : f4zero(); }
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const float nrm = sqrtf(row_sumsq<CPL>(v[r]));      // pow(sum(pow(|x|, 2)), 1/2)
      if (base[r] == nullptr || (op[r] == CON_CLIP && nrm < 1.0f)) continue;
      const float den = (op[r] == CON_CLIP) ? fmaxf(nrm, 1e-9f) : nrm;
#pragma unroll
      for (int u = 0; u < CPL; ++u) {
        const int c = lane + 32 * u;
        if (c < nch[r]) base[r][c] = div4(v[r][u], den);
      }
    }
  }
}
inline int launch_constrain_kernel(const ConstrainParams& cp, int sm_count, cudaStream_t st) {
  int max_ch = 1;
  for (int j = 0; j < cp.n_jobs; ++j) max_ch = ((cp.job[j].cols + 3) / 4 > max_ch) ? (cp.job[j].cols + 3) / 4 : max_ch;
  const int cpl = max_ch <= 32 ? 1 : (max_ch <= 64 ? 2 : (max_ch <= 128 ? 4 : 8));
  const int rb = cpl <= 2 ? 4 : (cpl == 4 ? 2 : 1);
  int64_t blocks = (cp.total_rows + 8 * rb - 1) / (8 * rb);
  const int64_t cap = (int64_t) sm_count * 8;
  if (blocks > cap) blocks = cap;
  switch (cpl) {
    case 1: k_constrain<1><<<(unsigned) blocks, 256, 0, st>>>(cp); break;
    case 2: k_constrain<2><<<(unsigned) blocks, 256, 0, st>>>(cp); break;
    case 4: k_constrain<4><<<(unsigned) blocks, 256, 0, st>>>(cp); break;
    default: k_constrain<8><<<(unsigned) blocks, 256, 0, st>>>(cp); break;
  }
  return cudaGetLastError() == cudaSuccess ? KGE_OK : KGE_ECUDA;
}

__global__ void __launch_bounds__(1024) k_reduce(const ReduceParams p) {
  // Two levels, fixed shape: CTA c sums the contiguous slice [c * S, (c + 1) * S) of every quantity (thread t takes
  // i = t, t + 1024, ... ascending; all quantities in one sweep so the loads of a thread stay in flight together),
  // publishes ONE value per quantity, and the last CTA to finish (ticket) adds the gridDim.x values in CTA order.
  // The single-CTA version was bound by one SM's L2 bandwidth: 45 us for the 1.3 MB of partials at B = 65 536.
  __shared__ float sm[N_PART_MAX][32];
  __shared__ int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int64_t S = (p.B + gridDim.x - 1) / gridDim.x;
  const int64_t lo = (int64_t) blockIdx.x * S, hi = (lo + S < p.B) ? lo + S : p.B;
  float a[N_PART_MAX];
#pragma unroll
  for (int q = 0; q < N_PART_MAX; ++q) a[q] = 0.f;
#pragma unroll 4
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
#pragma unroll
    for (int q = 0; q < N_PART_MAX; ++q)
      if (q < p.n_part) a[q] += p.part[(size_t) q * p.B + i];
  }
#pragma unroll
  for (int q = 0; q < N_PART_MAX; ++q) {
    a[q] = warp_sum(a[q]);
    if (lane == 0) sm[q][warp] = a[q];
  }
  __syncthreads();
  if (warp < N_PART_MAX) {
    float b = (lane < nwarp) ? sm[warp][lane] : 0.f;
    b = warp_sum(b);
    if (lane == 0) p.stage[warp * RED_CTAS + blockIdx.x] = b;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(p.done, 1u);
    s_last = (t == gridDim.x - 1);
    if (s_last) *p.done = 0u;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (warp < p.n_part && warp < N_PART_MAX) {
    const int q = warp;
    float b = 0.f;
    for (int c = 0; c < (int) gridDim.x; ++c) b += __ldcg(p.stage + q * RED_CTAS + c);      // CTA order (warp-uniform)
    if (lane == 0) {
      if (q == 0) { p.scal[8] = b; *p.loss_out = b; publish_loss_to_host(p.dyn, b); }
      else {
        p.scal[4 + q - 1] = b;
        // tf.clip_by_norm: g * clip / max(||g||, clip)      (BaseModel.py:327)
        p.scal[q - 1] = p.clip_norm / fmaxf(sqrtf(b), p.clip_norm);
      }
    }
  }
}

// ---- row update: segmented record sum + clip + TF-semantics Adam --------------------
// Dense regularisers: TransH constraint=True (TransH.py:204-211) and
// RESCAL constraint=True (RESCAL.py:191-198) add a gradient to EVERY row, the variable's
// gradient becomes dense and tf.clip_by_norm sees the densified (summed) tensor.  Those
// tables take two passes: DENSE=1 computes per-row ||sum records + reg||^2 and the
// regulariser's loss term, DENSE=2 applies the update with the regulariser added.
enum { DK_NONE = 0, DK_TRANSH_ENT = 1, DK_TRANSH_REL = 2, DK_TRANSH_HYPER = 3, DK_RESCAL_ENT = 4 };
struct UpdateTable {
  float *w, *m, *v;
  const float* rec;      // gradient records of this table
  int64_t rows;
  int cols, ld;
  int rec_ld;            // row stride of rec
  // factored records (StepParams.fac_coef): record ids >= fac_first are negatives whose row is
  // c * fac_q[j * rec_ld] with {c, j} = the float / uint32 pair fac_coef[2 * (id - fac_first)]
  const float* fac_coef; // nullable
  const float* fac_q;
  uint32_t fac_first;
  int fac_k, fac_ht;
  uint32_t key_base;     // key of row 0 in the sorted key space
  uint32_t rec_base;     // subtract from a sorted value to get this table's record row
  int slot;              // table slot (index into scal)
  int dense_kind;        // DK_*
  const float* other_w;  // TransH: the paired relation table (rel_hyper for rel_emb and vice versa)
  int other_ld;
  const float* w_in;     // where the row's current weights are read: w, or the shadow wn when it is valid
  float* wn_out;         // shadow output: normalise(new row) for the next batch's gathers (or NULL)
  // data-parallel halves, [rows x ld] each:  gradient half (DENSE / MODE 3) stores the row sum to
  // all n_peer destinations (one local buffer, or every rank's staging slot over NVLink);
  // apply half adds the n_peer sources to the record sum in index order.
  float* peer[KGE_MAX_PEERS];
  int n_peer;
  // owner-partitioned exchange (kge_grads.owner_rows): gradient half sends row r to peer[r / owner_rows]
  // only; apply half works on rows [row_lo, row_lo + rows) and stores the new row to all n_out tables
  int owner_rows;
  int64_t row_lo;
  int n_out;
  float* w_outs[KGE_MAX_PEERS];
  float* wn_outs[KGE_MAX_PEERS];
  // NVSwitch multicast addresses of the tables (kge_grads.w_mc): ONE multimem.st reaches every replica
  float* w_mc;
  float* wn_mc;
  int w_private;         // kge_grads.w_private: the new w row stays with its owner, only wn travels
};
template <int CPL>
__device__ __forceinline__ void store_peers(const UpdateTable& T, int64_t row, int nch, int lane, const float4 (&g)[CPL]) {
  if (T.owner_rows > 0) {                          // reduce-scatter: the row's owner only
    float4* o = reinterpret_cast<float4*>(T.peer[row / T.owner_rows] + row * (int64_t) T.ld);
#pragma unroll
    for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; if (c < nch) o[c] = g[u]; }
    return;
  }
  for (int j = 0; j < T.n_peer; ++j) {
    float4* o = reinterpret_cast<float4*>(T.peer[j] + row * (int64_t) T.ld);
#pragma unroll
    for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; if (c < nch) o[c] = g[u]; }
  }
}
template <int CPL>
__device__ __forceinline__ void add_peers(const UpdateTable& T, int64_t row, int nch, int lane, float4 (&g)[CPL]) {
  for (int j = 0; j < T.n_peer; ++j) {
    const float4* gs = reinterpret_cast<const float4*>(T.peer[j] + row * (int64_t) T.ld);
#pragma unroll
    for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; if (c < nch) g[u] += __ldg(gs + c); }
  }
}
// where a table's gradient records live: materialised rows, or (factored) coefficient x shared row
struct RecSrc {
  const float* rec; int rec_ld;
  const float* fac_coef; const float* fac_q;
  uint32_t fac_first; int fac_k, fac_ht;
};
__device__ __forceinline__ RecSrc rec_src(const UpdateTable& T) {
  return RecSrc{T.rec, T.rec_ld, T.fac_coef, T.fac_q, T.fac_first, T.fac_k, T.fac_ht};
}
constexpr int SCAL_ABORT = 12;        // scal[12] != 0: a cross-GPU barrier gave up (k_dp_scales), the apply half must not run
struct UpdateParams {
  UpdateTable tab[KGE_MAX_TABLES];
  int n_tabs;
  int64_t total_rows;
  const uint32_t* keys;  // sorted
  const uint32_t* vals;  // record ids, stable within equal keys
  int64_t n_keys;
  const uint32_t* seg_lo;  // per key: first sorted position, one past the last (k_segments)
  const uint32_t* seg_hi;
  uint32_t seg_key0;       // k_update: seg_lo / seg_hi are indexed by key - seg_key0 (row-lazy mode keeps the relation range only)
  const float* scal;
  float lr_t, beta1, beta2, eps;
  const float* lr_dyn;   // nullable: lr_t lives on the device (captured step graph, DynScal)
  int opt;               // kge_opt_kind
  float cw;              // constraint_weight
  float* rowpart;        // DENSE=1: [2][total_rows]  (norm^2, loss term) per row
  const uint32_t* hdr;   // k_update_fast: per key [count, first position, first 6 record ids] (k_canon)
  const uint32_t* heavy_keys;  // keys with more than UPD_HEAVY records (k_canon): k_update_heavy's work list
  const uint32_t* n_heavy;
};

__device__ __forceinline__ int64_t lower_bound_u32(const uint32_t* a, int64_t n, uint32_t key) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// ---- segment table: first / one-past-last sorted position of every key -----------------
// (replaces two binary searches per row; both arrays are zeroed first, absent keys keep lo == hi)
__global__ void __launch_bounds__(256) k_segments(const uint32_t* __restrict__ keys, int64_t n, uint32_t* seg_lo,
                                                  uint32_t* seg_hi) {
  const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t k = keys[i];
  if (i == 0 || keys[i - 1] != k) seg_lo[k] = (uint32_t) i;
  if (i == n - 1 || keys[i + 1] != k) seg_hi[k] = (uint32_t) (i + 1);
}

// Sum of n <= 32 records whose ids sit one per lane in `mine`, added to the lane's chunks
// c = lane + 32 u strictly in lane (= sorted, stable) order.  The row loads of J records x CPL
// chunks are independent, so many are in flight; the ragged last batch is predicated, not serial.
// All 32 lanes must call.
// records in flight: light rows (a handful of records) / heavy slices (hundreds: the adds are a serial
// chain by definition, so the only lever is how many row loads are outstanding)
template <int CPL> struct RecDepth {
  static constexpr int LIGHT = CPL <= 2 ? 2 : (CPL == 4 ? 2 : 1);
  static constexpr int HEAVY = CPL == 1 ? 16 : (CPL == 2 ? 8 : (CPL == 4 ? 4 : 2));
};
__device__ __forceinline__ float4 scale4(float s, float4 a) {
  return f4(__fmul_rn(s, a.x), __fmul_rn(s, a.y), __fmul_rn(s, a.z), __fmul_rn(s, a.w));
}
template <int CPL, int J = RecDepth<CPL>::LIGHT>
__device__ __forceinline__ void sum_chunk(uint32_t mine, int n, const RecSrc& R, int nch, int lane, float4 (&g)[CPL]) {
  for (int j = 0; j < n; j += J) {
    const float4* r[J];
    float cf[J];
    bool fc[J];
#pragma unroll
    for (int x = 0; x < J; ++x) {
      const uint32_t rid = __shfl_sync(FULL, mine, (j + x) & 31);
      fc[x] = R.fac_coef != nullptr && rid >= R.fac_first && j + x < n;
      cf[x] = 0.f;
      if (fc[x]) {                    // {coefficient, shared-row index} written by the step kernel
        const float2 cq = __ldg(reinterpret_cast<const float2*>(R.fac_coef) + (rid - R.fac_first));
        cf[x] = cq.x;
        r[x] = reinterpret_cast<const float4*>(R.fac_q + (int64_t) __float_as_uint(cq.y) * R.rec_ld);
      } else r[x] = reinterpret_cast<const float4*>(R.rec + (int64_t) rid * R.rec_ld);
    }
    float4 a[CPL][J];
#pragma unroll
    for (int u = 0; u < CPL; ++u) {
      const int c = lane + 32 * u;
#pragma unroll
      for (int x = 0; x < J; ++x) a[u][x] = (c < nch && j + x < n) ? __ldg(r[x] + c) : f4zero();
    }
#pragma unroll
    for (int x = 0; x < J; ++x)
      if (j + x < n) {
        // fl(coef * q): the row the step kernel would have stored (explicit rounding, no FMA with the add)
#pragma unroll
        for (int u = 0; u < CPL; ++u) g[u] += fc[x] ? scale4(cf[x], a[u][x]) : a[u][x];
      }
  }
}

// Records at sorted positions [lo, hi), ids fetched 32 at a time.
template <int CPL, int J = RecDepth<CPL>::LIGHT>
__device__ __forceinline__ void sum_records(const uint32_t* __restrict__ vals, const RecSrc& R, uint32_t rec_base, int64_t lo,
                                            int64_t hi, int nch, int lane, float4 (&g)[CPL]) {
  for (int64_t q0 = lo; q0 < hi; q0 += WARP) {
    const int n = (int) ((hi - q0 < WARP) ? hi - q0 : WARP);
    const uint32_t mine = (lane < n) ? __ldg(vals + q0 + lane) - rec_base : 0u;
    sum_chunk<CPL, J>(mine, n, R, nch, lane, g);
  }
}
// plain materialised records (row-lazy / sharded kernels)
template <int CPL, int J = RecDepth<CPL>::LIGHT>
__device__ __forceinline__ void sum_records(const uint32_t* __restrict__ vals, const float* __restrict__ rec, int rec_ld,
                                            uint32_t rec_base, int64_t lo, int64_t hi, int nch, int lane, float4 (&g)[CPL]) {
  const RecSrc R{rec, rec_ld, nullptr, nullptr, 0u, 1, 0};
  sum_records<CPL, J>(vals, R, rec_base, lo, hi, nch, lane, g);
}

// Keras Adam, sparse path: m = b1*m (+ (1-b1) g on touched rows), likewise v; then
// w -= lr_t * m / (sqrt(v) + eps) for EVERY row.  One function for every
// update kernel so the arithmetic (and its FMA contraction) is the same everywhere.
__device__ __forceinline__ float adam1(float& w, float& m, float& v, float g, float b1, float b2, float lr_t, float eps) {
  // explicit rounding of every operation: no context-dependent FMA contraction, so every kernel
  // that updates a row produces the same bits
  m = __fadd_rn(__fmul_rn(b1, m), __fmul_rn(1.f - b1, g));
  v = __fadd_rn(__fmul_rn(b2, v), __fmul_rn(__fmul_rn(1.f - b2, g), g));
  w = __fsub_rn(w, __fdiv_rn(__fmul_rn(lr_t, m), __fadd_rn(sqrtf(v), eps)));
  return w;
}
// Keras SGD (momentum 0) and Adagrad on the summed row gradient; a zero gradient leaves the row as it
// is, so the every-row passes of the Adam path stay correct (only superfluous) for them
__device__ __forceinline__ void sgd1(float& w, float g, float lr) { w = __fsub_rn(w, __fmul_rn(lr, g)); }
__device__ __forceinline__ void adagrad1(float& w, float& acc, float g, float lr, float eps) {
  acc = __fadd_rn(acc, __fmul_rn(g, g));
  w = __fsub_rn(w, __fdiv_rn(__fmul_rn(lr, g), __fadd_rn(sqrtf(acc), eps)));
}
__device__ __forceinline__ void adam_chunk(float4& w, float4& m, float4& v, const float4 gg, float b1, float b2, float lr_t,
                                           float eps, int opt = KGE_OPTIM_ADAM) {
  if (opt == KGE_OPTIM_SGD) { sgd1(w.x, gg.x, lr_t); sgd1(w.y, gg.y, lr_t); sgd1(w.z, gg.z, lr_t); sgd1(w.w, gg.w, lr_t); return; }
  if (opt == KGE_OPTIM_ADAGRAD) {
    adagrad1(w.x, v.x, gg.x, lr_t, eps); adagrad1(w.y, v.y, gg.y, lr_t, eps);
    adagrad1(w.z, v.z, gg.z, lr_t, eps); adagrad1(w.w, v.w, gg.w, lr_t, eps);
    return;
  }
  adam1(w.x, m.x, v.x, gg.x, b1, b2, lr_t, eps);
  adam1(w.y, m.y, v.y, gg.y, b1, b2, lr_t, eps);
  adam1(w.z, m.z, v.z, gg.z, b1, b2, lr_t, eps);
  adam1(w.w, m.w, v.w, gg.w, b1, b2, lr_t, eps);
}
// keep the zero-padding invariant on the ragged last chunk
__device__ __forceinline__ void mask_tail(float4& w, float4& m, float4& v, int c, int cols) {
  const int e0 = 4 * c;
  if (e0 + 1 >= cols) { w.y = 0.f; m.y = 0.f; v.y = 0.f; }
  if (e0 + 2 >= cols) { w.z = 0.f; m.z = 0.f; v.z = 0.f; }
  if (e0 + 3 >= cols) { w.w = 0.f; m.w = 0.f; v.w = 0.f; }
}


// One warp per table row (8 rows per CTA); CPL = float4 chunks per lane (row width <= 128*CPL floats).
template <int DENSE, int CPL>
__global__ void __launch_bounds__(256) k_update(const UpdateParams p) {
  __shared__ float4 s_part[8][CPL * WARP];
  __shared__ int s_lo[8], s_hi[8], s_tab[8], s_heavy[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t gw = (int64_t) blockIdx.x * 8 + warp;
  const bool valid = gw < p.total_rows;
  int64_t row = valid ? gw : 0;
  int j = 0;
  while (j < p.n_tabs - 1 && row >= p.tab[j].rows) { row -= p.tab[j].rows; ++j; }
  const UpdateTable& T = p.tab[j];
  uint32_t lo = 0, hi = 0;
  if (valid && T.rec != nullptr && p.n_keys > 0) {
    const uint32_t key = T.key_base + (uint32_t) row;
    lo = __ldg(p.seg_lo + (key - p.seg_key0));
    hi = __ldg(p.seg_hi + (key - p.seg_key0));
  }
  const int nch = (T.cols + 3) >> 2;
  const bool heavy = (hi - lo) > (uint32_t) UPD_HEAVY;
  if (lane == 0) { s_lo[warp] = (int) lo; s_hi[warp] = (int) hi; s_tab[warp] = j; s_heavy[warp] = heavy; }
  __syncthreads();

  float4 g[CPL];
#pragma unroll
  for (int u = 0; u < CPL; ++u) g[u] = f4zero();
  if (!heavy) sum_records<CPL>(p.vals, rec_src(T), T.rec_base, lo, hi, nch, lane, g);
  for (int w2 = 0; w2 < 8; ++w2) {                     // CTA-uniform loop
    if (!s_heavy[w2]) continue;
    const UpdateTable& H = p.tab[s_tab[w2]];
    const int hn = (H.cols + 3) >> 2;
    const int64_t L = s_hi[w2] - s_lo[w2];
    float4 part[CPL];
#pragma unroll
    for (int u = 0; u < CPL; ++u) part[u] = f4zero();
    sum_records<CPL, RecDepth<CPL>::HEAVY>(p.vals, rec_src(H), H.rec_base, s_lo[w2] + L * warp / 8,
                                           s_lo[w2] + L * (warp + 1) / 8, hn, lane, part);
#pragma unroll
    for (int u = 0; u < CPL; ++u) s_part[warp][u * WARP + lane] = part[u];
    __syncthreads();
    if (warp == w2) {
      for (int w3 = 0; w3 < 8; ++w3)                   // fixed order: slices are contiguous in sorted order
#pragma unroll
        for (int u = 0; u < CPL; ++u) g[u] += s_part[w3][u * WARP + lane];
    }
    __syncthreads();
  }
  if (!valid) return;

  if (DENSE == 3) {            // gradient only: every row is written (zeros where nothing was looked up)
    store_peers<CPL>(T, row, nch, lane, g);
    return;
  }
  add_peers<CPL>(T, row, nch, lane, g);        // n_peer == 0 outside the data-parallel apply

  const float b1 = p.beta1, b2 = p.beta2;
  float4* w4 = reinterpret_cast<float4*>(T.w + row * (int64_t) T.ld);

  // dense-regulariser coefficients of this row:  g += ca * w + cb * other
  float ca = 0.f, cb = 0.f, lterm = 0.f;
  const float4* o4 = nullptr;
  if (DENSE != 0 && T.dense_kind != DK_NONE) {
    if (T.dense_kind == DK_TRANSH_ENT) {
      float ss = 0.f;
      for (int c = lane; c < nch; c += WARP) { float4 x = w4[c]; ss += dot4(x, x); }
      ss = warp_sum(ss);
      const float nn = sqrtf(ss);
      const float val = nn * nn - 1.0f;              // pow(norm, 2) - value   (constraint.py:62-67)
      if (val >= 0.f) { ca = 2.0f * p.cw; lterm = p.cw * val; }
    } else if (T.dense_kind == DK_RESCAL_ENT) {
      float ss = 0.f;
      for (int c = lane; c < nch; c += WARP) { float4 x = w4[c]; ss += dot4(x, x); }
      ss = warp_sum(ss);
      ca = 2.0f * p.cw / (float) T.rows;             // d/de of cw * mean_i ||e_i||^2
      lterm = p.cw * ss / (float) T.rows;
    } else {
      // TransH.py:205-209: q = (w . r) / ||r||, term = clip(q^2 - 1e-18, 0, inf)
      o4 = reinterpret_cast<const float4*>(T.other_w + row * (int64_t) T.other_ld);
      const bool is_r = (T.dense_kind == DK_TRANSH_REL);     // this row is r, other is w (or vice versa)
      float wr = 0.f, rr = 0.f;
      for (int c = lane; c < nch; c += WARP) {
        float4 x = w4[c], o = o4[c];
        wr += dot4(x, o);
        rr += is_r ? dot4(x, x) : dot4(o, o);
      }
      warp_sum2(wr, rr);
      const float nr = sqrtf(rr);
      const float qv = wr / nr;
      if (qv * qv - 1e-18f >= 0.f) {
        if (is_r) { ca = -p.cw * 2.0f * qv * wr / (nr * nr * nr); cb = p.cw * 2.0f * qv / nr; lterm = p.cw * (qv * qv - 1e-18f); }
        else { ca = 0.f; cb = p.cw * 2.0f * qv / nr; }
      }
    }
  }

  if (DENSE == 1) {
    float n2 = 0.f;
    if (T.dense_kind != DK_NONE) {
#pragma unroll
      for (int u = 0; u < CPL; ++u) {
        const int c = lane + 32 * u;
        if (c < nch) {
          float4 gg = g[u];
          gg += ca * w4[c];
          if (o4) gg += cb * o4[c];
          n2 += dot4(gg, gg);
        }
      }
      n2 = warp_sum(n2);
    }
    if (lane == 0) { p.rowpart[gw] = n2; p.rowpart[p.total_rows + gw] = lterm; }
    return;
  }

  const float scale = p.scal[T.slot];
  float4* m4 = reinterpret_cast<float4*>(T.m + row * (int64_t) T.ld);
  float4* v4 = reinterpret_cast<float4*>(T.v + row * (int64_t) T.ld);
#pragma unroll
  for (int u = 0; u < CPL; ++u) {
    const int c = lane + 32 * u;
    if (c >= nch) continue;
    float4 gg = g[u];
    float4 w = w4[c];
    if (DENSE == 2 && T.dense_kind != DK_NONE) {
      gg += ca * w;
      if (o4) gg += cb * o4[c];
    }
    gg = scale4(scale, gg);
    float4 m = m4[c], v = v4[c];
    adam_chunk(w, m, v, gg, b1, b2, p.lr_dyn ? __ldg(p.lr_dyn) : p.lr_t, p.eps, p.opt);
    mask_tail(w, m, v, c, T.cols);
    m4[c] = m; v4[c] = v; w4[c] = w;
  }
}

// one 16-byte store through a multicast (NVLS) address: the switch writes every replica
__device__ __forceinline__ void st_mc(float* addr, float4 v) {
  asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// Clip scale + Adam on one row held by a warp (g = summed, unscaled gradient); optionally also
// writes normalise(new row) to the shadow table the next batch gathers from.
template <int CPL>
__device__ __forceinline__ void adam_row(const UpdateParams& p, const UpdateTable& T, int64_t row, int nch, int lane,
                                         const float4 (&g)[CPL]) {
  const float4* wi = reinterpret_cast<const float4*>(T.w_in + row * (int64_t) T.ld);
  float4* w4 = reinterpret_cast<float4*>(T.w + row * (int64_t) T.ld);
  float4* m4 = reinterpret_cast<float4*>(T.m + row * (int64_t) T.ld);
  float4* v4 = reinterpret_cast<float4*>(T.v + row * (int64_t) T.ld);
  const float scale = p.scal[T.slot];
  const float lr_t = p.lr_dyn ? __ldg(p.lr_dyn) : p.lr_t;
  float4 wnew[CPL];
#pragma unroll
  for (int u = 0; u < CPL; ++u) {
    const int c = lane + 32 * u;
    wnew[u] = f4zero();
    if (c >= nch) continue;
    float4 w = wi[c], m = m4[c], v = v4[c];
    adam_chunk(w, m, v, scale4(scale, g[u]), p.beta1, p.beta2, lr_t, p.eps, p.opt);
    if (T.cols & 3) mask_tail(w, m, v, c, T.cols);
    m4[c] = m; v4[c] = v;
    if (T.n_out == 0 || (T.w_private && T.wn_out != nullptr)) w4[c] = w;
    else if (T.w_mc != nullptr) st_mc(T.w_mc + row * (int64_t) T.ld + 4 * c, w);                                   // all-gather by ONE multicast store
    else for (int j = 0; j < T.n_out; ++j) reinterpret_cast<float4*>(T.w_outs[j] + row * (int64_t) T.ld)[c] = w;   // all-gather by stores
    wnew[u] = w;
  }
  if (T.wn_out != nullptr) {                       // warp-uniform
    const float nrm = sqrtf(row_sumsq<CPL>(wnew));
    if (T.n_out == 0) {
      float4* o = reinterpret_cast<float4*>(T.wn_out + row * (int64_t) T.ld);
#pragma unroll
      for (int u = 0; u < CPL; ++u) { const int c = lane + 32 * u; if (c < nch) o[c] = div4(wnew[u], nrm); }
    } else {
#pragma unroll
      for (int u = 0; u < CPL; ++u) {
        const int c = lane + 32 * u;
        if (c >= nch) continue;
        const float4 q = div4(wnew[u], nrm);
        if (T.wn_mc != nullptr) st_mc(T.wn_mc + row * (int64_t) T.ld + 4 * c, q);
        else for (int j = 0; j < T.n_out; ++j) reinterpret_cast<float4*>(T.wn_outs[j] + row * (int64_t) T.ld)[c] = q;
      }
    }
  }
}

// The common training case (no dense regulariser): one warp per table row, warps fully
// independent (no CTA barrier), dependent-load chain shortened to header -> records -> w/m/v:
//   * the w / m / v lines are requested into L2 first (no registers held),
//   * one 32-byte header per key gives the record count, the first sorted position and the first
//     six record ids, so most rows never read the id array,
//   * rows with more than UPD_HEAVY records are left to k_update_heavy (whole-CTA summation),
//     which runs concurrently on the side stream.
// Summation order and update arithmetic are those of k_update: results are bit-identical.
// MODE 0: fused step.  MODE 3: data-parallel gradient half (row sum -> every T.peer destination, tables untouched).
// MODE 4: data-parallel apply half (gradient = sum of the T.peer rows, no records).
template <int CPL, int MODE>
__global__ void __launch_bounds__(64, CPL <= 2 ? 18 : 1) k_update_fast(const UpdateParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t gw = (int64_t) blockIdx.x * 2 + (threadIdx.x >> 5);
  if (gw >= p.total_rows) return;
  // last table first: the relation rows (few rows, many records each: the long-running warps) start at once
  // instead of forming the kernel's tail behind thousands of short entity rows
  int64_t row = p.total_rows - 1 - gw;
  int j = 0;
  while (j < p.n_tabs - 1 && row >= p.tab[j].rows) { row -= p.tab[j].rows; ++j; }
  const UpdateTable& T = p.tab[j];
  if (row >= T.rows) return;                       // owner-partitioned apply: the last table may own no rows
  pdl_wait();                                      // everything above is launch-parameter arithmetic
  if (MODE == 4 && p.scal[SCAL_ABORT] != 0.f) return;   // a peer never reached the barrier: leave the tables alone
  row += T.row_lo;
  const int nch = (T.cols + 3) >> 2;
  const float4* w4 = reinterpret_cast<const float4*>(T.w_in + row * (int64_t) T.ld);
  const float4* m4 = reinterpret_cast<const float4*>(T.m + row * (int64_t) T.ld);
  const float4* v4 = reinterpret_cast<const float4*>(T.v + row * (int64_t) T.ld);
  if (MODE != 3) {
#pragma unroll
    for (int u = 0; u < CPL; ++u) {
      const int c = lane + 32 * u;
      if (c < nch && (c & 7) == 0) {                 // one request per 128-byte line
        asm volatile("prefetch.global.L2 [%0];" ::"l"(w4 + c));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(m4 + c));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(v4 + c));
      }
    }
  }
  uint32_t hv = 0;
  if (MODE != 4 && T.rec != nullptr && lane < 8) hv = __ldg(p.hdr + ((int64_t) T.key_base + row) * 8 + lane);
  const uint32_t L = __shfl_sync(FULL, hv, 0), lo = __shfl_sync(FULL, hv, 1);
  if (L > (uint32_t) UPD_HEAVY) return;             // k_update_heavy's row
  float4 g[CPL];
#pragma unroll
  for (int u = 0; u < CPL; ++u) g[u] = f4zero();
  if (L > 0) {
    const int n0 = (int) (L < 32u ? L : 32u);
    uint32_t mine = __shfl_sync(FULL, hv, (lane + 2) & 31);           // lanes 0..5: ids from the header
    if (lane >= 6 && lane < n0) mine = __ldg(p.vals + lo + lane);
    mine = (lane < n0) ? mine - T.rec_base : 0u;
    const RecSrc RS = rec_src(T);
    sum_chunk<CPL>(mine, n0, RS, nch, lane, g);
    if (L > 32u) sum_records<CPL>(p.vals, RS, T.rec_base, (int64_t) lo + 32, (int64_t) lo + L, nch, lane, g);
  }
  if (MODE == 3) { store_peers<CPL>(T, row, nch, lane, g); return; }
  if (MODE == 4) add_peers<CPL>(T, row, nch, lane, g);
  adam_row<CPL>(p, T, row, nch, lane, g);
}

// k_update_fast<1, 0> for tables whose rows are <= 64 floats (C1: d = 50): at 13 of 32 lanes per row a warp is mostly
// idle, so a warp takes TWO rows, one per 16-lane half.  Same header / record / Adam sequence per row, the records
// added strictly in sorted order, the shadow row's norm by the same butterfly (the upper half of k_update_fast's
// warp holds zeros): bit-identical results (tests/test_gpu_step.py::test_two_rows_per_warp_update_equals_k_update_fast).
// Plain fused step only (MODE 0, materialised records, no data-parallel outputs).
__global__ void __launch_bounds__(64, 18) k_update_narrow(const UpdateParams p) {
  const int lane = threadIdx.x & 31, sub = lane >> 4, hl = lane & 15;
  const int64_t idx = (((int64_t) blockIdx.x * 2 + (threadIdx.x >> 5)) << 1) + sub;
  const bool in_range = idx < p.total_rows;
  int64_t row = in_range ? p.total_rows - 1 - idx : 0;       // last table first, as k_update_fast
  int j = 0;
  while (j < p.n_tabs - 1 && row >= p.tab[j].rows) { row -= p.tab[j].rows; ++j; }
  const UpdateTable& T = p.tab[j];
  const bool valid = in_range && row < T.rows;
  pdl_wait();
  const int nch = (T.cols + 3) >> 2;
  const bool mine_l = hl < nch;
  const float4* wi = reinterpret_cast<const float4*>(T.w_in + row * (int64_t) T.ld);
  float4* w4 = reinterpret_cast<float4*>(T.w + row * (int64_t) T.ld);
  float4* m4 = reinterpret_cast<float4*>(T.m + row * (int64_t) T.ld);
  float4* v4 = reinterpret_cast<float4*>(T.v + row * (int64_t) T.ld);
  if (valid && mine_l && (hl & 7) == 0) {                     // one request per 128-byte line
    asm volatile("prefetch.global.L2 [%0];" ::"l"(wi + hl));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(m4 + hl));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(v4 + hl));
  }
  uint32_t hv = 0;
  if (valid && T.rec != nullptr && hl < 8) hv = __ldg(p.hdr + ((int64_t) T.key_base + row) * 8 + hl);
  const uint32_t L = __shfl_sync(FULL, hv, 0, 16), lo = __shfl_sync(FULL, hv, 1, 16);
  const bool active = valid && L <= (uint32_t) UPD_HEAVY;     // heavier rows are k_update_heavy's
  const uint32_t Lm = active ? L : 0u;
  const uint32_t hid = __shfl_sync(FULL, hv, (hl + 2) & 15, 16);      // lanes 0..5: the header's record ids
  float4 g = f4zero();
  for (uint32_t base = 0; __any_sync(FULL, base < Lm); base += 16) {
    const int n = base < Lm ? (int) (Lm - base < 16u ? Lm - base : 16u) : 0;
    uint32_t mine = 0;
    if (hl < n) {
      const uint32_t pos = base + hl;
      mine = (pos < 6u ? hid : __ldg(p.vals + lo + pos)) - T.rec_base;
    }
    const int n_other = __shfl_xor_sync(FULL, n, 16);
    const int nmax = n > n_other ? n : n_other;
    for (int x = 0; x < nmax; ++x) {
      const uint32_t rid = __shfl_sync(FULL, mine, x, 16);
      if (x < n && mine_l) g += __ldg(reinterpret_cast<const float4*>(T.rec + (int64_t) rid * T.rec_ld) + hl);
    }
  }
  float4 wnew = f4zero();
  if (active && mine_l) {
    const float scale = p.scal[T.slot];
    const float lr_t = p.lr_dyn ? __ldg(p.lr_dyn) : p.lr_t;
    float4 w = wi[hl], m = m4[hl], v = v4[hl];
    adam_chunk(w, m, v, scale4(scale, g), p.beta1, p.beta2, lr_t, p.eps, p.opt);
    if (T.cols & 3) mask_tail(w, m, v, hl, T.cols);
    m4[hl] = m; v4[hl] = v; w4[hl] = w;
    wnew = w;
  }
  // shadow row: normalise(new row) for the next batch's gathers (row_sumsq + div4, over the half)
  float ss = 0.f;
  ss = __fmaf_rn(wnew.x, wnew.x, ss); ss = __fmaf_rn(wnew.y, wnew.y, ss);
  ss = __fmaf_rn(wnew.z, wnew.z, ss); ss = __fmaf_rn(wnew.w, wnew.w, ss);
  const float nrm = sqrtf(warp_sum_w<16>(ss));
  if (active && mine_l && T.wn_out != nullptr) reinterpret_cast<float4*>(T.wn_out + row * (int64_t) T.ld)[hl] = div4(wnew, nrm);
}
inline bool update_narrow_ok(const UpdateParams& up, int cpl) {
  if (cpl != 1 || getenv("KGE_UPDATE_NARROW_OFF")) return false;
  for (int j = 0; j < up.n_tabs; ++j)
    if (up.tab[j].cols > 64 || up.tab[j].fac_coef != nullptr || up.tab[j].n_out != 0 || up.tab[j].n_peer != 0 || up.tab[j].row_lo != 0) return false;
  return true;
}

// Rows with many records (hot entities / relations): one CTA per (heavy key, table), the eight
// warps sum contiguous slices of the sorted record list, partials are combined in slice order —
// the order k_update uses for its heavy rows, so the bits agree.
template <int CPL, int MODE>
__global__ void __launch_bounds__(256) k_update_heavy(const UpdateParams p) {
  __shared__ float4 s_part[8][CPL * WARP];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const UpdateTable& T = p.tab[blockIdx.y];
  const uint32_t n_queued = *p.n_heavy;
  for (uint32_t item = blockIdx.x; item < n_queued; item += gridDim.x) {     // few CTAs, few items
  __syncthreads();
  const uint32_t key = p.heavy_keys[item];
  if (T.rec == nullptr || key < T.key_base || (int64_t) (key - T.key_base) >= T.rows) continue;
  const int64_t row = key - T.key_base;
  const int nch = (T.cols + 3) >> 2;
  const int64_t lo = __ldg(p.seg_lo + key), L = (int64_t) __ldg(p.seg_hi + key) - lo;
  float4 part[CPL];
#pragma unroll
  for (int u = 0; u < CPL; ++u) part[u] = f4zero();
  sum_records<CPL, RecDepth<CPL>::HEAVY>(p.vals, rec_src(T), T.rec_base, lo + L * warp / 8, lo + L * (warp + 1) / 8, nch,
                                         lane, part);
#pragma unroll
  for (int u = 0; u < CPL; ++u) s_part[warp][u * WARP + lane] = part[u];
  __syncthreads();
  if (warp != 0) continue;
  float4 g[CPL];
#pragma unroll
  for (int u = 0; u < CPL; ++u) {
    g[u] = f4zero();
    for (int w3 = 0; w3 < 8; ++w3) g[u] += s_part[w3][u * WARP + lane];
  }
  if (MODE == 3) store_peers<CPL>(T, row, nch, lane, g);
  else adam_row<CPL>(p, T, row, nch, lane, g);
  }
}

template <int MODE>
inline int launch_update_heavy(const UpdateParams& up, int cpl, int64_t max_heavy, cudaStream_t st) {
  const dim3 grid((unsigned) (max_heavy < 148 ? max_heavy : 148), (unsigned) up.n_tabs, 1);
  switch (cpl) {
    case 1: k_update_heavy<1, MODE><<<grid, 256, 0, st>>>(up); break;
    case 2: k_update_heavy<2, MODE><<<grid, 256, 0, st>>>(up); break;
    case 4: k_update_heavy<4, MODE><<<grid, 256, 0, st>>>(up); break;
    default: k_update_heavy<8, MODE><<<grid, 256, 0, st>>>(up); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("k_update_heavy launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
  return KGE_OK;
}

// KGE_PDL=1 launches the row update with programmatic stream serialization behind the step kernel (its launch
// latency then overlaps the step kernel's last CTAs).  Measured on C2 (round 2, scripts/gpu_pdl.sh): 80.9 vs 83.3 us
// per step with L2 flushed, but 871 vs 883 M triples/s in the steady state and no change end to end; with an
// explicit griddepcontrol.launch_dependents in k_step (at its start / after the negatives) 90.3 / 87.2 us — the
// parked update CTAs take the slots the next batch's grouping chain runs in.  Off by default.
inline bool pdl_enabled() {
  static const int on = [] { const char* e = getenv("KGE_PDL"); return e ? atoi(e) : 0; }();
  return on != 0;
}
template <class K, class P>
inline cudaError_t launch_pdl(K kern, unsigned grid, unsigned block, cudaStream_t st, const P& prm) {
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3(grid); lc.blockDim = dim3(block); lc.dynamicSmemBytes = 0; lc.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  lc.attrs = at; lc.numAttrs = 1;
  return cudaLaunchKernelEx(&lc, kern, prm);
}

template <int MODE>
inline int launch_update_fast(const UpdateParams& up, int cpl, cudaStream_t st) {
  if (MODE == 0 && update_narrow_ok(up, cpl)) {
    cudaError_t e;
    if (pdl_enabled()) e = launch_pdl(k_update_narrow, (unsigned) ((up.total_rows + 3) / 4), 64, st, up);
    else { k_update_narrow<<<(unsigned) ((up.total_rows + 3) / 4), 64, 0, st>>>(up); e = cudaGetLastError(); }
    if (e != cudaSuccess) { set_error("k_update_narrow launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
    return KGE_OK;
  }
  const unsigned blocks = (unsigned) ((up.total_rows + 1) / 2);
  if (MODE == 0 && pdl_enabled()) {
    cudaError_t e;
    switch (cpl) {
      case 1: e = launch_pdl(k_update_fast<1, MODE>, blocks, 64, st, up); break;
      case 2: e = launch_pdl(k_update_fast<2, MODE>, blocks, 64, st, up); break;
      case 4: e = launch_pdl(k_update_fast<4, MODE>, blocks, 64, st, up); break;
      default: e = launch_pdl(k_update_fast<8, MODE>, blocks, 64, st, up); break;
    }
    if (e != cudaSuccess) { set_error("k_update_fast launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
    return KGE_OK;
  }
  switch (cpl) {
    case 1: k_update_fast<1, MODE><<<blocks, 64, 0, st>>>(up); break;
    case 2: k_update_fast<2, MODE><<<blocks, 64, 0, st>>>(up); break;
    case 4: k_update_fast<4, MODE><<<blocks, 64, 0, st>>>(up); break;
    default: k_update_fast<8, MODE><<<blocks, 64, 0, st>>>(up); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("k_update_fast launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
  return KGE_OK;
}

template <int DENSE>
inline int launch_update(const UpdateParams& up, int cpl, int64_t blocks, cudaStream_t st) {
  switch (cpl) {
    case 1: k_update<DENSE, 1><<<(unsigned) blocks, 256, 0, st>>>(up); break;
    case 2: k_update<DENSE, 2><<<(unsigned) blocks, 256, 0, st>>>(up); break;
    case 4: k_update<DENSE, 4><<<(unsigned) blocks, 256, 0, st>>>(up); break;
    default: k_update<DENSE, 8><<<(unsigned) blocks, 256, 0, st>>>(up); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("k_update launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
  return KGE_OK;
}

// data-parallel apply: clip scales and loss from the [norm^2 x4, loss] blocks — one block already
// summed over ranks, or every rank's block (summed here in rank order)
struct ScalSlots { const float* s[KGE_MAX_PEERS]; int n; };
// abort (nullable): the error word of kge_peer_barrier.  KGE_PEER_DEAD => scal[SCAL_ABORT] = 1: the update
// kernels of the apply half return before touching a row, and the loss is NaN.
__global__ void k_dp_scales(const ScalSlots in, float clip_norm, float* scal, float* loss_out, const uint32_t* abort,
                            const DynScal* dyn = nullptr) {
  const int t = threadIdx.x;
  if (t > KGE_MAX_TABLES) return;
  const bool dead = abort != nullptr && *reinterpret_cast<const volatile uint32_t*>(abort) == KGE_PEER_DEAD;
  float a = 0.f;
  for (int j = 0; j < in.n; ++j) a += in.s[j][t];
  if (t < KGE_MAX_TABLES) {
    scal[4 + t] = a;
    scal[t] = clip_norm / fmaxf(sqrtf(a), clip_norm);
    if (t == 0) scal[SCAL_ABORT] = dead ? 1.f : 0.f;
  } else {
    scal[8] = a;
    const float loss = dead ? __int_as_float(0x7fc00000) : a;
    *loss_out = loss;
    publish_loss_to_host(dyn, loss);         // captured data-parallel step: the host may wait for exactly this
  }
}
// data-parallel gradient half with peer staging: this rank's [norm^2 x4, loss] block to every rank
__global__ void k_push_scal(const float* __restrict__ src, ScalSlots out) {
  const int t = threadIdx.x;
  if (t < 5) for (int j = 0; j < out.n; ++j) const_cast<float*>(out.s[j])[t] = src[t];
}

// ---- reduction of the dense pass: per-table norm^2 -> clip scale, regulariser -> loss ----
struct ReduceDenseParams {
  const float* rowpart;      // [2][total_rows] from k_update<1>
  int64_t total_rows;
  int64_t row_begin[KGE_MAX_TABLES + 1];   // table j owns rows [row_begin[j], row_begin[j+1])
  int slot[KGE_MAX_TABLES];
  int is_dense[KGE_MAX_TABLES];
  int n_tabs;
  const float* matpart;      // nullable: [2][n_mat] from k_update_mat<1>
  int64_t n_mat;
  int mat_slot;
  float mat_loss_coef;       // cw / n_rel
  float clip_norm;
  int write_scales;
  float* scal;
  float* loss_out;
  const DynScal* dyn;        // see ReduceParams.dyn: with a dense regulariser THIS kernel makes the loss final
};
__device__ __forceinline__ float block_sum_1024(float a, float* sm) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  a = warp_sum(a);
  __syncthreads();
  if (lane == 0) sm[warp] = a;
  __syncthreads();
  float b = (threadIdx.x < 32) ? sm[threadIdx.x] : 0.f;
  if (warp == 0) b = warp_sum(b);
  return b;                     // valid in warp 0
}
__global__ void __launch_bounds__(1024) k_reduce_dense(const ReduceDenseParams p) {
  __shared__ float sm[32];
  float loss_add = 0.f;
  for (int j = 0; j < p.n_tabs; ++j) {
    if (!p.is_dense[j]) continue;
    float a = 0.f, b = 0.f;
    for (int64_t i = p.row_begin[j] + threadIdx.x; i < p.row_begin[j + 1]; i += blockDim.x) {
      a += p.rowpart[i];
      b += p.rowpart[p.total_rows + i];
    }
    a = block_sum_1024(a, sm);
    b = block_sum_1024(b, sm);
    if (threadIdx.x == 0) {
      loss_add += b;
      if (p.write_scales) {
        p.scal[4 + p.slot[j]] = a;
        p.scal[p.slot[j]] = p.clip_norm / fmaxf(sqrtf(a), p.clip_norm);
      }
    }
  }
  if (p.matpart) {
    float a = 0.f, b = 0.f;
    for (int64_t i = threadIdx.x; i < p.n_mat; i += blockDim.x) { a += p.matpart[i]; b += p.matpart[p.n_mat + i]; }
    a = block_sum_1024(a, sm);
    b = block_sum_1024(b, sm);
    if (threadIdx.x == 0) {
      loss_add += p.mat_loss_coef * b;
      if (p.write_scales) {
        p.scal[4 + p.mat_slot] = a;
        p.scal[p.mat_slot] = p.clip_norm / fmaxf(sqrtf(a), p.clip_norm);
      }
    }
  }
  if (threadIdx.x == 0) { p.scal[8] += loss_add; *p.loss_out = p.scal[8]; publish_loss_to_host(p.dyn, p.scal[8]); }
}

}  // namespace kge
