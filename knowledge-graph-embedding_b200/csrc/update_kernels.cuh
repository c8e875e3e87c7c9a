// update_kernels.cuh — constraint, partial reduction and row update kernels.
// All are model independent: they see tables, records and sorted keys only.
#pragma once
#include "common.cuh"

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

__global__ void __launch_bounds__(256) k_constrain(const ConstrainParams p) {
  const int lane = threadIdx.x & 31;
  int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (; gw < p.total_rows; gw += nw) {
    int64_t row = gw;
    int j = 0;
    while (j < p.n_jobs - 1 && row >= p.job[j].rows) { row -= p.job[j].rows; ++j; }
    const ConstrainJob& J = p.job[j];
    float4* base = reinterpret_cast<float4*>(J.w + row * (int64_t) J.ld);
    const int nch = (J.cols + 3) >> 2;
    float ss = 0.f;
    for (int c = lane; c < nch; c += WARP) { float4 v = base[c]; ss += dot4(v, v); }
    ss = warp_sum(ss);
    const float nrm = sqrtf(ss);                 // pow(sum(pow(|x|, 2)), 1/2)
    if (J.op == CON_CLIP && nrm < 1.0f) continue;
    const float den = (J.op == CON_CLIP) ? fmaxf(nrm, 1e-9f) : nrm;
    for (int c = lane; c < nch; c += WARP) {
      float4 v = base[c];
      base[c] = f4(v.x / den, v.y / den, v.z / den, v.w / den);   // X / norm * 1
    }
  }
}

// ---- fixed-order reduction of per-positive partials ---------------------------------
// scal[0..3] = clip scale per table, scal[4..7] = norm^2 per table, scal[8] = loss.
struct ReduceParams {
  const float* part;     // [n_part][B]
  int64_t B;
  int n_part;
  float clip_norm;
  float* scal;
  float* loss_out;
};

__global__ void __launch_bounds__(1024) k_reduce(const ReduceParams p) {
  __shared__ float sm[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int q = 0; q < p.n_part; ++q) {
    const float* src = p.part + (size_t) q * p.B;
    float a = 0.f;
    for (int64_t i = threadIdx.x; i < p.B; i += blockDim.x) a += src[i];
    a = warp_sum(a);
    if (lane == 0) sm[warp] = a;
    __syncthreads();
    if (warp == 0) {
      float b = (lane < nwarp) ? sm[lane] : 0.f;
      b = warp_sum(b);
      if (lane == 0) {
        if (q == 0) { p.scal[8] = b; *p.loss_out = b; }
        else {
          p.scal[4 + q - 1] = b;
          // tf.clip_by_norm: g * clip / max(||g||, clip)      (BaseModel.py:327)
          p.scal[q - 1] = p.clip_norm / fmaxf(sqrtf(b), p.clip_norm);
        }
      }
    }
    __syncthreads();
  }
}

// ---- row update: segmented record sum + clip + TF-semantics Adam --------------------
struct UpdateTable {
  float *w, *m, *v;
  const float* rec;      // gradient records of this table
  int64_t rows;
  int cols, ld;
  int rec_ld;            // row stride of rec
  uint32_t key_base;     // key of row 0 in the sorted key space
  uint32_t rec_base;     // subtract from a sorted value to get this table's record row
  int slot;              // table slot (index into scal)
};
struct UpdateParams {
  UpdateTable tab[KGE_MAX_TABLES];
  int n_tabs;
  int64_t total_rows;
  const uint32_t* keys;  // sorted
  const uint32_t* vals;  // record ids, stable within equal keys
  int64_t n_keys;
  const float* scal;
  float lr_t, beta1, beta2, eps;
};

__device__ __forceinline__ int64_t lower_bound_u32(const uint32_t* a, int64_t n, uint32_t key) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(256) k_update(const UpdateParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (gw >= p.total_rows) return;
  int64_t row = gw;
  int j = 0;
  while (j < p.n_tabs - 1 && row >= p.tab[j].rows) { row -= p.tab[j].rows; ++j; }
  const UpdateTable& T = p.tab[j];
  const uint32_t key = T.key_base + (uint32_t) row;
  const int64_t lo = lower_bound_u32(p.keys, p.n_keys, key);
  const int64_t hi = lower_bound_u32(p.keys, p.n_keys, key + 1);
  const float scale = p.scal[T.slot];
  const float b1 = p.beta1, b2 = p.beta2;
  const int nch = (T.cols + 3) >> 2;
  float4* w4 = reinterpret_cast<float4*>(T.w + row * (int64_t) T.ld);
  float4* m4 = reinterpret_cast<float4*>(T.m + row * (int64_t) T.ld);
  float4* v4 = reinterpret_cast<float4*>(T.v + row * (int64_t) T.ld);
  for (int c = lane; c < nch; c += WARP) {
    float4 g = f4zero();
    for (int64_t q = lo; q < hi; ++q) {
      const int64_t rr = (int64_t) (__ldg(p.vals + q) - T.rec_base);
      g += __ldg(reinterpret_cast<const float4*>(T.rec + rr * (int64_t) T.rec_ld) + c);
    }
    g = scale * g;
    float4 m = m4[c], v = v4[c], w = w4[c];
    // Keras Adam, sparse path: m = b1*m (+ (1-b1) g on touched rows), likewise v; then
    // w -= lr_t * m / (sqrt(v) + eps) for EVERY row.
    m = f4(b1 * m.x + (1.f - b1) * g.x, b1 * m.y + (1.f - b1) * g.y,
           b1 * m.z + (1.f - b1) * g.z, b1 * m.w + (1.f - b1) * g.w);
    v = f4(b2 * v.x + (1.f - b2) * g.x * g.x, b2 * v.y + (1.f - b2) * g.y * g.y,
           b2 * v.z + (1.f - b2) * g.z * g.z, b2 * v.w + (1.f - b2) * g.w * g.w);
    w = f4(w.x - p.lr_t * m.x / (sqrtf(v.x) + p.eps), w.y - p.lr_t * m.y / (sqrtf(v.y) + p.eps),
           w.z - p.lr_t * m.z / (sqrtf(v.z) + p.eps), w.w - p.lr_t * m.w / (sqrtf(v.w) + p.eps));
    // keep the zero-padding invariant on the ragged last chunk
    const int e0 = 4 * c;
    if (e0 + 1 >= T.cols) { w.y = 0.f; m.y = 0.f; v.y = 0.f; }
    if (e0 + 2 >= T.cols) { w.z = 0.f; m.z = 0.f; v.z = 0.f; }
    if (e0 + 3 >= T.cols) { w.w = 0.f; m.w = 0.f; v.w = 0.f; }
    m4[c] = m; v4[c] = v; w4[c] = w;
  }
}

}  // namespace kge
