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
  uint32_t key_base;     // key of row 0 in the sorted key space
  uint32_t rec_base;     // subtract from a sorted value to get this table's record row
  int slot;              // table slot (index into scal)
  int dense_kind;        // DK_*
  const float* other_w;  // TransH: the paired relation table (rel_hyper for rel_emb and vice versa)
  int other_ld;
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
  float cw;              // constraint_weight
  float* rowpart;        // DENSE=1: [2][total_rows]  (norm^2, loss term) per row
};

__device__ __forceinline__ int64_t lower_bound_u32(const uint32_t* a, int64_t n, uint32_t key) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// sum of the row's records for float4 chunk c, in sorted (stable) order; record ids are
// fetched 32 at a time and the row loads are independent, so several are in flight
// (all 32 lanes must call; lanes whose chunk c is past the row clamp it and discard)
__device__ __forceinline__ float4 sum_records(const UpdateParams& p, const UpdateTable& T, int64_t lo, int64_t hi,
                                              int c, int lane) {
  float4 g = f4zero();
  for (int64_t q0 = lo; q0 < hi; q0 += WARP) {
    const int n = (int) ((hi - q0 < WARP) ? hi - q0 : WARP);
    const uint32_t mine = (lane < n) ? __ldg(p.vals + q0 + lane) - T.rec_base : 0u;
    int j = 0;
    for (; j + 4 <= n; j += 4) {
      const uint32_t r0 = __shfl_sync(FULL, mine, j), r1 = __shfl_sync(FULL, mine, j + 1),
                     r2 = __shfl_sync(FULL, mine, j + 2), r3 = __shfl_sync(FULL, mine, j + 3);
      const float4 a0 = __ldg(reinterpret_cast<const float4*>(T.rec + (int64_t) r0 * T.rec_ld) + c);
      const float4 a1 = __ldg(reinterpret_cast<const float4*>(T.rec + (int64_t) r1 * T.rec_ld) + c);
      const float4 a2 = __ldg(reinterpret_cast<const float4*>(T.rec + (int64_t) r2 * T.rec_ld) + c);
      const float4 a3 = __ldg(reinterpret_cast<const float4*>(T.rec + (int64_t) r3 * T.rec_ld) + c);
      g += a0; g += a1; g += a2; g += a3;
    }
    for (; j < n; ++j) {
      const uint32_t r0 = __shfl_sync(FULL, mine, j);
      g += __ldg(reinterpret_cast<const float4*>(T.rec + (int64_t) r0 * T.rec_ld) + c);
    }
  }
  return g;
}

template <int DENSE>
__global__ void __launch_bounds__(256) k_update(const UpdateParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (gw >= p.total_rows) return;
  int64_t row = gw;
  int j = 0;
  while (j < p.n_tabs - 1 && row >= p.tab[j].rows) { row -= p.tab[j].rows; ++j; }
  const UpdateTable& T = p.tab[j];
  const uint32_t key = T.key_base + (uint32_t) row;
  int64_t lo = 0, hi = 0;
  if (T.rec != nullptr && p.n_keys > 0) {
    lo = lower_bound_u32(p.keys, p.n_keys, key);
    hi = lower_bound_u32(p.keys, p.n_keys, key + 1);
  }
  const float b1 = p.beta1, b2 = p.beta2;
  const int nch = (T.cols + 3) >> 2;
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
      for (int c0 = 0; c0 < nch; c0 += WARP) {
        const bool act = c0 + lane < nch;
        const int c = act ? c0 + lane : nch - 1;
        float4 g = sum_records(p, T, lo, hi, c, lane);
        g += ca * w4[c];
        if (o4) g += cb * o4[c];
        if (act) n2 += dot4(g, g);
      }
      n2 = warp_sum(n2);
    }
    if (lane == 0) { p.rowpart[gw] = n2; p.rowpart[p.total_rows + gw] = lterm; }
    return;
  }

  const float scale = p.scal[T.slot];
  float4* m4 = reinterpret_cast<float4*>(T.m + row * (int64_t) T.ld);
  float4* v4 = reinterpret_cast<float4*>(T.v + row * (int64_t) T.ld);
  for (int c0 = 0; c0 < nch; c0 += WARP) {
    const bool act = c0 + lane < nch;
    const int c = act ? c0 + lane : nch - 1;
    float4 g = sum_records(p, T, lo, hi, c, lane);
    if (!act) continue;
    float4 w = w4[c];
    if (DENSE == 2 && T.dense_kind != DK_NONE) {
      g += ca * w;
      if (o4) g += cb * o4[c];
    }
    g = scale * g;
    float4 m = m4[c], v = v4[c];
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
  if (threadIdx.x == 0) { p.scal[8] += loss_add; *p.loss_out = p.scal[8]; }
}

}  // namespace kge
