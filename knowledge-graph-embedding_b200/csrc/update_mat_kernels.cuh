// update_mat_kernels.cuh — per-relation matrix gradient + TF-semantics Adam for the
// matrix tables of SE / TransR / RESCAL.
//
// grid = (tiles, n_rel, n_tables); a CTA owns one 64x64 tile of one relation's matrix:
//   G_tile = sum over the relation's positives (sorted, stable => occurrence order) and
//            their pair slots of  rowvec (x) colvec           (grouped small GEMM, K = #pairs)
//   then   g = clip_scale * (G + dense_reg);  m,v,w <- Adam     (EVERY relation: Keras Adam
//   decays/moves all rows).
// MODE 1 (norm pass, only for dense-regulariser models) writes sum((G + reg)^2) and the
// regulariser's loss term per CTA instead of updating.  No atomics: deterministic.
#pragma once
#include "update_kernels.cuh"
#include <stdlib.h>

namespace kge {

constexpr int MU_TILE = 64, MU_KC = 16, MU_THREADS = 256;

struct MatUpdTable {
  float *w, *m, *v;
  int rows_m, cols_m;      // matrix dims (row-major inside the table row)
  int ld;                  // table row stride (floats)
  int slot;                // kge table slot (index into scal)
  int use_sh, use_st;      // shared-h / shared-t pair slots feed this table
  int neg_mode;            // per-negative pair slots: 0 none, 1 all, 2 head-corrupted, 3 tail-corrupted
  int ent_row_sh, ent_row_st, ent_row_neg;   // 1: entity row indexes matrix rows (P indexes columns)
  float* gout;             // MODE 2 (data-parallel gradient half): the relation's summed gradient goes here ([n_rel x ld]), tables untouched
};

// optimizer step of one matrix element (the arithmetic of k_update_mat's epilogue; k_mat_apply shares it)
__device__ __forceinline__ void mat_opt1(float& w, float& m, float& v, float g, int opt, float lr_t, float beta1, float beta2, float eps) {
  if (opt == KGE_OPTIM_SGD) w = w - lr_t * g;
  else if (opt == KGE_OPTIM_ADAGRAD) {
    const float acc = v + g * g;
    v = acc;
    w = w - lr_t * g / (sqrtf(acc) + eps);
  } else {
    m = beta1 * m + (1.f - beta1) * g;
    v = beta2 * v + (1.f - beta2) * g * g;
    w = w - lr_t * m / (sqrtf(v) + eps);
  }
}

struct MatUpdParams {
  MatUpdTable tab[2];
  int n_tabs;
  int64_t n_rel;
  const float* ent_w; int ent_ld, d_ent;
  const float* recP; int ldp, dp;
  const void* triples; int id64;
  const int64_t* neg_ids;
  int64_t B; int k, side_mode;
  const uint32_t* keys; const uint32_t* vals; int64_t n_keys;
  const uint32_t* seg_lo; const uint32_t* seg_hi;      // k_segments
  uint32_t key_base, rec_base;      // relation keys are key_base + r; vals - rec_base = positive index
  const float* scal;
  float lr_t, beta1, beta2, eps;
  const float* lr_dyn;   // nullable: lr_t on the device (captured step graph)
  int opt;               // kge_opt_kind
  float reg_coef;                   // dense regulariser: g += reg_coef * w   (RESCAL.py:191-198), 0 = none
  float* part;                      // MODE 1: [2][n_tabs * n_rel * tiles] (norm^2, sum w^2)
};

// number of per-negative slots of table T for one positive, and the c of the j-th
__device__ __forceinline__ int mu_neg_count(int neg_mode, int side_mode, int k) {
  if (neg_mode == 0) return 0;
  if (neg_mode == 1) return k;
  const bool want_head = (neg_mode == 2);
  if (side_mode == KGE_SIDE_HT) return k / 2;
  return ((side_mode == KGE_SIDE_H) == want_head) ? k : 0;
}
__device__ __forceinline__ int mu_neg_c(int neg_mode, int side_mode, int j) {
  if (neg_mode == 1 || side_mode != KGE_SIDE_HT) return j;
  return 2 * j + (neg_mode == 2 ? 0 : 1);       // 'h+t': even c = head corruption
}

// TC = columns per tile.  128-column tiles (a 100 x 100 matrix as 2 CTAs of 4 x 8 outputs per thread: one wave of
// 474 CTAs instead of 1.6 waves of 948) were measured SLOWER on C4 (update phase TransR 66 -> 84 us, SE 81 -> 108 us):
// the kernel is bound by the per-CTA chain of pair chunks, and halving the CTAs halves the chains in flight.
// The wide instantiation stays compiled for the A/B (KGE_MU_WIDE=1).
inline int mu_tile_cols(int cols_m) {
  static const bool wide = getenv("KGE_MU_WIDE") != nullptr;
  return (wide && cols_m > MU_TILE) ? 2 * MU_TILE : MU_TILE;
}
template <int MODE, int TC>
__global__ void __launch_bounds__(MU_THREADS, TC == MU_TILE ? 4 : 2) k_update_mat(const MatUpdParams p) {
  constexpr int NB = TC / 16;          // output columns per thread: 4 + 4 (second set 64 columns on) for TC = 128
  __shared__ float s_row[MU_KC][MU_TILE];
  __shared__ float s_col[MU_KC][TC];
  __shared__ float s_red[2][MU_THREADS / 32];
  __shared__ int64_t s_rec[2][MU_KC], s_ent[2][MU_KC];
  __shared__ int s_entrow[2][MU_KC];
  const MatUpdTable& T = p.tab[blockIdx.z];
  const int64_t r = blockIdx.y;
  const int tiles_c = (T.cols_m + TC - 1) / TC;
  const int tr = blockIdx.x / tiles_c, tc = blockIdx.x % tiles_c;
  if (tr * MU_TILE >= T.rows_m) return;
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  const int row0 = tr * MU_TILE, col0 = tc * TC;

  const uint32_t key = p.key_base + (uint32_t) r;
  const int64_t lo = p.n_keys > 0 ? (int64_t) __ldg(p.seg_lo + key) : 0;
  const int64_t hi = p.n_keys > 0 ? (int64_t) __ldg(p.seg_hi + key) : 0;
  const int n_neg = mu_neg_count(T.neg_mode, p.side_mode, p.k);
  const int S = T.use_sh + T.use_st + n_neg;
  const int64_t n_pairs = (hi - lo) * S;

  float acc[4][NB];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < NB; ++b) acc[a][b] = 0.f;

  // Software pipeline over chunks of MU_KC pairs.  The dependent lookups of a pair (sorted position -> positive ->
  // entity / record id) are done by MU_KC threads TWO chunks ahead, the 2 * MU_KC * 64 vector loads of the next
  // chunk are issued into registers before this chunk's FMAs, so a chunk costs one store-to-shared, its FMAs and
  // two barriers instead of two dependent global round trips (the update of C4 TransR was latency-bound on them).
  auto lookup = [&](int64_t u0, int buf) {
    if (threadIdx.x < MU_KC) {
      const int64_t u = u0 + threadIdx.x;
      int64_t rec = -1, ent = 0;
      int ent_row = 0;
      if (u < n_pairs) {
        const int64_t pos = (int64_t) (__ldg(p.vals + lo + u / S) - p.rec_base);
        const int s = (int) (u % S);
        if (T.use_sh && s == 0) { rec = pos; ent = load_id(p.triples, 3 * pos, p.id64); ent_row = T.ent_row_sh; }
        else if (T.use_st && s == T.use_sh) { rec = p.B + pos; ent = load_id(p.triples, 3 * pos + 2, p.id64); ent_row = T.ent_row_st; }
        else {
          const int c = mu_neg_c(T.neg_mode, p.side_mode, s - T.use_sh - T.use_st);
          rec = 2 * p.B + pos * p.k + c; ent = __ldg(p.neg_ids + pos * p.k + c); ent_row = T.ent_row_neg;
        }
      }
      s_rec[buf][threadIdx.x] = rec; s_ent[buf][threadIdx.x] = ent; s_entrow[buf][threadIdx.x] = ent_row;
    }
  };
  constexpr int NROWV = MU_KC * MU_TILE;                          // row-vector elements of a chunk; column vectors follow
  constexpr int NLD = MU_KC * (MU_TILE + TC) / MU_THREADS;        // vector elements per thread and chunk
  float stage[NLD];
  auto fetch = [&](int buf) {
#pragma unroll
    for (int x = 0; x < NLD; ++x) {
      const int e = x * MU_THREADS + threadIdx.x;
      const int which = e >= NROWV;                       // 0 row vectors, 1 column vectors
      const int e2 = which ? e - NROWV : e;
      const int kk = which ? e2 / TC : e2 / MU_TILE, off = which ? e2 % TC : e2 % MU_TILE;
      const int64_t rec = s_rec[buf][kk];
      float val = 0.f;
      if (rec >= 0) {
        const bool from_ent = (which == 0) == (s_entrow[buf][kk] != 0);
        const int idx = (which == 0 ? row0 : col0) + off;
        const int lim = which == 0 ? T.rows_m : T.cols_m;
        if (idx < lim)
          val = from_ent ? __ldg(p.ent_w + s_ent[buf][kk] * (int64_t) p.ent_ld + idx) : __ldg(p.recP + rec * (int64_t) p.ldp + idx);
      }
      stage[x] = val;
    }
  };
  if (n_pairs > 0) {
    lookup(0, 0);
    __syncthreads();
    fetch(0);
    lookup(MU_KC, 1);
  }
  int cb = 0;
  for (int64_t u0 = 0; u0 < n_pairs; u0 += MU_KC, cb ^= 1) {
#pragma unroll
    for (int x = 0; x < NLD; ++x) {
      const int e = x * MU_THREADS + threadIdx.x;
      const int which = e >= NROWV;
      const int e2 = which ? e - NROWV : e;
      if (which) s_col[e2 / TC][e2 % TC] = stage[x];
      else s_row[e2 / MU_TILE][e2 % MU_TILE] = stage[x];
    }
    __syncthreads();                                        // the chunk is staged; lookups of chunk + 1 are visible
    if (u0 + MU_KC < n_pairs) {
      fetch(cb ^ 1);                                        // next chunk's vectors, in flight under the FMAs
      lookup(u0 + 2 * MU_KC, cb);                           // this chunk's lookup slots are free: its loads have landed
    }
#pragma unroll
    for (int kk = 0; kk < MU_KC; ++kk) {
      float a4[4], b4[NB];
#pragma unroll
      for (int a = 0; a < 4; ++a) a4[a] = s_row[kk][ty * 4 + a];
#pragma unroll
      for (int b = 0; b < NB; ++b) b4[b] = s_col[kk][(b >> 2) * 64 + tx * 4 + (b & 3)];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < NB; ++b) acc[a][b] = fmaf(a4[a], b4[b], acc[a][b]);
    }
    __syncthreads();
  }

  float* wr = T.w + r * (int64_t) T.ld;
  float* mr = T.m + r * (int64_t) T.ld;
  float* vr = T.v + r * (int64_t) T.ld;
  const float scale = (MODE == 0) ? p.scal[T.slot] : 1.f;     // MODE 2 never reads it
  const float lr_t = p.lr_dyn ? __ldg(p.lr_dyn) : p.lr_t;
  float n2 = 0.f, w2 = 0.f;
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int ra = row0 + ty * 4 + a;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const int cb = col0 + (b >> 2) * 64 + tx * 4 + (b & 3);
      if (ra < T.rows_m && cb < T.cols_m) {
        const int64_t o = (int64_t) ra * T.cols_m + cb;
        if (MODE == 2) { T.gout[r * (int64_t) T.ld + o] = acc[a][b]; continue; }      // gradient half: raw sum of this rank's pairs
        float w = wr[o];
        float g = acc[a][b] + p.reg_coef * w;
        if (MODE == 1) { n2 += g * g; w2 += w * w; }
        else {
          g *= scale;
          float m = p.opt == KGE_OPTIM_ADAM ? mr[o] : 0.f, v = p.opt != KGE_OPTIM_SGD ? vr[o] : 0.f;
          mat_opt1(w, m, v, g, p.opt, lr_t, p.beta1, p.beta2, p.eps);
          if (p.opt == KGE_OPTIM_ADAM) mr[o] = m;
          if (p.opt != KGE_OPTIM_SGD) vr[o] = v;
          wr[o] = w;
        }
      }
    }
  }
  if (MODE == 1) {
    warp_sum2(n2, w2);
    if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = n2; s_red[1][threadIdx.x >> 5] = w2; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float a = 0.f, b = 0.f;
      for (int w2i = 0; w2i < MU_THREADS / 32; ++w2i) { a += s_red[0][w2i]; b += s_red[1][w2i]; }
      const int64_t n_cta = (int64_t) gridDim.x * gridDim.y * gridDim.z;
      const int64_t me = ((int64_t) blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
      p.part[me] = a;
      p.part[n_cta + me] = b;
    }
  }
}

// Data-parallel apply half for a matrix table: g = clip scale * (rank-summed gradient), optimizer step on EVERY
// element (Keras Adam moves all rows, App. B.2).  One thread per element of the [n_rel x rows_m * cols_m] table.
struct MatApplyParams {
  float *w, *m, *v;
  const float* gsum;       // [n_rel x ld]
  int64_t n_rel; int elems, ld, slot;
  const float* scal;
  float lr_t, beta1, beta2, eps;
  const float* lr_dyn;
  int opt;
};
__global__ void __launch_bounds__(256) k_mat_apply(const MatApplyParams p) {
  const int64_t total = p.n_rel * (int64_t) p.elems;
  const float scale = p.scal[p.slot];
  const float lr_t = p.lr_dyn ? __ldg(p.lr_dyn) : p.lr_t;
  for (int64_t x = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; x < total; x += (int64_t) gridDim.x * blockDim.x) {
    const int64_t r = x / p.elems;
    const int64_t o = r * p.ld + (x - r * p.elems);
    float w = p.w[o], m = p.opt == KGE_OPTIM_ADAM ? p.m[o] : 0.f, v = p.opt != KGE_OPTIM_SGD ? p.v[o] : 0.f;
    const float g = p.gsum[o] * scale;
    mat_opt1(w, m, v, g, p.opt, lr_t, p.beta1, p.beta2, p.eps);
    if (p.opt == KGE_OPTIM_ADAM) p.m[o] = m;
    if (p.opt != KGE_OPTIM_SGD) p.v[o] = v;
    p.w[o] = w;
  }
}

}  // namespace kge
