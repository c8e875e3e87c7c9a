// rank_kernels.cuh — batch scoring and the generic (any vector-family model) filtered
// ranking kernel.  Replaces score_hrt on id vectors (BaseModel.py:410-430 + models) and
// get_rank (BaseModel.py:620-654): for each query, count candidates on one side whose
// score is strictly greater than the true triple's, skipping known positives.
//
// The true triple's score is computed by the same code path as the candidates', so the
// true entity never out-scores itself by rounding.
#pragma once
#include "models.cuh"

namespace kge {

struct ScoreParams {
  ModelDev md;
  const void *h, *r, *t;
  int id64;
  int64_t n;
  float* out;
};

template <class M, int NV>
__global__ void __launch_bounds__(256) k_score(const ScoreParams p) {
  using Ent = typename M::template Ent<NV>;
  using Rel = typename M::template Rel<NV>;
  const int lane = threadIdx.x & 31;
  int64_t w = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (; w < p.n; w += nw) {
    Ent H, T;
    Rel R;
    H.load(p.md, load_id(p.h, w, p.id64), lane);
    T.load(p.md, load_id(p.t, w, p.id64), lane);
    R.load(p.md, load_id(p.r, w, p.id64), lane);
    Row<NV> x, y;
    FwdSaved<NV> sv;
    const float s = M::template forward<NV>(p.md, H, R, T, lane, x, y, sv);
    if (lane == 0) p.out[w] = s;
  }
}

constexpr int RANK_THREADS = 256;
constexpr int RANK_WARPS = RANK_THREADS / WARP;

struct RankParams {
  ModelDev md;
  const void* queries;       // (Q, 3)
  int id64;
  int64_t Q;
  int side;                  // KGE_SIDE_H: candidates replace h; KGE_SIDE_T: replace t
  const int64_t* f_off;      // nullable filter: ids[f_off[q] .. f_end[q])   (kge_filter; packed CSR: f_end = f_off + 1)
  const int64_t* f_end;
  const int64_t* f_ids;
  int64_t ent_lo, ent_hi;    // candidate shard
  int chunk;                 // candidates per CTA
  int32_t* count;            // [Q], accumulated with integer atomics (order independent)
};

template <class M, int NV>
__global__ void __launch_bounds__(RANK_THREADS) k_rank_generic(const RankParams p) {
  using Ent = typename M::template Ent<NV>;
  using Rel = typename M::template Rel<NV>;
  __shared__ int s_cnt[RANK_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t q = blockIdx.y;
  const int64_t h = load_id(p.queries, 3 * q + 0, p.id64);
  const int64_t r = load_id(p.queries, 3 * q + 1, p.id64);
  const int64_t t = load_id(p.queries, 3 * q + 2, p.id64);
  Ent H, T;
  Rel R;
  H.load(p.md, h, lane);
  T.load(p.md, t, lane);
  R.load(p.md, r, lane);
  Row<NV> x, y;
  FwdSaved<NV> sv;
  const float pos = M::template forward<NV>(p.md, H, R, T, lane, x, y, sv);

  const int64_t f0 = p.f_off ? p.f_off[q] : 0, f1 = p.f_off ? p.f_end[q] : 0;
  const int64_t c0 = p.ent_lo + (int64_t) blockIdx.x * p.chunk;
  const int64_t c1 = (c0 + p.chunk < p.ent_hi) ? c0 + p.chunk : p.ent_hi;
  const bool head = (p.side == KGE_SIDE_H);
  int cnt = 0;
  for (int64_t e = c0 + warp; e < c1; e += RANK_WARPS) {
    Ent N;
    N.load(p.md, e, lane);
    const float s = M::template forward<NV>(p.md, head ? N : H, R, head ? T : N, lane, x, y, sv);
    if (s > pos) {                       // strict '>' (BaseModel.py:654); NaN compares false
      bool filtered = false;
      for (int64_t f = f0; f < f1; f += WARP) {
        const bool hit = (f + lane < f1) && (p.f_ids[f + lane] == e);
        if (__any_sync(FULL, hit)) { filtered = true; break; }
      }
      cnt += filtered ? 0 : 1;
    }
  }
  if (lane == 0) s_cnt[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
#pragma unroll
    for (int w2 = 0; w2 < RANK_WARPS; ++w2) tot += s_cnt[w2];
    if (tot) atomicAdd(p.count + q, tot);
  }
}

struct AuxLaunch {
  int nv;
  dim3 grid;
  cudaStream_t stream;
};

template <class M>
int launch_score_model(const ScoreParams& p, const AuxLaunch& L) {
  switch (L.nv) {
    case 1: k_score<M, 1><<<L.grid, 256, 0, L.stream>>>(p); break;
    case 2: k_score<M, 2><<<L.grid, 256, 0, L.stream>>>(p); break;
    case 4: k_score<M, 4><<<L.grid, 256, 0, L.stream>>>(p); break;
    case 8: k_score<M, 8><<<L.grid, 256, 0, L.stream>>>(p); break;
    default: set_error("unsupported row width (nv=%d)", L.nv); return KGE_EUNSUPPORTED;
  }
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}
template <class M>
int launch_rank_model(const RankParams& p, const AuxLaunch& L) {
  switch (L.nv) {
    case 1: k_rank_generic<M, 1><<<L.grid, RANK_THREADS, 0, L.stream>>>(p); break;
    case 2: k_rank_generic<M, 2><<<L.grid, RANK_THREADS, 0, L.stream>>>(p); break;
    case 4: k_rank_generic<M, 4><<<L.grid, RANK_THREADS, 0, L.stream>>>(p); break;
    case 8: k_rank_generic<M, 8><<<L.grid, RANK_THREADS, 0, L.stream>>>(p); break;
    default: set_error("unsupported row width (nv=%d)", L.nv); return KGE_EUNSUPPORTED;
  }
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}

#define KGE_DECL_AUX(name)                                      \
  int launch_score_##name(const ScoreParams&, const AuxLaunch&); \
  int launch_rank_##name(const RankParams&, const AuxLaunch&);
KGE_DECL_AUX(um)
KGE_DECL_AUX(transe)
KGE_DECL_AUX(transh)
KGE_DECL_AUX(transd)
KGE_DECL_AUX(rotate)
KGE_DECL_AUX(distmult)
#undef KGE_DECL_AUX

}  // namespace kge
