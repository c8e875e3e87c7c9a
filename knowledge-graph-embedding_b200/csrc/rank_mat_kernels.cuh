// rank_mat_kernels.cuh — batch scoring and filtered ranking for the matrix-relation
// models (SE, TransR, RESCAL).  k_rank_mat: one CTA per (query, candidate chunk) stages
// the query relation's matrix in shared memory, projects the query side once per warp and
// then one candidate per warp pass.  RESCAL needs no per-candidate mat-vec:
// score(e) = e . (R t)  or  (h^T R) . e.
#pragma once
#include "rank_kernels.cuh"
#include "mat_models.cuh"

namespace kge {

struct MatAux {
  int dpad;      // scratch floats per warp
  int stage;     // matrices staged in shared memory (rank only)
};

template <class M, int NV>
__device__ __forceinline__ float mat_score_triple(const ModelDev& md, const MatCtx& cx, const Row<NV>& Eh,
                                                  const Row<NV>& Et, const Row<NV>& Rv, int lane) {
  if constexpr (M::BILINEAR) {
    Row<NV> U;
    matvec_colin(U, Et, cx.M0, cx.ld0, md.d_ent, md.d_ent, cx.scratch, lane);
    return warp_sum(row_dot_partial(Eh, U));
  } else {
    Row<NV> CH, CT, x;
    M::template proj_fwd<NV>(md, cx, 0, Eh, CH, lane);
    M::template proj_fwd<NV>(md, cx, 1, Et, CT, lane);
#pragma unroll
    for (int u = 0; u < NV; ++u) x.v[u] = CH.v[u] + Rv.v[u];
    ScoreSaved sv;
    return score_fwd<NV, false>(md.sc, x, CT, sv);
  }
}

template <class M>
__device__ __forceinline__ void mat_ctx_global(const ModelDev& md, int64_t r, MatCtx& cx) {
  const int s0 = (M::KIND == KGE_TRANSR) ? KGE_T_REL_B : KGE_T_REL_A;
  const int mcols = (M::KIND == KGE_TRANSR) ? md.d_rel : md.d_ent;
  cx.M0 = md.t[s0].w + r * (int64_t) md.t[s0].ld;
  cx.M1 = M::TWO_MATS ? md.t[KGE_T_REL_B].w + r * (int64_t) md.t[KGE_T_REL_B].ld : nullptr;
  cx.ld0 = cx.ld1 = mcols;
}

template <class M, int NV>
__global__ void __launch_bounds__(256) k_score_mat(const ScoreParams p, const MatAux a) {
  extern __shared__ float4 smem4[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const ModelDev& md = p.md;
  int64_t w = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  const int din = M::d_in(md), dout = M::d_out(md);
  for (; w < p.n; w += nw) {
    const int64_t h = load_id(p.h, w, p.id64), r = load_id(p.r, w, p.id64), t = load_id(p.t, w, p.id64);
    MatCtx cx;
    mat_ctx_global<M>(md, r, cx);
    cx.scratch = reinterpret_cast<float*>(smem4) + warp * a.dpad;
    Row<NV> Eh, Et, Rv;
    sload(Eh, md.t[KGE_T_ENT].w + h * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
    sload(Et, md.t[KGE_T_ENT].w + t * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
    if (M::HAS_REL_VEC) sload(Rv, md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld, dout, lane);
    else Rv.zero();
    const float s = mat_score_triple<M, NV>(md, cx, Eh, Et, Rv, lane);
    if (lane == 0) p.out[w] = s;
  }
}

template <class M, int NV>
__global__ void __launch_bounds__(RANK_THREADS) k_rank_mat(const RankParams p, const MatAux a) {
  extern __shared__ float4 smem4[];
  __shared__ int s_cnt[RANK_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const ModelDev& md = p.md;
  const int64_t q = blockIdx.y;
  const int64_t h = load_id(p.queries, 3 * q + 0, p.id64);
  const int64_t r = load_id(p.queries, 3 * q + 1, p.id64);
  const int64_t t = load_id(p.queries, 3 * q + 2, p.id64);
  const int din = M::d_in(md), dout = M::d_out(md);
  float* sm = reinterpret_cast<float*>(smem4);
  MatCtx cx;
  mat_ctx_global<M>(md, r, cx);
  cx.scratch = sm + warp * a.dpad;
  if (a.stage) {
    float* s_mat = sm + RANK_WARPS * a.dpad;
    const int mrows = md.d_ent, mcols = cx.ld0, lds = mcols | 1;
    for (int e = threadIdx.x; e < mrows * mcols; e += RANK_THREADS) {
      const int ra = e / mcols, cb = e - ra * mcols;
      s_mat[ra * lds + cb] = __ldg(cx.M0 + e);
      if (M::TWO_MATS) s_mat[mrows * lds + ra * lds + cb] = __ldg(cx.M1 + e);
    }
    cx.M0 = s_mat; cx.M1 = s_mat + mrows * lds; cx.ld0 = cx.ld1 = lds;
  }
  __syncthreads();
  const bool head = (p.side == KGE_SIDE_H);
  Row<NV> Eh, Et, Rv, Cq, x;
  sload(Eh, md.t[KGE_T_ENT].w + h * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
  sload(Et, md.t[KGE_T_ENT].w + t * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
  if (M::HAS_REL_VEC) sload(Rv, md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld, dout, lane);
  else Rv.zero();
  // query-side vector: the projection of the entity that stays
  if constexpr (M::BILINEAR) {
    if (head) matvec_colin(Cq, Et, cx.M0, cx.ld0, din, din, cx.scratch, lane);     // U = R t
    else matvec_rowin(Cq, Eh, cx.M0, cx.ld0, din, din, cx.scratch, lane);          // V = h^T R
  } else {
    M::template proj_fwd<NV>(md, cx, head ? 1 : 0, head ? Et : Eh, Cq, lane);
  }
  ScoreSaved sv;
  auto cand_score = [&](const Row<NV>& N) -> float {
    if constexpr (M::BILINEAR) {
      return warp_sum(row_dot_partial(N, Cq));
    } else {
      Row<NV> CN;
      M::template proj_fwd<NV>(md, cx, head ? 0 : 1, N, CN, lane);
#pragma unroll
      for (int u = 0; u < NV; ++u) x.v[u] = (head ? CN.v[u] : Cq.v[u]) + Rv.v[u];
      return score_fwd<NV, false>(md.sc, x, head ? Cq : CN, sv);
    }
  };
  const float pos = cand_score(head ? Eh : Et);       // same code path as the candidates

  const int64_t f0 = p.f_off ? p.f_off[q] : 0, f1 = p.f_off ? p.f_end[q] : 0;
  const int64_t c0 = p.ent_lo + (int64_t) blockIdx.x * p.chunk;
  const int64_t c1 = (c0 + p.chunk < p.ent_hi) ? c0 + p.chunk : p.ent_hi;
  int cnt = 0;
  for (int64_t e = c0 + warp; e < c1; e += RANK_WARPS) {
    Row<NV> N;
    sload(N, md.t[KGE_T_ENT].w + e * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
    const float s = cand_score(N);
    if (s > pos) {
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

template <class M>
int launch_score_mat_model(const ScoreParams& p, const AuxLaunch& L, const MatAux& a, size_t smem) {
  switch (L.nv) {
    case 1: k_score_mat<M, 1><<<L.grid, 256, smem, L.stream>>>(p, a); break;
    case 2: k_score_mat<M, 2><<<L.grid, 256, smem, L.stream>>>(p, a); break;
    default: set_error("matrix models support embedding sizes up to 256 (nv=%d)", L.nv); return KGE_EUNSUPPORTED;
  }
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}
template <class M>
int launch_rank_mat_model(const RankParams& p, const AuxLaunch& L, const MatAux& a, size_t smem) {
  if (L.nv == 1) {
    if (smem > 48 * 1024) KGE_CUDA_OK(cudaFuncSetAttribute(k_rank_mat<M, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    k_rank_mat<M, 1><<<L.grid, RANK_THREADS, smem, L.stream>>>(p, a);
  } else if (L.nv == 2) {
    if (smem > 48 * 1024) KGE_CUDA_OK(cudaFuncSetAttribute(k_rank_mat<M, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    k_rank_mat<M, 2><<<L.grid, RANK_THREADS, smem, L.stream>>>(p, a);
  } else { set_error("matrix models support embedding sizes up to 256 (nv=%d)", L.nv); return KGE_EUNSUPPORTED; }
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}

#define KGE_DECL_MAT(name)                                                                  \
  int launch_score_##name(const ScoreParams&, const AuxLaunch&, const MatAux&, size_t);     \
  int launch_rank_##name(const RankParams&, const AuxLaunch&, const MatAux&, size_t);
KGE_DECL_MAT(se)
KGE_DECL_MAT(transr)
KGE_DECL_MAT(rescal)
#undef KGE_DECL_MAT

}  // namespace kge
