// rank_lp_tile.cuh — filtered ranking for translation models scored with an Lp distance whose
// candidate side is a plain entity row (TransE, UM with LpDistance / LpDistancePow):
//
//      score(e) = -g( S(e) ),   S(e) = sum_j |q_j - e_j|^p   (max_j for p = inf),   g monotone
//      count_q  = #{ e : S(e) < S(true) }                 (strict, BaseModel.py:651-653)
//
// with the query vector q = h + r (tail candidates) or t - r (head candidates).  |q - e| has no
// GEMM form, so this is a register-tiled CUDA-core kernel shaped like one: a CTA owns a
// 128 x 128 (query x entity) tile, stages 32-dimension slabs of both operands transposed in
// shared memory, and each thread accumulates an 8 x 8 micro-tile: 4 LDS.128 feed 128 ALU ops.
// The entity tile is read once per 128 queries (bytes = N * row per query tile).
//
// Exactness.  Every S is accumulated by ONE thread in ascending j with the same two operations
// (lp_acc), in this kernel, in k_rank_lp_prep (S of the true triple) and in k_rank_lp_filter
// (known positives): the three agree bit for bit, the true entity never beats itself, and the
// integer counts equal an all-sequential evaluation.  The clip of score.py:61 (sum clipped to
// 1e-9 before the root) is applied to both sides of the comparison.
#pragma once
#include "rank_kernels.cuh"

namespace kge {

enum { LPK_L1 = 0, LPK_L2 = 1, LPK_LINF = 2, LPK_GEN = 3 };

template <int KIND>
__device__ __forceinline__ float lp_acc(float acc, float q, float e, float p) {
  const float d = __fsub_rn(q, e);
  if (KIND == LPK_L1) return __fadd_rn(acc, fabsf(d));
  if (KIND == LPK_L2) return __fmaf_rn(d, d, acc);
  if (KIND == LPK_LINF) return fmaxf(acc, fabsf(d));
  return __fadd_rn(acc, powf(fabsf(d), p));
}
// complex rows (RotatE, re/im interleaved): |z| is the modulus (tf.abs on complex64, score.py:61);
// one MUFU.SQRT per complex element — the same function in all three kernels
__device__ __forceinline__ float sqrt_fast(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
template <int KIND>
__device__ __forceinline__ float lp_acc_c(float acc, float qre, float qim, float ere, float eim, float p) {
  const float dr = __fsub_rn(qre, ere), di = __fsub_rn(qim, eim);
  const float m2 = __fmaf_rn(di, di, __fmul_rn(dr, dr));
  if (KIND == LPK_L2) return __fadd_rn(acc, m2);
  const float m = sqrt_fast(m2);
  if (KIND == LPK_L1) return __fadd_rn(acc, m);
  if (KIND == LPK_LINF) return fmaxf(acc, m);
  return __fadd_rn(acc, powf(m, p));
}
__device__ __forceinline__ float lp_clip(float s, int kind) { return kind == LPK_LINF ? s : fmaxf(s, 1e-9f); }

struct LpParams {
  ModelDev md;
  const void* queries; int id64; int64_t Q; int side;
  float* Qv; int ldq;              // [Qpad x ldq] query vectors, zero padded
  float* thr;                      // [Q] clipped S of the true triple
  const float* E; int ldE; int d;  // candidate table; d = floats per row (2 x complex dimension for RotatE)
  int64_t ent_lo, ent_hi;
  const int64_t* f_off; const int64_t* f_end; const int64_t* f_ids;      // kge_filter (f_end = f_off + 1 for a packed CSR)
  int32_t* count;
  float p;
};

// one thread per query: q vector and the threshold
template <int KIND, bool CPLX>
__global__ void __launch_bounds__(128) k_rank_lp_prep(const LpParams p) {
  const int64_t q = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= p.Q) return;
  const ModelDev& md = p.md;
  const int64_t h = load_id(p.queries, 3 * q + 0, p.id64), r = load_id(p.queries, 3 * q + 1, p.id64),
                t = load_id(p.queries, 3 * q + 2, p.id64);
  const bool head = (p.side == KGE_SIDE_H);
  const float* keep = p.E + (head ? t : h) * (int64_t) p.ldE;
  const float* truth = p.E + (head ? h : t) * (int64_t) p.ldE;
  const float* rel = (md.model == KGE_UM) ? nullptr : md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld;
  float* qv = p.Qv + q * (int64_t) p.ldq;
  float s = 0.f;
  if constexpr (CPLX) {
    // tail candidates: q = h o r.  Head candidates: |e o r - t| = |e - t o conj(r)| (|r| = 1)
    for (int c = 0; c < p.d / 2; ++c) {
      float sn, cs;
      sincosf(rel[c] * md.rotate_scale, &sn, &cs);              // r_emb / limit * pi, RotatE.py:160
      if (head) sn = -sn;
      const float kr = keep[2 * c], ki = keep[2 * c + 1];
      const float qre = kr * cs - ki * sn, qim = kr * sn + ki * cs;
      qv[2 * c] = qre; qv[2 * c + 1] = qim;
      s = lp_acc_c<KIND>(s, qre, qim, truth[2 * c], truth[2 * c + 1], p.p);
    }
  } else {
    for (int j = 0; j < p.d; ++j) {
      float v = keep[j];
      if (rel) v = head ? __fsub_rn(v, rel[j]) : __fadd_rn(v, rel[j]);     // t - r | h + r
      qv[j] = v;
      s = lp_acc<KIND>(s, v, truth[j], p.p);
    }
  }
  p.thr[q] = lp_clip(s, KIND);
}

constexpr int LP_T = 128, LP_DK = 32, LP_THREADS = 256;

template <int KIND, bool CPLX>
__global__ void __launch_bounds__(LP_THREADS) k_rank_lp_tile(const LpParams p) {
  __shared__ __align__(16) float sQ[LP_DK][LP_T];
  __shared__ __align__(16) float sE[LP_DK][LP_T];
  __shared__ int s_cnt[LP_T];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;        // tx: entity group, ty: query group
  const int64_t q0 = (int64_t) blockIdx.y * LP_T, e0 = p.ent_lo + (int64_t) blockIdx.x * LP_T;
  float acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
  if (tid < LP_T) s_cnt[tid] = 0;

  for (int j0 = 0; j0 < p.d; j0 += LP_DK) {
    __syncthreads();
    // stage transposed: thread reads one float4 of a row (4 dims), writes 4 scalars
    for (int it = tid; it < LP_T * (LP_DK / 4); it += LP_THREADS) {
      const int row = it % LP_T, c4 = it / LP_T;          // consecutive threads -> consecutive smem words
      const int j = j0 + 4 * c4;
      float4 qv = f4zero(), ev = f4zero();
      if (j < p.d) {                                     // rows are zero padded to a multiple of 4
        if (q0 + row < p.Q) qv = *reinterpret_cast<const float4*>(p.Qv + (q0 + row) * (int64_t) p.ldq + j);
        if (e0 + row < p.ent_hi) ev = __ldg(reinterpret_cast<const float4*>(p.E + (e0 + row) * (int64_t) p.ldE + j));
      }
      sQ[4 * c4 + 0][row] = qv.x; sQ[4 * c4 + 1][row] = qv.y; sQ[4 * c4 + 2][row] = qv.z; sQ[4 * c4 + 3][row] = qv.w;
      sE[4 * c4 + 0][row] = ev.x; sE[4 * c4 + 1][row] = ev.y; sE[4 * c4 + 2][row] = ev.z; sE[4 * c4 + 3][row] = ev.w;
    }
    __syncthreads();
    const int jn = (p.d - j0 < LP_DK) ? p.d - j0 : LP_DK;         // padding dims must not enter (|0-0| is fine, pow is not)
    if constexpr (CPLX) {
      for (int j = 0; j < jn; j += 2) {                  // floats 2c, 2c+1 = re, im of complex element c
        const float4 ra = *reinterpret_cast<const float4*>(&sQ[j][ty * 4]), rb = *reinterpret_cast<const float4*>(&sQ[j][64 + ty * 4]);
        const float4 ia = *reinterpret_cast<const float4*>(&sQ[j + 1][ty * 4]), ib = *reinterpret_cast<const float4*>(&sQ[j + 1][64 + ty * 4]);
        const float4 ua = *reinterpret_cast<const float4*>(&sE[j][tx * 4]), ub = *reinterpret_cast<const float4*>(&sE[j][64 + tx * 4]);
        const float4 va = *reinterpret_cast<const float4*>(&sE[j + 1][tx * 4]), vb = *reinterpret_cast<const float4*>(&sE[j + 1][64 + tx * 4]);
        const float qr[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w}, qi[8] = {ia.x, ia.y, ia.z, ia.w, ib.x, ib.y, ib.z, ib.w};
        const float er[8] = {ua.x, ua.y, ua.z, ua.w, ub.x, ub.y, ub.z, ub.w}, ei[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int a = 0; a < 8; ++a)
#pragma unroll
          for (int b = 0; b < 8; ++b) acc[a][b] = lp_acc_c<KIND>(acc[a][b], qr[a], qi[a], er[b], ei[b], p.p);
      }
    } else {
    for (int j = 0; j < jn; ++j) {
      // micro-tile rows / columns: 4 at lid*4 and 4 at 64 + lid*4 (conflict-free 128-bit reads)
      const float4 qa = *reinterpret_cast<const float4*>(&sQ[j][ty * 4]), qb = *reinterpret_cast<const float4*>(&sQ[j][64 + ty * 4]);
      const float4 ea = *reinterpret_cast<const float4*>(&sE[j][tx * 4]), eb = *reinterpret_cast<const float4*>(&sE[j][64 + tx * 4]);
      const float qq[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
      const float ee[8] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y, eb.z, eb.w};
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = lp_acc<KIND>(acc[a][b], qq[a], ee[b], p.p);
    }
    }
  }
  // strict "closer than the true triple" counts
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    const int ql = (a < 4) ? ty * 4 + a : 64 + ty * 4 + (a - 4);
    const int64_t q = q0 + ql;
    if (q >= p.Q) continue;
    const float thr = p.thr[q];
    int c = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int el = (b < 4) ? tx * 4 + b : 64 + tx * 4 + (b - 4);
      c += (e0 + el < p.ent_hi && lp_clip(acc[a][b], KIND) < thr) ? 1 : 0;
    }
    if (c) atomicAdd(&s_cnt[ql], c);
  }
  __syncthreads();
  if (tid < LP_T && q0 + tid < p.Q && s_cnt[tid]) atomicAdd(p.count + q0 + tid, s_cnt[tid]);
}

// known positives: one thread per (query, filter entry); undo the ones the tile kernel counted
template <int KIND, bool CPLX>
__global__ void __launch_bounds__(256) k_rank_lp_filter(const LpParams p) {
  const int64_t q = blockIdx.x;
  const int64_t lo = p.f_off[q], hi = p.f_end[q];
  const float thr = p.thr[q];
  const float* qv = p.Qv + q * (int64_t) p.ldq;
  int c = 0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const int64_t e = p.f_ids[i];
    if (e < p.ent_lo || e >= p.ent_hi) continue;
    const float* er = p.E + e * (int64_t) p.ldE;
    float s = 0.f;
    if constexpr (CPLX) { for (int j = 0; j < p.d; j += 2) s = lp_acc_c<KIND>(s, qv[j], qv[j + 1], __ldg(er + j), __ldg(er + j + 1), p.p); }
    else { for (int j = 0; j < p.d; ++j) s = lp_acc<KIND>(s, qv[j], __ldg(er + j), p.p); }
    c += (lp_clip(s, KIND) < thr) ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicSub(p.count + q, c);
}

// ---- host side ----------------------------------------------------------------------------------
inline bool rank_lp_eligible(const kge_cfg* c) {
  return (c->model == KGE_TRANSE || c->model == KGE_UM || c->model == KGE_ROTATE) &&
         (c->score == KGE_SCORE_LP || c->score == KGE_SCORE_LP_POW);
}
constexpr int64_t LP_QBLK = 65536;
struct LpWs { size_t off_q, off_thr, total; int ldq; int64_t qblk; };
inline void layout_lp_ws(const kge_cfg* c, int64_t Q, LpWs& L) {
  L.ldq = ((c->model == KGE_ROTATE ? 2 : 1) * c->d_ent + 3) / 4 * 4;
  const int64_t qpad = (Q + LP_T - 1) / LP_T * LP_T;
  L.qblk = qpad < LP_QBLK ? qpad : LP_QBLK;
  L.off_q = 0;
  L.off_thr = (sizeof(float) * (size_t) L.qblk * L.ldq + 255) / 256 * 256;
  L.total = L.off_thr + (sizeof(float) * (size_t) L.qblk + 255) / 256 * 256;
}

template <int KIND, bool CPLX>
inline int rank_lp_launch2(const LpParams& p, int64_t span, cudaStream_t st) {
  k_rank_lp_prep<KIND, CPLX><<<(unsigned) ((p.Q + 127) / 128), 128, 0, st>>>(p);
  const dim3 grid((unsigned) ((span + LP_T - 1) / LP_T), (unsigned) ((p.Q + LP_T - 1) / LP_T));
  k_rank_lp_tile<KIND, CPLX><<<grid, LP_THREADS, 0, st>>>(p);
  if (p.f_off) k_rank_lp_filter<KIND, CPLX><<<(unsigned) p.Q, 256, 0, st>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("Lp ranking launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
  return KGE_OK;
}
template <int KIND>
inline int rank_lp_launch(const LpParams& p, int64_t span, cudaStream_t st) {
  return p.md.model == KGE_ROTATE ? rank_lp_launch2<KIND, true>(p, span, st) : rank_lp_launch2<KIND, false>(p, span, st);
}

inline int rank_lp_tiled(const kge_cfg* cfg, const kge_tables* tables, const void* queries, int64_t Q, int side,
                         const kge_filter* filter, int64_t ent_lo, int64_t ent_hi, int32_t* count_out, void* ws,
                         size_t ws_bytes, cudaStream_t st) {
  LpWs L;
  layout_lp_ws(cfg, Q, L);
  KGE_REQUIRE(ws && ws_bytes >= L.total, KGE_EWORKSPACE, "rank workspace %zu < required %zu bytes", ws_bytes, L.total);
  KGE_REQUIRE(((uintptr_t) ws & 255) == 0, KGE_EINVAL, "workspace must be 256-byte aligned");
  char* wsb = reinterpret_cast<char*>(ws);
  LpParams p;
  fill_model_dev(cfg, tables, p.md);
  p.id64 = cfg->id_dtype; p.side = side;
  p.Qv = reinterpret_cast<float*>(wsb + L.off_q); p.ldq = L.ldq;
  p.thr = reinterpret_cast<float*>(wsb + L.off_thr);
  p.E = tables->t[KGE_T_ENT].w; p.ldE = tables->t[KGE_T_ENT].ld;
  p.d = (cfg->model == KGE_ROTATE ? 2 : 1) * cfg->d_ent;
  p.ent_lo = ent_lo; p.ent_hi = ent_hi;
  p.p = cfg->score_p;
  const int kind = isinf(cfg->score_p) ? LPK_LINF : (cfg->score_p == 1.f ? LPK_L1 : (cfg->score_p == 2.f ? LPK_L2 : LPK_GEN));
  for (int64_t q0 = 0; q0 < Q; q0 += L.qblk) {
    p.Q = (Q - q0 < L.qblk) ? Q - q0 : L.qblk;
    p.queries = (const char*) queries + (size_t) q0 * 3 * (cfg->id_dtype ? 8 : 4);
    p.count = count_out + q0;
    p.f_off = (filter && filter->off) ? filter->off + q0 : nullptr;
    p.f_end = (filter && filter->off) ? (filter->end ? filter->end : filter->off + 1) + q0 : nullptr;
    p.f_ids = filter ? filter->ids : nullptr;
    int rc;
    switch (kind) {
      case LPK_L1: rc = rank_lp_launch<LPK_L1>(p, ent_hi - ent_lo, st); break;
      case LPK_L2: rc = rank_lp_launch<LPK_L2>(p, ent_hi - ent_lo, st); break;
      case LPK_LINF: rc = rank_lp_launch<LPK_LINF>(p, ent_hi - ent_lo, st); break;
      default: rc = rank_lp_launch<LPK_GEN>(p, ent_hi - ent_lo, st);
    }
    if (rc) return rc;
  }
  return KGE_OK;
}

}  // namespace kge
