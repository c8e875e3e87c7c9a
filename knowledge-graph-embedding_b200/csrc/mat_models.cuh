// mat_models.cuh — "matrix family" models: SE, TransR, RESCAL (one d x d' matrix per
// relation).  One warp evaluates one triple; vectors live in registers as Row<NV> in a
// STRIDED layout (register slot t of a lane holds element lane + 32*t), matrices are
// read through a generic pointer (shared memory when the CTA staged the relation's
// matrix, global/L2 otherwise).  Never materialises per-triple matrices (the reference
// gathers (n, d, d') tensors: TransR.py:182, SE.py:155-161, RESCAL.py:164-174).
//
// forward() restates each model's score_hrt; backward() is the hand-derived gradient of
// SURVEY.md Appendix E and also returns the two "pair vectors" PH, PT whose outer
// products with the h / t entity rows form the matrix gradient (see k_update_mat).
#pragma once
#include "models.cuh"

namespace kge {

// ---- strided-layout helpers ---------------------------------------------------------
template <int NV>
__device__ __forceinline__ float& elt(Row<NV>& r, int t) {
  float4& c = r.v[t >> 2];
  switch (t & 3) { case 0: return c.x; case 1: return c.y; case 2: return c.z; default: return c.w; }
}
template <int NV>
__device__ __forceinline__ float elt(const Row<NV>& r, int t) {
  const float4& c = r.v[t >> 2];
  switch (t & 3) { case 0: return c.x; case 1: return c.y; case 2: return c.z; default: return c.w; }
}
template <int NV>
__device__ __forceinline__ void sload(Row<NV>& r, const float* __restrict__ base, int cols, int lane) {
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) { const int e = lane + 32 * t; elt(r, t) = (e < cols) ? base[e] : 0.f; }
}
template <int NV>
__device__ __forceinline__ void sstore(const Row<NV>& r, float* __restrict__ base, int cols, int lane) {
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) { const int e = lane + 32 * t; if (e < cols) base[e] = elt(r, t); }
}
template <int NV>
__device__ __forceinline__ void smask(Row<NV>& r, int cols, int lane) {
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) if (lane + 32 * t >= cols) elt(r, t) = 0.f;
}

// Matrix access context of one warp.
struct MatCtx {
  const float* M0;    // SE: rel_proj_h[r]; TransR: rel_proj[r]; RESCAL: rel_inter[r]
  const float* M1;    // SE: rel_proj_t[r]
  int ld0, ld1;       // row strides of M0 / M1 (odd when staged in shared memory)
  float* scratch;     // per-warp vector used to broadcast a matvec operand (>= dmax floats)
};

// out[e] = sum_{i < n_in} vec[i] * M[i*ld + e],  e = lane + 32 t < n_out   (vec . M)
template <int NV>
__device__ __forceinline__ void matvec_rowin(Row<NV>& out, const Row<NV>& vec, const float* M, int ld, int n_in,
                                             int n_out, float* scratch, int lane) {
  __syncwarp();
  sstore(vec, scratch, n_in, lane);
  __syncwarp();
  float acc[4 * NV];
  int col[4 * NV];
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) { acc[t] = 0.f; col[t] = min(lane + 32 * t, n_out - 1); }
#pragma unroll 4
  for (int i = 0; i < n_in; ++i) {
    const float v = scratch[i];
    const float* Mi = M + (size_t) i * ld;
#pragma unroll
    for (int t = 0; t < 4 * NV; ++t) acc[t] = fmaf(v, Mi[col[t]], acc[t]);
  }
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) elt(out, t) = (lane + 32 * t < n_out) ? acc[t] : 0.f;
}

// out[e] = sum_{j < n_in} M[e*ld + j] * vec[j],  e = lane + 32 t < n_out   (M . vec)
template <int NV>
__device__ __forceinline__ void matvec_colin(Row<NV>& out, const Row<NV>& vec, const float* M, int ld, int n_out,
                                             int n_in, float* scratch, int lane) {
  __syncwarp();
  sstore(vec, scratch, n_in, lane);
  __syncwarp();
  float acc[4 * NV];
  const float* row[4 * NV];
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) { acc[t] = 0.f; row[t] = M + (size_t) min(lane + 32 * t, n_out - 1) * ld; }
#pragma unroll 4
  for (int j = 0; j < n_in; ++j) {
    const float v = scratch[j];
#pragma unroll
    for (int t = 0; t < 4 * NV; ++t) acc[t] = fmaf(row[t][j], v, acc[t]);
  }
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) elt(out, t) = (lane + 32 * t < n_out) ? acc[t] : 0.f;
}

// clip_constraint(p=2, value=1) on a warp-distributed vector (constraint.py:93-99);
// returns the norm when clipped, 0 when passed through.
template <int NV>
__device__ __forceinline__ float clip1(Row<NV>& p, bool on) {
  if (!on) return 0.f;
  const float n = sqrtf(warp_sum(row_dot_partial(p, p)));
  if (n < 1.0f) return 0.f;
  const float inv = 1.0f / fmaxf(n, 1e-9f);
#pragma unroll
  for (int i = 0; i < NV; ++i) p.v[i] = inv * p.v[i];
  return n;
}
// back through c = p/n:  g_p = (g_c - (g_c . c) c) / n
template <int NV>
__device__ __forceinline__ void clip1_bwd(Row<NV>& g, const Row<NV>& c, float n) {
  if (n <= 0.f) return;
  const float d = warp_sum(row_dot_partial(g, c));
  const float inv = 1.0f / n;
#pragma unroll
  for (int i = 0; i < NV; ++i) g.v[i] = inv * (g.v[i] - d * c.v[i]);
}

// Entity / relation-vector bundles in the strided layout (records are plain row-major).
template <int NV> struct SEnt {
  Row<NV> e;
  __device__ __forceinline__ void load(const ModelDev& md, int64_t id, int lane) {
    sload(e, md.t[KGE_T_ENT].w + id * (int64_t) md.t[KGE_T_ENT].ld, md.t[KGE_T_ENT].cols, lane);
  }
};

struct MatSaved {
  ScoreSaved sc;
  float nh, nt;       // clip norms (0 = not clipped)
};

// ---- projection models: x = f_h(e_h) [+ r], y = f_t(e_t), score_fn(x, y) -------------
// proj_fwd(side, e) -> c (and the clip norm); proj_bwd(side, g_c, c, n) -> pair vector P
// and the entity gradient.  side 0 = head role, 1 = tail role.
struct MTransR {        // TransR.py:178-191: c = clip1(e M_r), x = c_h + r, y = c_t
  static constexpr int KIND = KGE_TRANSR;
  static constexpr bool BILINEAR = false, HAS_REL_VEC = true, TWO_MATS = false;
  static __device__ __forceinline__ int d_in(const ModelDev& md) { return md.d_ent; }
  static __device__ __forceinline__ int d_out(const ModelDev& md) { return md.d_rel; }
  template <int NV>
  static __device__ __forceinline__ float proj_fwd(const ModelDev& md, const MatCtx& cx, int, const Row<NV>& e,
                                                   Row<NV>& c, int lane) {
    matvec_rowin(c, e, cx.M0, cx.ld0, md.d_ent, md.d_rel, cx.scratch, lane);
    return clip1(c, md.constraint != 0);
  }
  template <int NV>
  static __device__ __forceinline__ void proj_bwd(const ModelDev& md, const MatCtx& cx, int, Row<NV>& P,
                                                  const Row<NV>& c, float n, Row<NV>& ge, int lane) {
    clip1_bwd(P, c, n);
    matvec_colin(ge, P, cx.M0, cx.ld0, md.d_ent, md.d_rel, cx.scratch, lane);
  }
  // squared norm of this triple's matrix-gradient occurrence(s): one lookup row holds
  // e_h (x) PH + e_t (x) PT  (rel_proj is looked up once per triple, TransR.py:182)
  template <int NV>
  static __device__ __forceinline__ void mat_norm(const Row<NV>& eh, const Row<NV>& et, const Row<NV>& PH,
                                                  const Row<NV>& PT, float* matn) {
    float a = row_dot_partial(eh, eh), b = row_dot_partial(PH, PH);
    float c = row_dot_partial(et, et), d = row_dot_partial(PT, PT);
    float e = row_dot_partial(eh, et), f = row_dot_partial(PH, PT);
    warp_sum2(a, b); warp_sum2(c, d); warp_sum2(e, f);
    matn[1] += a * b + c * d + 2.f * e * f;       // warp-uniform; matn[0] -> slot REL_A, matn[1] -> REL_B
  }
};

struct MSE {            // SE.py:155-161: x = R^h_r e_h, y = R^t_r e_t
  static constexpr int KIND = KGE_SE;
  static constexpr bool BILINEAR = false, HAS_REL_VEC = false, TWO_MATS = true;
  static __device__ __forceinline__ int d_in(const ModelDev& md) { return md.d_ent; }
  static __device__ __forceinline__ int d_out(const ModelDev& md) { return md.d_ent; }
  template <int NV>
  static __device__ __forceinline__ float proj_fwd(const ModelDev& md, const MatCtx& cx, int side, const Row<NV>& e,
                                                   Row<NV>& c, int lane) {
    matvec_colin(c, e, side ? cx.M1 : cx.M0, side ? cx.ld1 : cx.ld0, md.d_ent, md.d_ent, cx.scratch, lane);
    return 0.f;
  }
  template <int NV>
  static __device__ __forceinline__ void proj_bwd(const ModelDev& md, const MatCtx& cx, int side, Row<NV>& P,
                                                  const Row<NV>&, float, Row<NV>& ge, int lane) {
    matvec_rowin(ge, P, side ? cx.M1 : cx.M0, side ? cx.ld1 : cx.ld0, md.d_ent, md.d_ent, cx.scratch, lane);
  }
  template <int NV>
  static __device__ __forceinline__ void mat_norm(const Row<NV>& eh, const Row<NV>& et, const Row<NV>& PH,
                                                  const Row<NV>& PT, float* matn) {
    float a = row_dot_partial(eh, eh), b = row_dot_partial(PH, PH);
    float c = row_dot_partial(et, et), d = row_dot_partial(PT, PT);
    warp_sum2(a, b); warp_sum2(c, d);
    matn[0] += a * b;
    matn[1] += c * d;
  }
};

struct MRESCAL {        // RESCAL.py:164-174: f = e_h^T R_r e_t
  static constexpr int KIND = KGE_RESCAL;
  static constexpr bool BILINEAR = true, HAS_REL_VEC = false, TWO_MATS = false;
  static __device__ __forceinline__ int d_in(const ModelDev& md) { return md.d_ent; }
  static __device__ __forceinline__ int d_out(const ModelDev& md) { return md.d_ent; }
};

}  // namespace kge
