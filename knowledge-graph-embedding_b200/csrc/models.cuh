// models.cuh — "vector family" models (every table row is a vector): UM, TransE,
// TransH, TransD, RotatE, DistMult.  One warp evaluates one triple; rows live in
// registers as Row<NV>.  forward() restates each model's score_hrt, backward() is
// the hand-derived gradient of SURVEY.md Appendix E (the oracle uses autograd).
#pragma once
#include "score.cuh"

namespace kge {

struct TabDev {
  float* w;
  int64_t rows;
  int cols, ld;
};
struct ModelDev {
  int model, constraint;
  int d_ent, d_rel;
  float cw;            // constraint_weight
  float rotate_scale;  // pi / limit
  ScoreDev sc;
  TabDev t[KGE_MAX_TABLES];
  int rld[KGE_MAX_TABLES];   // row stride of the gradient-record buffers (round4(cols))
  // entity tables sharded over peer GPUs (kge_shard): row id -> owner id / shard_rows, base shard_w[slot][owner]
  uint32_t shard_rows;       // 0 = not sharded
  const float* shard_w[2][KGE_MAX_PEERS];
};
// base pointer and local row of an entity-table row (slot 0 / 1)
__device__ __forceinline__ const float* ent_base(const ModelDev& md, int slot, int64_t& id) {
  if (md.shard_rows == 0) return md.t[slot].w;
  const uint32_t o = (uint32_t) id / md.shard_rows;
  id -= (int64_t) o * md.shard_rows;
  return md.shard_w[slot][o];
}

// ---- register bundles ------------------------------------------------------------
template <int NV> struct Ent1 {
  Row<NV> e;
  static constexpr int NROWS = 1;
  template <class F> __device__ __forceinline__ void each(F&& f) { f(e, 0); }
  __device__ __forceinline__ void zero() { e.zero(); }
  __device__ __forceinline__ void add(const Ent1& o) { row_add(e, o.e); }
  __device__ __forceinline__ void sumsq(float* n) const { n[KGE_T_ENT] += row_dot_partial(e, e); }
  __device__ __forceinline__ void load(const ModelDev& md, int64_t id, int lane) {
    const float* b = ent_base(md, KGE_T_ENT, id);
    load_row(e, b, id, md.t[KGE_T_ENT].ld, md.t[KGE_T_ENT].cols, lane);
  }
  // store as gradient record `rec` of the entity record buffers
  __device__ __forceinline__ void store(const ModelDev& md, float* const* recs, int64_t rec, int lane) const {
    store_row(e, recs[KGE_T_ENT], rec, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane);
  }
  __device__ __forceinline__ void store_to(const ModelDev& md, float* b0, float*, int64_t rec, int lane) const {
    store_row(e, b0, rec, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane);
  }
};
template <int NV> struct Ent2 {
  Row<NV> e, p;
  static constexpr int NROWS = 2;
  template <class F> __device__ __forceinline__ void each(F&& f) { f(e, 0); f(p, 1); }
  __device__ __forceinline__ void zero() { e.zero(); p.zero(); }
  __device__ __forceinline__ void add(const Ent2& o) { row_add(e, o.e); row_add(p, o.p); }
  __device__ __forceinline__ void sumsq(float* n) const {
    n[KGE_T_ENT] += row_dot_partial(e, e);
    n[KGE_T_ENT_PROJ] += row_dot_partial(p, p);
  }
  __device__ __forceinline__ void load(const ModelDev& md, int64_t id, int lane) {
    int64_t id2 = id;
    const float* b = ent_base(md, KGE_T_ENT, id);
    const float* b2 = ent_base(md, KGE_T_ENT_PROJ, id2);
    load_row(e, b, id, md.t[KGE_T_ENT].ld, md.t[KGE_T_ENT].cols, lane);
    load_row(p, b2, id2, md.t[KGE_T_ENT_PROJ].ld, md.t[KGE_T_ENT_PROJ].cols, lane);
  }
  __device__ __forceinline__ void store(const ModelDev& md, float* const* recs, int64_t rec, int lane) const {
    store_row(e, recs[KGE_T_ENT], rec, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane);
    store_row(p, recs[KGE_T_ENT_PROJ], rec, md.rld[KGE_T_ENT_PROJ], md.t[KGE_T_ENT_PROJ].cols, lane);
  }
  __device__ __forceinline__ void store_to(const ModelDev& md, float* b0, float* b1, int64_t rec, int lane) const {
    store_row(e, b0, rec, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane);
    store_row(p, b1, rec, md.rld[KGE_T_ENT_PROJ], md.t[KGE_T_ENT_PROJ].cols, lane);
  }
};
template <int NV> struct Rel0 {
  static constexpr int NROWS = 0;
  template <class F> __device__ __forceinline__ void each(F&&) {}
  __device__ __forceinline__ void zero() {}
  __device__ __forceinline__ void add(const Rel0&) {}
  __device__ __forceinline__ void sumsq(float*) const {}
  __device__ __forceinline__ void load(const ModelDev&, int64_t, int) {}
  __device__ __forceinline__ void store(const ModelDev&, float* const*, int64_t, int) const {}
};
template <int NV> struct Rel1 {
  Row<NV> a;
  static constexpr int NROWS = 1;
  template <class F> __device__ __forceinline__ void each(F&& f) { f(a, 0); }
  __device__ __forceinline__ void zero() { a.zero(); }
  __device__ __forceinline__ void add(const Rel1& o) { row_add(a, o.a); }
  __device__ __forceinline__ void sumsq(float* n) const { n[KGE_T_REL_A] += row_dot_partial(a, a); }
  __device__ __forceinline__ void load(const ModelDev& md, int64_t id, int lane) {
    load_row(a, md.t[KGE_T_REL_A].w, id, md.t[KGE_T_REL_A].ld, md.t[KGE_T_REL_A].cols, lane);
  }
  __device__ __forceinline__ void store(const ModelDev& md, float* const* recs, int64_t rec, int lane) const {
    store_row(a, recs[KGE_T_REL_A], rec, md.rld[KGE_T_REL_A], md.t[KGE_T_REL_A].cols, lane);
  }
};
template <int NV> struct Rel2 {
  Row<NV> a, b;
  static constexpr int NROWS = 2;
  template <class F> __device__ __forceinline__ void each(F&& f) { f(a, 0); f(b, 1); }
  __device__ __forceinline__ void zero() { a.zero(); b.zero(); }
  __device__ __forceinline__ void add(const Rel2& o) { row_add(a, o.a); row_add(b, o.b); }
  __device__ __forceinline__ void sumsq(float* n) const {
    n[KGE_T_REL_A] += row_dot_partial(a, a);
    n[KGE_T_REL_B] += row_dot_partial(b, b);
  }
  __device__ __forceinline__ void load(const ModelDev& md, int64_t id, int lane) {
    load_row(a, md.t[KGE_T_REL_A].w, id, md.t[KGE_T_REL_A].ld, md.t[KGE_T_REL_A].cols, lane);
    load_row(b, md.t[KGE_T_REL_B].w, id, md.t[KGE_T_REL_B].ld, md.t[KGE_T_REL_B].cols, lane);
  }
  __device__ __forceinline__ void store(const ModelDev& md, float* const* recs, int64_t rec, int lane) const {
    store_row(a, recs[KGE_T_REL_A], rec, md.rld[KGE_T_REL_A], md.t[KGE_T_REL_A].cols, lane);
    store_row(b, recs[KGE_T_REL_B], rec, md.rld[KGE_T_REL_B], md.t[KGE_T_REL_B].cols, lane);
  }
};
// RotatE relation row: phases live at float2 granularity under each complex chunk.
// As parameters: a = (cos0, cos1, sin0, sin1) of the two complex elements the lane's
// chunk holds.  As gradient: a = (gphi0, gphi1, 0, 0).
template <int NV> struct RelRot {
  Row<NV> a;
  static constexpr int NROWS = 1;
  template <class F> __device__ __forceinline__ void each(F&& f) { f(a, 0); }
  __device__ __forceinline__ void zero() { a.zero(); }
  __device__ __forceinline__ void add(const RelRot& o) { row_add(a, o.a); }
  __device__ __forceinline__ void sumsq(float* n) const { n[KGE_T_REL_A] += row_dot_partial(a, a); }
  __device__ __forceinline__ void load(const ModelDev& md, int64_t id, int lane) {
    const float* base = md.t[KGE_T_REL_A].w + id * (int64_t) md.t[KGE_T_REL_A].ld;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      int c = lane + i * WARP;               // complex elements 2c, 2c+1
      float p0 = (2 * c < md.d_ent) ? __ldg(base + 2 * c) : 0.f;
      float p1 = (2 * c + 1 < md.d_ent) ? __ldg(base + 2 * c + 1) : 0.f;
      float s0, c0, s1, c1;
      sincosf(p0 * md.rotate_scale, &s0, &c0);     // r_emb / limit * pi, RotatE.py:160
      sincosf(p1 * md.rotate_scale, &s1, &c1);
      a.v[i] = f4(c0, c1, s0, s1);
    }
  }
  __device__ __forceinline__ void store(const ModelDev& md, float* const* recs, int64_t rec, int lane) const {
    float* base = recs[KGE_T_REL_A] + rec * (int64_t) md.rld[KGE_T_REL_A];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      int c = lane + i * WARP;
      if (2 * c < md.d_ent) base[2 * c] = a.v[i].x;
      if (2 * c + 1 < md.d_ent) base[2 * c + 1] = a.v[i].y;
    }
  }
};

template <int NV> struct FwdSaved {
  ScoreSaved sc;
  float s0, s1, s2, s3;   // model scalars (dots, norms)
};

// =============================== UM (UM.py:134-137) ===============================
struct MUM {
  static constexpr int KIND = KGE_UM;
  static constexpr bool CPLX = false;
  template <int NV> using Ent = Ent1<NV>;
  template <int NV> using Rel = Rel0<NV>;
  template <int NV>
  static __device__ __forceinline__ float forward(const ModelDev& md, const Ent<NV>& H, const Rel<NV>&,
                                                  const Ent<NV>& T, int, Row<NV>& x, Row<NV>& y, FwdSaved<NV>& sv) {
    x = H.e; y = T.e;
    return score_fwd<NV, false>(md.sc, x, y, sv.sc);
  }
  template <int NV>
  static __device__ __forceinline__ void backward(const ModelDev& md, const Ent<NV>&, const Rel<NV>&, const Ent<NV>&,
                                                  const Row<NV>& x, const Row<NV>& y, const FwdSaved<NV>& sv, float g,
                                                  int, Ent<NV>& gH, Rel<NV>&, Ent<NV>& gT) {
    score_bwd<NV, false>(md.sc, x, y, sv.sc, g, gH.e, gT.e);
  }
};

// ============================ TransE (TransE.py:151-155) ==========================
struct MTransE {
  static constexpr int KIND = KGE_TRANSE;
  static constexpr bool CPLX = false;
  template <int NV> using Ent = Ent1<NV>;
  template <int NV> using Rel = Rel1<NV>;
  template <int NV>
  static __device__ __forceinline__ float forward(const ModelDev& md, const Ent<NV>& H, const Rel<NV>& R,
                                                  const Ent<NV>& T, int, Row<NV>& x, Row<NV>& y, FwdSaved<NV>& sv) {
#pragma unroll
    for (int i = 0; i < NV; ++i) { x.v[i] = H.e.v[i] + R.a.v[i]; y.v[i] = T.e.v[i]; }
    return score_fwd<NV, false>(md.sc, x, y, sv.sc);
  }
  template <int NV>
  static __device__ __forceinline__ void backward(const ModelDev& md, const Ent<NV>&, const Rel<NV>&, const Ent<NV>&,
                                                  const Row<NV>& x, const Row<NV>& y, const FwdSaved<NV>& sv, float g,
                                                  int, Ent<NV>& gH, Rel<NV>& gR, Ent<NV>& gT) {
    score_bwd<NV, false>(md.sc, x, y, sv.sc, g, gH.e, gT.e);
    gR.a = gH.e;
  }
};

// ============================ TransH (TransH.py:173-185) ==========================
// R.a = rel_emb row r, R.b = rel_hyper row w.
struct MTransH {
  static constexpr int KIND = KGE_TRANSH;
  static constexpr bool CPLX = false;
  template <int NV> using Ent = Ent1<NV>;
  template <int NV> using Rel = Rel2<NV>;
  template <int NV>
  static __device__ __forceinline__ float forward(const ModelDev& md, const Ent<NV>& H, const Rel<NV>& R,
                                                  const Ent<NV>& T, int, Row<NV>& x, Row<NV>& y, FwdSaved<NV>& sv) {
    float wh = row_dot_partial(R.b, H.e), wt = row_dot_partial(R.b, T.e);
    warp_sum2(wh, wt);
    sv.s0 = wh; sv.s1 = wt;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      x.v[i] = (H.e.v[i] - wh * R.b.v[i]) + R.a.v[i];
      y.v[i] = T.e.v[i] - wt * R.b.v[i];
    }
    return score_fwd<NV, false>(md.sc, x, y, sv.sc);
  }
  template <int NV>
  static __device__ __forceinline__ void backward(const ModelDev& md, const Ent<NV>& H, const Rel<NV>& R, const Ent<NV>& T,
                                                  const Row<NV>& x, const Row<NV>& y, const FwdSaved<NV>& sv, float g,
                                                  int, Ent<NV>& gH, Rel<NV>& gR, Ent<NV>& gT) {
    Row<NV> gx, gy;
    score_bwd<NV, false>(md.sc, x, y, sv.sc, g, gx, gy);
    float wgx = row_dot_partial(R.b, gx), wgy = row_dot_partial(R.b, gy);
    warp_sum2(wgx, wgy);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      gH.e.v[i] = gx.v[i] - wgx * R.b.v[i];
      gT.e.v[i] = gy.v[i] - wgy * R.b.v[i];
      gR.a.v[i] = gx.v[i];
      gR.b.v[i] = -((sv.s0 * gx.v[i] + wgx * H.e.v[i]) + (sv.s1 * gy.v[i] + wgy * T.e.v[i]));
    }
  }
};

// ============================ TransD (TransD.py:194-222) ==========================
// (r_p e_p^T + I) e == (e_p . e) r_p + pad/trunc(e -> d_r)  (rank-1 identity).
// R.a = rel_emb r, R.b = rel_proj r_p; Ent = {e, p}.
struct MTransD {
  static constexpr int KIND = KGE_TRANSD;
  static constexpr bool CPLX = false;
  template <int NV> using Ent = Ent2<NV>;
  template <int NV> using Rel = Rel2<NV>;

  template <int NV>
  static __device__ __forceinline__ float forward(const ModelDev& md, const Ent<NV>& H, const Rel<NV>& R,
                                                  const Ent<NV>& T, int lane, Row<NV>& x, Row<NV>& y, FwdSaved<NV>& sv) {
    float ah = row_dot_partial(H.p, H.e), at = row_dot_partial(T.p, T.e);
    warp_sum2(ah, at);
    Row<NV> eh = H.e, et = T.e;
    if (md.d_rel < md.d_ent) { row_mask(eh, md.d_rel, lane); row_mask(et, md.d_rel, lane); }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      x.v[i] = ah * R.b.v[i] + eh.v[i];
      y.v[i] = at * R.b.v[i] + et.v[i];
    }
    float nh = row_dot_partial(x, x), nt = row_dot_partial(y, y);
    warp_sum2(nh, nt);
    nh = sqrtf(nh); nt = sqrtf(nt);
    sv.s0 = ah; sv.s1 = at;
    // clip_constraint(p=2, value=1): constraint.py:93-99 (TransD.py:218-220)
    sv.s2 = (md.constraint && !(nh < 1.0f)) ? nh : 0.f;     // 0 = not clipped
    sv.s3 = (md.constraint && !(nt < 1.0f)) ? nt : 0.f;
    float ih = sv.s2 > 0.f ? 1.0f / fmaxf(nh, 1e-9f) : 1.0f;
    float it = sv.s3 > 0.f ? 1.0f / fmaxf(nt, 1e-9f) : 1.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      x.v[i] = ih * x.v[i] + R.a.v[i];      // c_h + r
      y.v[i] = it * y.v[i];                 // c_t
    }
    return score_fwd<NV, false>(md.sc, x, y, sv.sc);
  }
  template <int NV>
  static __device__ __forceinline__ void backward(const ModelDev& md, const Ent<NV>& H, const Rel<NV>& R, const Ent<NV>& T,
                                                  const Row<NV>& x, const Row<NV>& y, const FwdSaved<NV>& sv, float g,
                                                  int lane, Ent<NV>& gH, Rel<NV>& gR, Ent<NV>& gT) {
    Row<NV> gx, gy;
    score_bwd<NV, false>(md.sc, x, y, sv.sc, g, gx, gy);
    gR.a = gx;
    // back through the clip: c = p/n  =>  g_p = (g_c - (g_c . c) c) / n
    Row<NV> gph = gx, gpt = gy;
    if (sv.s2 > 0.f || sv.s3 > 0.f) {
      float dh = 0.f, dt = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        dh += dot4(gx.v[i], x.v[i] - R.a.v[i]);
        dt += dot4(gy.v[i], y.v[i]);
      }
      warp_sum2(dh, dt);
      if (sv.s2 > 0.f) {
        float inv = 1.0f / sv.s2;
#pragma unroll
        for (int i = 0; i < NV; ++i) gph.v[i] = inv * (gx.v[i] - dh * (x.v[i] - R.a.v[i]));
      }
      if (sv.s3 > 0.f) {
        float inv = 1.0f / sv.s3;
#pragma unroll
        for (int i = 0; i < NV; ++i) gpt.v[i] = inv * (gy.v[i] - dt * y.v[i]);
      }
    }
    float gah = row_dot_partial(R.b, gph), gat = row_dot_partial(R.b, gpt);
    warp_sum2(gah, gat);
    Row<NV> th = gph, tt = gpt;            // pad/trunc(g_p -> d_e)
    if (md.d_ent < md.d_rel) { row_mask(th, md.d_ent, lane); row_mask(tt, md.d_ent, lane); }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      gR.b.v[i] = sv.s0 * gph.v[i] + sv.s1 * gpt.v[i];
      gH.p.v[i] = gah * H.e.v[i];
      gT.p.v[i] = gat * T.e.v[i];
      gH.e.v[i] = gah * H.p.v[i] + th.v[i];
      gT.e.v[i] = gat * T.p.v[i] + tt.v[i];
    }
  }
};

// ============================ RotatE (RotatE.py:150-165) ==========================
// ent rows are (d, 2) re/im interleaved: chunk = (re0, im0, re1, im1).
struct MRotatE {
  static constexpr int KIND = KGE_ROTATE;
  static constexpr bool CPLX = true;
  template <int NV> using Ent = Ent1<NV>;
  template <int NV> using Rel = RelRot<NV>;
  template <int NV>
  static __device__ __forceinline__ float forward(const ModelDev& md, const Ent<NV>& H, const Rel<NV>& R,
                                                  const Ent<NV>& T, int, Row<NV>& x, Row<NV>& y, FwdSaved<NV>& sv) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float4 h = H.e.v[i], cs = R.a.v[i];
      x.v[i] = f4(h.x * cs.x - h.y * cs.z, h.x * cs.z + h.y * cs.x,
                  h.z * cs.y - h.w * cs.w, h.z * cs.w + h.w * cs.y);
      y.v[i] = T.e.v[i];
    }
    return score_fwd<NV, true>(md.sc, x, y, sv.sc);
  }
  template <int NV>
  static __device__ __forceinline__ void backward(const ModelDev& md, const Ent<NV>&, const Rel<NV>& R, const Ent<NV>&,
                                                  const Row<NV>& x, const Row<NV>& y, const FwdSaved<NV>& sv, float g,
                                                  int, Ent<NV>& gH, Rel<NV>& gR, Ent<NV>& gT) {
    Row<NV> gx;
    score_bwd<NV, true>(md.sc, x, y, sv.sc, g, gx, gT.e);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float4 q = gx.v[i], cs = R.a.v[i], xv = x.v[i];
      gH.e.v[i] = f4(q.x * cs.x + q.y * cs.z, -q.x * cs.z + q.y * cs.x,
                     q.z * cs.y + q.w * cs.w, -q.z * cs.w + q.w * cs.y);
      gR.a.v[i] = f4((-q.x * xv.y + q.y * xv.x) * md.rotate_scale,
                     (-q.z * xv.w + q.w * xv.z) * md.rotate_scale, 0.f, 0.f);
    }
  }
};

// =========================== DistMult (DistMult.py:142-146) =======================
struct MDistMult {
  static constexpr int KIND = KGE_DISTMULT;
  static constexpr bool CPLX = false;
  template <int NV> using Ent = Ent1<NV>;
  template <int NV> using Rel = Rel1<NV>;
  template <int NV>
  static __device__ __forceinline__ float forward(const ModelDev&, const Ent<NV>& H, const Rel<NV>& R,
                                                  const Ent<NV>& T, int, Row<NV>& x, Row<NV>& y, FwdSaved<NV>& sv) {
#pragma unroll
    for (int i = 0; i < NV; ++i) { x.v[i] = H.e.v[i] * R.a.v[i]; y.v[i] = T.e.v[i]; }
    float s = warp_sum(row_dot_partial(x, y));
    sv.sc.S = s; sv.sc.s = s; sv.sc.ties = 1.f;
    return s;
  }
  template <int NV>
  static __device__ __forceinline__ void backward(const ModelDev&, const Ent<NV>& H, const Rel<NV>& R, const Ent<NV>&,
                                                  const Row<NV>& x, const Row<NV>& y, const FwdSaved<NV>&, float g,
                                                  int, Ent<NV>& gH, Rel<NV>& gR, Ent<NV>& gT) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float4 gy = g * y.v[i];
      gH.e.v[i] = gy * R.a.v[i];
      gR.a.v[i] = gy * H.e.v[i];
      gT.e.v[i] = g * x.v[i];
    }
  }
};

}  // namespace kge
