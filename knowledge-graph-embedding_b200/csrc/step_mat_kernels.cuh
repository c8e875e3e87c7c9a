// step_mat_kernels.cuh — fused training step for the matrix-relation models (SE, TransR,
// RESCAL).  One CTA (8 warps) owns one positive and its k negatives: the relation's
// matrix (matrices) is staged once in shared memory (odd row stride: conflict-free in
// both matvec orientations), the positive's two projections are computed once (warp 0 /
// warp 1) and shared, each warp then runs whole negatives: 1 forward + 2 backward
// mat-vecs.  Emits the same entity / relation-vector gradient records as k_step plus one
// "pair vector" P per record slot; the matrix gradient  sum_pairs ent (x) P  is formed
// per relation by k_update_mat (grouped small GEMMs) — per-triple matrices
// are never materialised.
#pragma once
#include "step_kernels.cuh"
#include "mat_models.cuh"

namespace kge {

struct MatStepParams {
  StepParams s;
  float* recP;        // pair vectors, [(2B + B*k) x ldp], same slot numbering as entity records
  int ldp;
  int stage;          // matrices fit in shared memory
  int dpad;           // floats per scratch / shared-projection vector (multiple of 4)
  int proj;           // TransR / SE: the phase-structured kernel (step_proj_kernels.cuh) instead of k_step_mat
};

constexpr int MAT_ACC_ROWS = 5;   // GH, GT, GR, SPH, SPT

template <int NV>
__device__ __forceinline__ void rscale(Row<NV>& o, float a, const Row<NV>& x) {
#pragma unroll
  for (int i = 0; i < NV; ++i) o.v[i] = a * x.v[i];
}

template <class M, int NV, bool TRAIN>
__global__ void __launch_bounds__(STEP_THREADS) k_step_mat(const MatStepParams q) {
  extern __shared__ float4 smem4[];
  const StepParams& p = q.s;
  const ModelDev& md = p.md;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = blockIdx.x;
  const int k = p.k_eff;
  const int din = M::d_in(md), dout = M::d_out(md);
  const int s0 = (M::KIND == KGE_TRANSR) ? KGE_T_REL_B : KGE_T_REL_A;   // slot of the first matrix table

  const int64_t h = load_id(p.triples, 3 * i + 0, p.id64);
  const int64_t r = load_id(p.triples, 3 * i + 1, p.id64);
  const int64_t t = load_id(p.triples, 3 * i + 2, p.id64);

  // ---- shared-memory carve-up -------------------------------------------------------
  float* sm = reinterpret_cast<float*>(smem4);
  float* s_scr = sm;                              sm += STEP_WARPS * q.dpad;
  float* s_ch = sm;                               sm += q.dpad;
  float* s_ct = sm;                               sm += q.dpad;
  float* s_red = sm;                              sm += STEP_WARPS * 8;
  float* s_n = sm;                                sm += 8;
  float* s_neg = sm;                              sm += (k + 3) & ~3;
  float4* s_acc = reinterpret_cast<float4*>(sm);  sm += TRAIN ? STEP_WARPS * MAT_ACC_ROWS * NV * WARP * 4 : 0;
  float* s_mat = sm;

  MatCtx cx;
  const int mrows = (M::KIND == KGE_TRANSR) ? md.d_ent : md.d_ent;
  const int mcols = (M::KIND == KGE_TRANSR) ? md.d_rel : md.d_ent;
  const float* g0 = md.t[s0].w + r * (int64_t) md.t[s0].ld;
  const float* g1 = M::TWO_MATS ? md.t[KGE_T_REL_B].w + r * (int64_t) md.t[KGE_T_REL_B].ld : nullptr;
  if (q.stage) {
    const int lds = mcols | 1;
    for (int e = threadIdx.x; e < mrows * mcols; e += STEP_THREADS) {
      const int a = e / mcols, b = e - a * mcols;
      s_mat[a * lds + b] = __ldg(g0 + e);
      if (M::TWO_MATS) s_mat[mrows * lds + a * lds + b] = __ldg(g1 + e);
    }
    cx.M0 = s_mat; cx.M1 = s_mat + mrows * lds; cx.ld0 = cx.ld1 = lds;
  } else {
    cx.M0 = g0; cx.M1 = g1; cx.ld0 = cx.ld1 = mcols;
  }
  cx.scratch = s_scr + warp * q.dpad;

  Row<NV> Eh, Et, Rv;
  sload(Eh, md.t[KGE_T_ENT].w + h * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
  sload(Et, md.t[KGE_T_ENT].w + t * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
  if (M::HAS_REL_VEC) sload(Rv, md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld, dout, lane);
  else Rv.zero();
  __syncthreads();

  // ---- the positive's two projections, computed once and shared ----------------------
  // projection models: CH = f_h(e_h), CT = f_t(e_t); RESCAL: CH = V = e_h^T R, CT = U = R e_t
  if (warp == 0) {
    Row<NV> c; float n = 0.f;
    if constexpr (M::BILINEAR) matvec_rowin(c, Eh, cx.M0, cx.ld0, din, din, cx.scratch, lane);
    else n = M::template proj_fwd<NV>(md, cx, 0, Eh, c, lane);
    sstore(c, s_ch, dout, lane);
    if (lane == 0) s_n[0] = n;
  } else if (warp == 1) {
    Row<NV> c; float n = 0.f;
    if constexpr (M::BILINEAR) matvec_colin(c, Et, cx.M0, cx.ld0, din, din, cx.scratch, lane);
    else n = M::template proj_fwd<NV>(md, cx, 1, Et, c, lane);
    sstore(c, s_ct, dout, lane);
    if (lane == 0) s_n[1] = n;
  }
  __syncthreads();
  Row<NV> CH, CT;
  sload(CH, s_ch, dout, lane);
  sload(CT, s_ct, dout, lane);
  const float nh = s_n[0], nt = s_n[1];

  Row<NV> x, y;
  ScoreSaved sv;
  auto score_of = [&](const Row<NV>& ch, const Row<NV>& ct) -> float {
    if constexpr (M::BILINEAR) { return 0.f; }
    else {
#pragma unroll
      for (int u = 0; u < NV; ++u) { x.v[u] = ch.v[u] + Rv.v[u]; y.v[u] = ct.v[u]; }
      return score_fwd<NV, false>(md.sc, x, y, sv);
    }
  };
  float sp;
  if constexpr (M::BILINEAR) sp = warp_sum(row_dot_partial(Eh, CT));
  else sp = score_of(CH, CT);

  // forward of negative c (entity row N): returns score; CN / nN = its projection
  auto neg_forward = [&](bool head, const Row<NV>& N, Row<NV>& CN, float& nN) -> float {
    if constexpr (M::BILINEAR) {
      nN = 0.f;
      return warp_sum(head ? row_dot_partial(N, CT) : row_dot_partial(CH, N));
    } else {
      nN = M::template proj_fwd<NV>(md, cx, head ? 0 : 1, N, CN, lane);
      return head ? score_of(CN, CT) : score_of(CH, CN);
    }
  };

  float sm_max = 0.f, sm_inv = 1.f;
  const bool softmax = (p.loss.kind == KGE_LOSS_SELF_ADV);
  if (softmax) {              // loss.py:176-177 needs all k scores of this positive first
    for (int c = warp; c < k; c += STEP_WARPS) {
      const bool head = neg_is_head(p.sd, c);
      const int64_t e = p.neg_override ? p.neg_override[i * k + c] : sample_negative(p.sd, i, c, h, t);
      Row<NV> N, CN; float nN;
      sload(N, md.t[KGE_T_ENT].w + e * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
      const float sn = neg_forward(head, N, CN, nN);
      if (lane == 0) s_neg[c] = sn;
    }
    __syncthreads();
    float m = -INFINITY;
    for (int c = lane; c < k; c += WARP) m = fmaxf(m, p.loss.temperature * s_neg[c]);
    m = warp_max(m);
    float z = 0.f;
    for (int c = lane; c < k; c += WARP) z += expf(p.loss.temperature * s_neg[c] - m);
    z = warp_sum(z);
    sm_max = m; sm_inv = 1.0f / z;
  }

  Row<NV> GH, GT, GR, SPH, SPT;
  GH.zero(); GT.zero(); GR.zero(); SPH.zero(); SPT.zero();
  float nrm[KGE_MAX_TABLES] = {0.f, 0.f, 0.f, 0.f};   // per-lane partials (vector tables)
  float matn[2] = {0.f, 0.f};                         // warp-uniform (matrix tables)
  float lsum = 0.f, gpos = 0.f;
  const int64_t nE = 2 * p.B + p.B * (int64_t) k;
  float* recE = p.rec[KGE_T_ENT];
  const int ldE = md.rld[KGE_T_ENT];

  // backward of one scored triple (A, r, B) with d loss / d score = g
  auto triple_backward = [&](float g, bool is_pos, bool head, const Row<NV>& N, const Row<NV>& CN, float nN,
                             int64_t rec_slot) {
    if constexpr (M::BILINEAR) {
      if (is_pos) {
        Row<NV> a; rscale(a, g, CT); row_add(GH, a); nrm[KGE_T_ENT] += row_dot_partial(a, a);
        rscale(a, g, CH); row_add(GT, a); nrm[KGE_T_ENT] += row_dot_partial(a, a);
        rscale(a, g, Et); row_add(SPH, a);
        float hh = row_dot_partial(Eh, Eh), tt = row_dot_partial(Et, Et);
        warp_sum2(hh, tt);
        matn[0] += g * g * hh * tt;
      } else {
        Row<NV> gN, Q, gS;
        rscale(gN, g, head ? CT : CH);                  // d s / d N = U (head) or V (tail)
        sstore(gN, recE + rec_slot * (int64_t) ldE, din, lane);
        if (head) matvec_rowin(Q, N, cx.M0, cx.ld0, din, din, cx.scratch, lane);     // R^T N
        else matvec_colin(Q, N, cx.M0, cx.ld0, din, din, cx.scratch, lane);          // R N
        rscale(gS, g, Q);
        nrm[KGE_T_ENT] += row_dot_partial(gN, gN) + row_dot_partial(gS, gS);
        Row<NV> a; rscale(a, g, N);
        if (head) { row_add(GT, gS); row_add(SPT, a); } else { row_add(GH, gS); row_add(SPH, a); }
        float nn = row_dot_partial(N, N), ss = head ? row_dot_partial(Et, Et) : row_dot_partial(Eh, Eh);
        warp_sum2(nn, ss);
        matn[0] += g * g * nn * ss;
      }
    } else {
      Row<NV> PHr, PTr, geH, geT;
      score_bwd<NV, false>(md.sc, x, y, sv, g, PHr, PTr);         // gx, gy
      if (M::HAS_REL_VEC) { row_add(GR, PHr); nrm[KGE_T_REL_A] += row_dot_partial(PHr, PHr); }
      const bool hneg = !is_pos && head, tneg = !is_pos && !head;
      M::template proj_bwd<NV>(md, cx, 0, PHr, hneg ? CN : CH, hneg ? nN : nh, geH, lane);
      M::template proj_bwd<NV>(md, cx, 1, PTr, tneg ? CN : CT, tneg ? nN : nt, geT, lane);
      nrm[KGE_T_ENT] += row_dot_partial(geH, geH) + row_dot_partial(geT, geT);
      M::template mat_norm<NV>(hneg ? N : Eh, tneg ? N : Et, PHr, PTr, matn);
      if (hneg) {
        sstore(geH, recE + rec_slot * (int64_t) ldE, din, lane);
        sstore(PHr, q.recP + rec_slot * (int64_t) q.ldp, dout, lane);
      } else { row_add(GH, geH); row_add(SPH, PHr); }
      if (tneg) {
        sstore(geT, recE + rec_slot * (int64_t) ldE, din, lane);
        sstore(PTr, q.recP + rec_slot * (int64_t) q.ldp, dout, lane);
      } else { row_add(GT, geT); row_add(SPT, PTr); }
    }
  };

  for (int c = warp; c < k; c += STEP_WARPS) {
    const bool head = neg_is_head(p.sd, c);
    const int64_t slot = i * k + c;
    const int64_t e = p.neg_override ? p.neg_override[slot] : sample_negative(p.sd, i, c, h, t);
    Row<NV> N, CN; float nN;
    sload(N, md.t[KGE_T_ENT].w + e * (int64_t) md.t[KGE_T_ENT].ld, din, lane);
    if (lane == 0 && p.write_keys) {
      p.neg_ids[slot] = e;
      if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) e; p.vals[2 * p.B + slot] = (uint32_t)(2 * p.B + slot); }
    }
    const float sn = neg_forward(head, N, CN, nN);
    const float w = softmax ? expf(p.loss.temperature * sn - sm_max) * sm_inv : 1.f;
    float term, dneg, dpp;
    loss_neg(p.loss, sp, sn, w, term, dneg, dpp);
    lsum += term;
    gpos += dpp;
    if (TRAIN) triple_backward(dneg, false, head, N, CN, nN, 2 * p.B + slot);
  }

  if (lane == 0) s_red[warp * 8] = gpos;
  __syncthreads();
  if (warp == 0) {
    float gp = 0.f;
    for (int w2 = 0; w2 < STEP_WARPS; ++w2) gp += s_red[w2 * 8];
    float term, dpos;
    loss_pos(p.loss, sp, term, dpos);
    lsum += term;
    gp += dpos;
    if (TRAIN) {
      if constexpr (!M::BILINEAR) score_of(CH, CT);
      triple_backward(gp, true, false, Eh, CH, 0.f, 0);
    }
  }
  __syncthreads();

  // ---- cross-warp reduction (fixed order) ---------------------------------------------
  float red[8];
  red[0] = lsum;
#pragma unroll
  for (int u = 0; u < KGE_MAX_TABLES; ++u) red[1 + u] = TRAIN ? warp_sum(nrm[u]) : 0.f;
  red[1 + KGE_T_REL_A] += (M::KIND == KGE_TRANSR) ? 0.f : matn[0];
  red[1 + KGE_T_REL_B] += matn[1];
  if (lane == 0) {
#pragma unroll
    for (int u = 0; u < N_PART; ++u) s_red[warp * 8 + u] = red[u];
  }
  if (TRAIN) {
    float4* mine = s_acc + (size_t) warp * MAT_ACC_ROWS * NV * WARP;
    const Row<NV>* rows[MAT_ACC_ROWS] = {&GH, &GT, &GR, &SPH, &SPT};
#pragma unroll
    for (int a = 0; a < MAT_ACC_ROWS; ++a)
#pragma unroll
      for (int u = 0; u < NV; ++u) mine[(a * NV + u) * WARP + lane] = rows[a]->v[u];
  }
  __syncthreads();
  if (warp == 0) {
    if (lane < N_PART) {
      float a = 0.f;
      for (int w2 = 0; w2 < STEP_WARPS; ++w2) a += s_red[w2 * 8 + lane];
      p.part[(size_t) lane * p.B + i] = a;
    }
    if (p.tail && last_warp_done(p.done, (unsigned) p.B, lane)) tail_reduce_ctas(p.rp, (int) p.B, lane);
  }
  if (TRAIN && warp < MAT_ACC_ROWS) {
    Row<NV> acc;
    acc.zero();
    for (int w2 = 0; w2 < STEP_WARPS; ++w2) {
      const float4* src = s_acc + ((size_t) w2 * MAT_ACC_ROWS + warp) * NV * WARP;
#pragma unroll
      for (int u = 0; u < NV; ++u) acc.v[u] += src[u * WARP + lane];
    }
    switch (warp) {
      case 0: sstore(acc, recE + i * (int64_t) ldE, din, lane); break;
      case 1: sstore(acc, recE + (p.B + i) * (int64_t) ldE, din, lane); break;
      case 2: if (M::HAS_REL_VEC) sstore(acc, p.rec[KGE_T_REL_A] + i * (int64_t) md.rld[KGE_T_REL_A], dout, lane); break;
      case 3: sstore(acc, q.recP + i * (int64_t) q.ldp, dout, lane); break;
      default: sstore(acc, q.recP + (p.B + i) * (int64_t) q.ldp, dout, lane);
    }
  }
  if (TRAIN && threadIdx.x == 0 && p.write_keys) {
    p.keys[i] = (uint32_t) h;           p.vals[i] = (uint32_t) i;
    p.keys[p.B + i] = (uint32_t) t;     p.vals[p.B + i] = (uint32_t)(p.B + i);
    p.keys[nE + i] = (uint32_t)(md.t[KGE_T_ENT].rows + r);
    p.vals[nE + i] = (uint32_t)(nE + i);
  }
}

}  // namespace kge
