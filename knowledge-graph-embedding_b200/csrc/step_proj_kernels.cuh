// step_proj_kernels.cuh — fused training step of the projection models with one matrix per relation (TransR,
// SE): the phase-structured successor of k_step_mat for these two models (TransR.py:178-191, SE.py:155-161).
//
// k_step_mat gave every negative to one warp, which ran 1 forward + 2 backward mat-vecs of ONE vector each: 5
// shared-memory reads per 4 FMAs, and the kernel sat on the LDS pipe (C4 TransR: 188 us per step, 34 mat-vecs per
// positive).  Here all vectors of a positive that meet the same matrix in the same orientation go through it FOUR at
// a time (matvec_*4: 4 matrix reads + 1 broadcast read per 16 FMAs), so a positive with 10 negatives costs 9 units
// instead of 34 mat-vecs:
//   P0  stage the relation's matrix (odd row stride) and the k + 2 entity rows in shared memory; negative ids
//   P1  forward units: projections C_j (and clip norms) of 4 entity rows per warp                  -> s_C, s_n
//   P2  scores of the k + 1 triples, one warp each                                                 -> s_score
//   P3  loss terms, lane-parallel over the negatives (every warp, redundantly; self-adversarial weights)
//   P4  per triple, one warp: score backward -> (gx, gy); relation-vector gradient; clip backward; matrix-norm
//       term; pair vectors P' -> s_P (negatives' also straight into their records)
//   P5  backward units: entity gradients of 4 pair vectors per warp; negatives' rows into their records, the
//       positive's accumulate in registers
//   P6  cross-warp reduction in fixed order, per-positive records and partials, in-kernel tail (as k_step_mat)
// Every projected element is the same ascending fmaf chain as matvec_rowin / matvec_colin, so scores equal
// kge_score's and the ranking kernels' bit for bit; sums over negatives are taken in a fixed (unit, warp) order.
#pragma once
#include "step_mat_kernels.cuh"

namespace kge {

constexpr int PJ_R = 4;             // vectors per warp through one mat-vec
constexpr int PJ_MAX_ROWS = 64;     // k + 2 <= PJ_MAX_ROWS (row lists live in shared memory)

__device__ __forceinline__ float f4get(const float4& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }

// acc[r][t] = sum_{i < n_in} vec_r[i] * M[i*ld + col_t],  col_t = lane + 32 t   (matvec_rowin, PJ_R vectors)
template <int NV>
__device__ __forceinline__ void matvec_rowin4(float (&acc)[PJ_R][4 * NV], const float* const (&vec)[PJ_R], const float* M,
                                              int ld, int n_in, int n_out, int lane) {
  int col[4 * NV];
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) {
    col[t] = min(lane + 32 * t, n_out - 1);
#pragma unroll
    for (int r = 0; r < PJ_R; ++r) acc[r][t] = 0.f;
  }
  const int n4 = n_in & ~3;
  for (int i = 0; i < n4; i += 4) {
    float4 v[PJ_R];
#pragma unroll
    for (int r = 0; r < PJ_R; ++r) v[r] = *reinterpret_cast<const float4*>(vec[r] + i);
#pragma unroll
    for (int ii = 0; ii < 4; ++ii) {
      const float* Mi = M + (size_t) (i + ii) * ld;
      float m[4 * NV];
#pragma unroll
      for (int t = 0; t < 4 * NV; ++t) m[t] = Mi[col[t]];
#pragma unroll
      for (int r = 0; r < PJ_R; ++r) {
        const float vv = f4get(v[r], ii);
#pragma unroll
        for (int t = 0; t < 4 * NV; ++t) acc[r][t] = fmaf(vv, m[t], acc[r][t]);
      }
    }
  }
  for (int i = n4; i < n_in; ++i) {
    const float* Mi = M + (size_t) i * ld;
#pragma unroll
    for (int r = 0; r < PJ_R; ++r) {
      const float vv = vec[r][i];
#pragma unroll
      for (int t = 0; t < 4 * NV; ++t) acc[r][t] = fmaf(vv, Mi[col[t]], acc[r][t]);
    }
  }
}
// acc[r][t] = sum_{j < n_in} M[row_t*ld + j] * vec_r[j],  row_t = lane + 32 t   (matvec_colin, PJ_R vectors)
template <int NV>
__device__ __forceinline__ void matvec_colin4(float (&acc)[PJ_R][4 * NV], const float* const (&vec)[PJ_R], const float* M,
                                              int ld, int n_out, int n_in, int lane) {
  const float* row[4 * NV];
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) {
    row[t] = M + (size_t) min(lane + 32 * t, n_out - 1) * ld;
#pragma unroll
    for (int r = 0; r < PJ_R; ++r) acc[r][t] = 0.f;
  }
  const int n4 = n_in & ~3;
  for (int j = 0; j < n4; j += 4) {
    float4 v[PJ_R];
#pragma unroll
    for (int r = 0; r < PJ_R; ++r) v[r] = *reinterpret_cast<const float4*>(vec[r] + j);
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      float m[4 * NV];
#pragma unroll
      for (int t = 0; t < 4 * NV; ++t) m[t] = row[t][j + jj];
#pragma unroll
      for (int r = 0; r < PJ_R; ++r) {
        const float vv = f4get(v[r], jj);
#pragma unroll
        for (int t = 0; t < 4 * NV; ++t) acc[r][t] = fmaf(m[t], vv, acc[r][t]);
      }
    }
  }
  for (int j = n4; j < n_in; ++j) {
#pragma unroll
    for (int r = 0; r < PJ_R; ++r) {
      const float vv = vec[r][j];
#pragma unroll
      for (int t = 0; t < 4 * NV; ++t) acc[r][t] = fmaf(row[t][j], vv, acc[r][t]);
    }
  }
}
template <int NV>
__device__ __forceinline__ void acc_to_row(Row<NV>& o, const float (&a)[4 * NV], int n_out, int lane) {
#pragma unroll
  for (int t = 0; t < 4 * NV; ++t) elt(o, t) = (lane + 32 * t < n_out) ? a[t] : 0.f;
}

// shared memory (floats) of k_step_proj, see the carve-up in the kernel
inline size_t step_proj_smem_floats(int nv, int k, int dpad, bool train, int n_mats, int mrows, int mcols) {
  size_t rows = 2 * (size_t) (k + 2) * dpad;             // s_E, s_C
  if (train) {
    rows += 2 * (size_t) (k + 1) * dpad;                 // s_P
    const size_t acc = (size_t) STEP_WARPS * MAT_ACC_ROWS * nv * WARP * 4;   // s_acc lies over the rows once they are dead
    rows = rows > acc ? rows : acc;
  }
  return rows + (size_t) ((k + 2 + 3) & ~3)              // s_n
       + (size_t) ((k + 1 + 3) & ~3)                     // s_score
       + STEP_WARPS * 8 + 2 * PJ_MAX_ROWS                // s_red, row lists (as ints)
       + (size_t) n_mats * mrows * (mcols | 1);
}

template <class M, int NV, bool TRAIN>
__global__ void __launch_bounds__(STEP_THREADS, NV == 1 ? 3 : 1) k_step_proj(const MatStepParams q) {
  static_assert(!M::BILINEAR, "projection models only");
  extern __shared__ float4 smem4[];
  const StepParams& p = q.s;
  const ModelDev& md = p.md;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = blockIdx.x;
  const int k = p.k_eff, nrow = k + 2, ntri = k + 1;
  const int din = M::d_in(md), dout = M::d_out(md), dpad = q.dpad;
  constexpr bool TR = M::KIND == KGE_TRANSR;
  const int s0 = TR ? KGE_T_REL_B : KGE_T_REL_A;

  const int64_t h = load_id(p.triples, 3 * i + 0, p.id64);
  const int64_t r = load_id(p.triples, 3 * i + 1, p.id64);
  const int64_t t = load_id(p.triples, 3 * i + 2, p.id64);

  // ---- shared-memory carve-up (every array starts on a 16-byte boundary) ----------------
  float* sm = reinterpret_cast<float*>(smem4);
  float4* s_acc = reinterpret_cast<float4*>(sm);                                   // P6 only: over the rows of P0 - P5
  {
    size_t rows = 2 * (size_t) nrow * dpad + (TRAIN ? 2 * (size_t) ntri * dpad : 0);
    const size_t acc = TRAIN ? (size_t) STEP_WARPS * MAT_ACC_ROWS * NV * WARP * 4 : 0;
    sm += rows > acc ? rows : acc;
  }
  float* s_E = reinterpret_cast<float*>(smem4);                                    // raw entity rows: 0 = h, 1 = t, 2 + c = negative c
  float* s_C = s_E + (size_t) nrow * dpad;                                         // their projections
  float* s_P = s_C + (size_t) nrow * dpad;                                         // pair vectors [side][triple] (TRAIN)
  float* s_n = sm;                                sm += (nrow + 3) & ~3;           // clip norms (0 = not clipped)
  float* s_score = sm;                            sm += (ntri + 3) & ~3;           // triple 0 = positive, 1 + c = negative c
  float* s_red = sm;                              sm += STEP_WARPS * 8;
  int* s_list = reinterpret_cast<int*>(sm);       sm += 2 * PJ_MAX_ROWS;           // forward row lists: [0] head role, [1] tail role
  float* s_mat = sm;

  const int mrows = md.d_ent, mcols = TR ? md.d_rel : md.d_ent, lds = mcols | 1;
  {
    const float* g0 = md.t[s0].w + r * (int64_t) md.t[s0].ld;
    const float* g1 = M::TWO_MATS ? md.t[KGE_T_REL_B].w + r * (int64_t) md.t[KGE_T_REL_B].ld : nullptr;
    if ((mcols & 3) == 0) {          // 128-bit loads (matrix rows stay 16-byte aligned), one division per thread
      const int c4 = mcols >> 2, n4 = mrows * c4;
      int a = threadIdx.x / c4, b = threadIdx.x - a * c4;
      const int da = STEP_THREADS / c4, db = STEP_THREADS - da * c4;
      for (int e = threadIdx.x; e < n4; e += STEP_THREADS) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(g0) + e);
        float* d = s_mat + a * lds + 4 * b;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        if (M::TWO_MATS) {
          const float4 w = __ldg(reinterpret_cast<const float4*>(g1) + e);
          float* d1 = d + mrows * lds;
          d1[0] = w.x; d1[1] = w.y; d1[2] = w.z; d1[3] = w.w;
        }
        a += da; b += db;
        if (b >= c4) { b -= c4; ++a; }
      }
    } else {
      int a = threadIdx.x / mcols, b = threadIdx.x - a * mcols;
      const int da = STEP_THREADS / mcols, db = STEP_THREADS - da * mcols;
      for (int e = threadIdx.x; e < mrows * mcols; e += STEP_THREADS) {
        s_mat[a * lds + b] = __ldg(g0 + e);
        if (M::TWO_MATS) s_mat[mrows * lds + a * lds + b] = __ldg(g1 + e);
        a += da; b += db;
        if (b >= mcols) { b -= mcols; ++a; }
      }
    }
  }
  const float* M0 = s_mat;
  const float* M1 = M::TWO_MATS ? s_mat + mrows * lds : s_mat;

  // ---- P0: entity rows, negative ids, row lists -------------------------------------------
  for (int j = warp; j < nrow; j += STEP_WARPS) {
    int64_t e;
    if (j == 0) e = h;
    else if (j == 1) e = t;
    else {
      const int c = j - 2;
      const int64_t slot = i * k + c;
      e = p.neg_override ? p.neg_override[slot] : sample_negative(p.sd, i, c, h, t);
      if (lane == 0 && p.write_keys) {
        p.neg_ids[slot] = e;
        if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) e; p.vals[2 * p.B + slot] = (uint32_t) (2 * p.B + slot); }
      }
    }
    const float* src = md.t[KGE_T_ENT].w + e * (int64_t) md.t[KGE_T_ENT].ld;
    float* dst = s_E + (size_t) j * dpad;
    for (int x = lane; x < dpad; x += WARP) dst[x] = (x < din) ? src[x] : 0.f;
  }
  if (threadIdx.x == 0) {           // forward lists: one list for TransR (one matrix, one orientation), two roles for SE
    int na = 0, nb = 0;
    for (int j = 0; j < nrow; ++j) {
      const bool head_role = !M::TWO_MATS || j == 0 || (j >= 2 && neg_is_head(p.sd, j - 2));
      if (head_role) s_list[na++] = j; else s_list[PJ_MAX_ROWS + nb++] = j;
    }
    s_list[PJ_MAX_ROWS - 1] = na;              // k + 2 < PJ_MAX_ROWS: the last slots hold the counts
    s_list[2 * PJ_MAX_ROWS - 1] = nb;
  }
  Row<NV> Rv;
  if (M::HAS_REL_VEC) sload(Rv, md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld, dout, lane);
  else Rv.zero();
  __syncthreads();

  // ---- P1: forward units ---------------------------------------------------------------------
  {
    const int na = s_list[PJ_MAX_ROWS - 1], nb = s_list[2 * PJ_MAX_ROWS - 1];
    const int ua = (na + PJ_R - 1) / PJ_R, ub = (nb + PJ_R - 1) / PJ_R;
    for (int u = warp; u < ua + ub; u += STEP_WARPS) {
      const bool second = u >= ua;
      const int* list = s_list + (second ? PJ_MAX_ROWS : 0);
      const int first = (second ? u - ua : u) * PJ_R, cnt = min(PJ_R, (second ? nb : na) - first);
      int rows[PJ_R];
      const float* vec[PJ_R];
#pragma unroll
      for (int x = 0; x < PJ_R; ++x) { rows[x] = list[first + min(x, cnt - 1)]; vec[x] = s_E + (size_t) rows[x] * dpad; }
      float acc[PJ_R][4 * NV];
      if constexpr (TR) matvec_rowin4<NV>(acc, vec, M0, lds, din, dout, lane);
      else matvec_colin4<NV>(acc, vec, second ? M1 : M0, lds, dout, din, lane);
#pragma unroll
      for (int x = 0; x < PJ_R; ++x) {
        if (x >= cnt) break;
        Row<NV> c;
        acc_to_row<NV>(c, acc[x], dout, lane);
        const float n = TR ? clip1(c, md.constraint != 0) : 0.f;
        float* dst = s_C + (size_t) rows[x] * dpad;
        sstore(c, dst, dout, lane);
        if (lane == 0) s_n[rows[x]] = n;
      }
    }
  }
  __syncthreads();

  // triple j: head / tail row indices (0 = positive; 1 + c = negative c)
  auto tri_rows = [&](int j, int& hj, int& tj, bool& hneg, bool& tneg) {
    hneg = tneg = false; hj = 0; tj = 1;
    if (j > 0) { if (neg_is_head(p.sd, j - 1)) { hneg = true; hj = 1 + j; } else { tneg = true; tj = 1 + j; } }
  };
  auto tri_xy = [&](int hj, int tj, Row<NV>& x, Row<NV>& y) {
    Row<NV> ch;
    sload(ch, s_C + (size_t) hj * dpad, dout, lane);
    sload(y, s_C + (size_t) tj * dpad, dout, lane);
#pragma unroll
    for (int u = 0; u < NV; ++u) x.v[u] = ch.v[u] + Rv.v[u];
  };

  // ---- P2: scores ------------------------------------------------------------------------------
  for (int j = warp; j < ntri; j += STEP_WARPS) {
    int hj, tj; bool hneg, tneg;
    tri_rows(j, hj, tj, hneg, tneg);
    Row<NV> x, y;
    ScoreSaved sv;
    tri_xy(hj, tj, x, y);
    const float s = score_fwd<NV, false>(md.sc, x, y, sv);
    if (lane == 0) s_score[j] = s;
  }
  __syncthreads();

  // ---- P3: loss terms, lane-parallel over the negatives (every warp computes the same numbers) ---
  const float sp = s_score[0];
  const bool softmax = (p.loss.kind == KGE_LOSS_SELF_ADV);
  float sm_max = 0.f, sm_inv = 1.f;
  if (softmax) {
    float m = -INFINITY;
    for (int c = lane; c < k; c += WARP) m = fmaxf(m, p.loss.temperature * s_score[1 + c]);
    m = warp_max(m);
    float z = 0.f;
    for (int c = lane; c < k; c += WARP) z += expf(p.loss.temperature * s_score[1 + c] - m);
    z = warp_sum(z);
    sm_max = m; sm_inv = 1.0f / z;
  }
  float lsum = 0.f, gpos = 0.f;
  for (int c = lane; c < k; c += WARP) {
    const float sn = s_score[1 + c];
    const float w = softmax ? expf(p.loss.temperature * sn - sm_max) * sm_inv : 1.f;
    float term, dneg, dpp;
    loss_neg(p.loss, sp, sn, w, term, dneg, dpp);
    lsum += term; gpos += dpp;
  }
  lsum = warp_sum(lsum); gpos = warp_sum(gpos);
  {
    float term, dpos;
    loss_pos(p.loss, sp, term, dpos);
    lsum += term; gpos += dpos;
  }

  float nrm[KGE_MAX_TABLES] = {0.f, 0.f, 0.f, 0.f};   // per-lane partials (vector tables)
  float matn[2] = {0.f, 0.f};                         // warp-uniform (matrix tables)
  Row<NV> GH, GT, GR, SPH, SPT;
  GH.zero(); GT.zero(); GR.zero(); SPH.zero(); SPT.zero();
  float* recE = p.rec[KGE_T_ENT];
  const int ldE = md.rld[KGE_T_ENT];
  const int64_t nE = 2 * p.B + p.B * (int64_t) k;

  if constexpr (TRAIN) {
    // ---- P4: per triple: score backward, clip backward, pair vectors ------------------------------
    for (int j = warp; j < ntri; j += STEP_WARPS) {
      int hj, tj; bool hneg, tneg;
      tri_rows(j, hj, tj, hneg, tneg);
      Row<NV> x, y, PHr, PTr;
      ScoreSaved sv;
      tri_xy(hj, tj, x, y);
      score_fwd<NV, false>(md.sc, x, y, sv);
      float g = gpos;
      if (j > 0) {
        const float sn = s_score[j];
        const float w = softmax ? expf(p.loss.temperature * sn - sm_max) * sm_inv : 1.f;
        float term, dpp;
        loss_neg(p.loss, sp, sn, w, term, g, dpp);
      }
      score_bwd<NV, false>(md.sc, x, y, sv, g, PHr, PTr);         // gx, gy
      if (M::HAS_REL_VEC) { row_add(GR, PHr); nrm[KGE_T_REL_A] += row_dot_partial(PHr, PHr); }
      if constexpr (TR) {
        Row<NV> ch;                                              // the stored projection (x - r would round differently)
        sload(ch, s_C + (size_t) hj * dpad, dout, lane);
        clip1_bwd(PHr, ch, s_n[hj]);
        clip1_bwd(PTr, y, s_n[tj]);
      }
      Row<NV> eh, et;
      sload(eh, s_E + (size_t) hj * dpad, din, lane);
      sload(et, s_E + (size_t) tj * dpad, din, lane);
      M::template mat_norm<NV>(eh, et, PHr, PTr, matn);
      float* ph = s_P + (size_t) j * dpad;
      float* ptl = s_P + (size_t) (ntri + j) * dpad;
      Row<NV> a = PHr, b = PTr;
      smask(a, dout, lane); smask(b, dout, lane);
#pragma unroll
      for (int tt = 0; tt < 4 * NV; ++tt) { ph[lane + 32 * tt] = elt(a, tt); ptl[lane + 32 * tt] = elt(b, tt); }
      const int64_t slot = 2 * p.B + i * k + (j - 1);
      if (hneg) sstore(PHr, q.recP + slot * (int64_t) q.ldp, dout, lane); else row_add(SPH, PHr);
      if (tneg) sstore(PTr, q.recP + slot * (int64_t) q.ldp, dout, lane); else row_add(SPT, PTr);
    }
    __syncthreads();

    // ---- P5: backward units: side 0 = head-side vectors, side 1 = tail-side vectors ------------------
    const int us = (ntri + PJ_R - 1) / PJ_R;
    for (int u = warp; u < 2 * us; u += STEP_WARPS) {
      const int side = u >= us ? 1 : 0;
      const int first = (side ? u - us : u) * PJ_R, cnt = min(PJ_R, ntri - first);
      const float* vec[PJ_R];
#pragma unroll
      for (int x = 0; x < PJ_R; ++x) vec[x] = s_P + (size_t) (side * ntri + first + min(x, cnt - 1)) * dpad;
      float acc[PJ_R][4 * NV];
      if constexpr (TR) matvec_colin4<NV>(acc, vec, M0, lds, din, dout, lane);
      else matvec_rowin4<NV>(acc, vec, side ? M1 : M0, lds, dout, din, lane);
#pragma unroll
      for (int x = 0; x < PJ_R; ++x) {
        if (x >= cnt) break;
        const int j = first + x;
        Row<NV> ge;
        acc_to_row<NV>(ge, acc[x], din, lane);
        nrm[KGE_T_ENT] += row_dot_partial(ge, ge);
        const bool is_neg_side = j > 0 && (neg_is_head(p.sd, j - 1) == (side == 0));
        if (is_neg_side) sstore(ge, recE + (2 * p.B + i * k + (j - 1)) * (int64_t) ldE, din, lane);
        else if (side == 0) row_add(GH, ge);
        else row_add(GT, ge);
      }
    }
  }

  // ---- P6: cross-warp reduction (fixed order), records of the positive, partials ------------------------
  float red[8];
  red[0] = (warp == 0) ? lsum : 0.f;
#pragma unroll
  for (int u = 0; u < KGE_MAX_TABLES; ++u) red[1 + u] = TRAIN ? warp_sum(nrm[u]) : 0.f;
  red[1 + KGE_T_REL_A] += TR ? 0.f : matn[0];
  red[1 + KGE_T_REL_B] += matn[1];
  if (lane == 0) {
#pragma unroll
    for (int u = 0; u < N_PART; ++u) s_red[warp * 8 + u] = red[u];
  }
  if (TRAIN) {
    __syncthreads();                 // every warp is through with s_E / s_C / s_P: s_acc takes their place
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
    p.keys[p.B + i] = (uint32_t) t;     p.vals[p.B + i] = (uint32_t) (p.B + i);
    p.keys[nE + i] = (uint32_t) (md.t[KGE_T_ENT].rows + r);
    p.vals[nE + i] = (uint32_t) (nE + i);
  }
}

}  // namespace kge
