// step_narrow_kernels.cuh — training step of TransE / UM for NARROW rows (<= 64 floats: C1 is d = 50) with one
// warp-half per positive.
//
// k_step gives a positive a whole warp: lane l holds floats 4l .. 4l+3 of a row, so at d = 50 only 13 of 32 lanes
// carry data and C1's 4096 positives need 4096 warps — more than the 3552 the GPU holds at 80 registers, i.e. two
// waves of a latency-bound kernel (26 us for 8192 scored triples).  Here a warp runs TWO positives, one per 16-lane
// half: the same loads, the same per-element expressions, the same fixed butterfly sums (warp_sum_w<16>: with the
// upper half of a warp empty warp_sum and warp_sum_w<16> return the same bits), half as many warps.
// Covers exactly the cases in which k_step takes its TransE / UM fast path with one warp per positive
// (wpp == 1, one corrupted side, LpDistance / LpDistancePow scores, any loss but the self-adversarial one) and keeps
// k_step's outputs: records (also into a shard owner's inbox: rec_target), keys, per-positive partials, and the
// two-level in-kernel tail in k_step's own order (a CTA of 16 positives publishes the two values k_step's two CTAs
// of 8 would have).
#pragma once
#include "step_kernels.cuh"

namespace kge {

constexpr int NRW_PPC = 2 * STEP_WARPS;      // positives per CTA

template <class M, bool TRAIN>
__global__ void __launch_bounds__(STEP_THREADS, 4) k_step_narrow(const StepParams p) {
  static_assert(M::KIND == KGE_TRANSE || M::KIND == KGE_UM, "TransE / UM only");
  constexpr int HW = 16;
  __shared__ float s_red[NRW_PPC][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane >> 4, hl = lane & (HW - 1);
  const int slot_in_cta = warp * 2 + sub;
  const int64_t i_raw = (int64_t) blockIdx.x * NRW_PPC + slot_in_cta;
  const bool valid = i_raw < p.B;
  const int64_t i = valid ? i_raw : p.B - 1;             // an idle half repeats the last positive, stores suppressed
  const ModelDev& md = p.md;
  const int k = p.k_eff;
  const int cols = md.t[KGE_T_ENT].cols;
  const bool mine = 4 * hl < cols;                       // this lane holds a float4 of the row

  const int64_t h = load_id(p.triples, 3 * i + 0, p.id64);
  const int64_t r = load_id(p.triples, 3 * i + 1, p.id64);
  const int64_t t = load_id(p.triples, 3 * i + 2, p.id64);

  auto ent_row = [&](int64_t id) -> float4 {
    const float* b = ent_base(md, KGE_T_ENT, id);
    return mine ? __ldg(reinterpret_cast<const float4*>(b + id * (int64_t) md.t[KGE_T_ENT].ld) + hl) : f4zero();
  };
  Row<1> H, T, R;
  H.v[0] = ent_row(h);
  T.v[0] = ent_row(t);
  if constexpr (M::KIND == KGE_TRANSE)
    R.v[0] = (4 * hl < md.t[KGE_T_REL_A].cols) ? __ldg(reinterpret_cast<const float4*>(md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld) + hl) : f4zero();
  else R.zero();

  // negative ids: lane hl of the half draws the id of negative it0 + hl
  int64_t my_e = 0;
  auto draw = [&](int it0) {
    const int c = it0 + hl;
    if (c < k) my_e = p.neg_override ? p.neg_override[i * k + c] : sample_negative(p.sd, i, c, h, t);
  };
  if (k > 0) draw(0);

  Row<1> x, y;
  ScoreSaved svp;
  if constexpr (M::KIND == KGE_TRANSE) x.v[0] = H.v[0] + R.v[0]; else x.v[0] = H.v[0];
  y.v[0] = T.v[0];
  const float sp = score_fwd<1, false, HW>(md.sc, x, y, svp);

  const bool head = neg_is_head(p.sd, 0);                // one corrupted side per call (host condition)
  Row<1> Xq, A;
  A.zero();
  if constexpr (M::KIND == KGE_UM) Xq.v[0] = head ? T.v[0] : H.v[0];
  else Xq.v[0] = head ? T.v[0] - R.v[0] : H.v[0] + R.v[0];

  float nrm[KGE_MAX_TABLES] = {0.f, 0.f, 0.f, 0.f};
  float lsum = 0.f, gpos = 0.f;
  const int64_t nE = 2 * p.B + p.B * (int64_t) k;
  const int rld = md.rld[KGE_T_ENT];

  for (int it = 0; it < k; ++it) {
    if (it > 0 && (it & (HW - 1)) == 0) draw(it);
    const int64_t e = (int64_t) __shfl_sync(FULL, (uint32_t) my_e, it & (HW - 1), HW);
    const int64_t slot = i * k + it;
    Row<1> N;
    N.v[0] = ent_row(e);
    if (hl == 0 && valid && p.write_keys) {
      p.neg_ids[slot] = e;
      if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) e; p.vals[2 * p.B + slot] = (uint32_t) (2 * p.B + slot); }
    }
    ScoreSaved sv;
    const float sn = score_fwd<1, false, HW>(md.sc, Xq, N, sv);
    float term, dneg, dpp;
    loss_neg(p.loss, sp, sn, 1.f, term, dneg, dpp);
    lsum += term;
    gpos += dpp;
    if (TRAIN) {
      Row<1> gx, gy;
      score_bwd<1, false>(md.sc, Xq, N, sv, dneg, gx, gy);
      const float ss = row_dot_partial(gx, gx);
      nrm[KGE_T_ENT] += 2.0f * ss;                       // the negative's row and the kept entity's occurrence
      if constexpr (M::KIND == KGE_TRANSE) nrm[KGE_T_REL_A] += ss;
      row_add(A, gx);
      if (valid && mine) {
        const RecTarget rt = rec_target(p, 2 * p.B + slot);
        reinterpret_cast<float4*>(rt.b0 + rt.row * (int64_t) rld)[hl] = gy.v[0];
      }
    }
  }
  Row<1> GH, GT, GR;
  GH.zero(); GT.zero(); GR.zero();
  if (TRAIN) {
    if (head) GT.v[0] += A.v[0]; else GH.v[0] += A.v[0];
    if constexpr (M::KIND == KGE_TRANSE) GR.v[0] += head ? -A.v[0] : A.v[0];
  }
  {
    float term, dpos;
    float gp = 0.f;
    gp += gpos;                                          // k_step: sum over the positive's warps (one here)
    loss_pos(p.loss, sp, term, dpos);
    lsum += term;
    gp += dpos;
    if (TRAIN) {
      ScoreSaved sv2;
      score_fwd<1, false, HW>(md.sc, x, y, sv2);
      Row<1> gA, gB;
      score_bwd<1, false>(md.sc, x, y, sv2, gp, gA, gB);
      nrm[KGE_T_ENT] += row_dot_partial(gA, gA);
      nrm[KGE_T_ENT] += row_dot_partial(gB, gB);
      if constexpr (M::KIND == KGE_TRANSE) nrm[KGE_T_REL_A] += row_dot_partial(gA, gA);
      row_add(GH, gA); row_add(GT, gB);
      if constexpr (M::KIND == KGE_TRANSE) row_add(GR, gA);
    }
  }
  float red[N_PART];
  red[0] = lsum;
#pragma unroll
  for (int q = 0; q < KGE_MAX_TABLES; ++q) red[1 + q] = TRAIN ? warp_sum_w<HW>(nrm[q]) : 0.f;

  if (TRAIN && valid) {
    if (mine) {
      { const RecTarget rt = rec_target(p, i); reinterpret_cast<float4*>(rt.b0 + rt.row * (int64_t) rld)[hl] = GH.v[0]; }
      { const RecTarget rt = rec_target(p, p.B + i); reinterpret_cast<float4*>(rt.b0 + rt.row * (int64_t) rld)[hl] = GT.v[0]; }
    }
    if constexpr (M::KIND == KGE_TRANSE) {
      if (4 * hl < md.t[KGE_T_REL_A].cols)
        reinterpret_cast<float4*>(p.rec[KGE_T_REL_A] + i * (int64_t) md.rld[KGE_T_REL_A])[hl] = GR.v[0];
    }
    if (hl == 0 && p.write_keys) {
      p.keys[i] = (uint32_t) h;           p.vals[i] = (uint32_t) i;
      p.keys[p.B + i] = (uint32_t) t;     p.vals[p.B + i] = (uint32_t) (p.B + i);
      p.keys[nE + i] = (uint32_t) (md.t[KGE_T_ENT].rows + r);
      p.vals[nE + i] = (uint32_t) (nE + i);
    }
  }
  if (p.tail) {
    // k_step's two levels: its CTA b sums positives 8b .. 8b+7 in order and publishes part[q][b]; this CTA stands for
    // its CTAs 2 blockIdx.x and 2 blockIdx.x + 1; the last CTA adds the ceil(B / 8) values (tail_reduce_ctas)
    if (hl == 0) {
#pragma unroll
      for (int q = 0; q < N_PART; ++q) s_red[slot_in_cta][q] = valid ? red[q] : 0.f;
    }
    __syncthreads();
    if (warp == 0) {
      const int n_old = (int) ((p.B + STEP_WARPS - 1) / STEP_WARPS);
      if (lane < 2 * 8) {
        const int half = lane >> 3, q = lane & 7;
        const int64_t ob = 2 * (int64_t) blockIdx.x + half;          // k_step's CTA index
        if (q < N_PART && ob < n_old) {
          const int64_t left = p.B - ob * STEP_WARPS;
          const int nvalid = left < STEP_WARPS ? (int) left : STEP_WARPS;
          float a = 0.f;
          for (int g2 = 0; g2 < nvalid; ++g2) a += s_red[half * STEP_WARPS + g2][q];
          p.part[(size_t) q * n_old + ob] = a;
        }
      }
      if (last_warp_done(p.done, gridDim.x, lane)) tail_reduce_ctas(p.rp, n_old, lane);
    }
  } else if (hl == 0 && valid) {
#pragma unroll
    for (int q = 0; q < N_PART; ++q) p.part[(size_t) q * p.B + i] = red[q];
  }
}

template <class M>
inline int launch_step_narrow(const StepParams& p, const StepLaunch& L) {
  const unsigned grid = (unsigned) ((p.B + NRW_PPC - 1) / NRW_PPC);
  if (L.train) k_step_narrow<M, true><<<grid, STEP_THREADS, 0, L.stream>>>(p);
  else k_step_narrow<M, false><<<grid, STEP_THREADS, 0, L.stream>>>(p);
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}

}  // namespace kge
