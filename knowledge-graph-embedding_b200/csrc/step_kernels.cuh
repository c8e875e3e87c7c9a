// step_kernels.cuh — the fused per-batch training kernels (vector-family models).
//
//   k_step   : one warp group per positive: Philox corruption, gather of h/r/t rows
//              (128-bit loads, rows in registers), model projection, score, loss,
//              analytic backward.  Emits gradient *records* (one row per negative
//              entity, plus per-positive pre-reduced rows for the shared h, t and
//              relation rows), sort keys, and per-positive loss / norm^2 partials.
//   k_reduce : fixed-order reduction of the partials -> loss, per-table clip scales.
//   k_update : one warp per table row: sum the row's records in sorted (stable)
//              order, apply clip scale and the TF-semantics Adam update (dense
//              decay of m, v and move of w for every row).
// No atomics anywhere: results are bitwise reproducible run to run.
#pragma once
#include "models.cuh"
#include "loss.cuh"
#include "reduce_tail.cuh"
#include "tc_ptx.cuh"

namespace kge {

constexpr int STEP_THREADS = 256;
constexpr int STEP_WARPS = STEP_THREADS / WARP;
constexpr int STEP_MAX_WARPS = 14;           // DistMult (NV <= 2): the CTA size is picked per launch (StepParams.nwarps)
// launch bounds of k_step<M, NV>: threads per CTA the kernel may be launched with, CTAs per SM it is built for
template <class M, int NV> struct StepBounds {
  static constexpr bool WIDE = (M::KIND == KGE_DISTMULT && NV <= 2);
  static constexpr int THREADS = WIDE ? STEP_MAX_WARPS * WARP : STEP_THREADS;
  static constexpr int MIN_CTAS = WIDE ? 1 : (M::KIND == KGE_TRANSD ? (NV == 1 ? 2 : 1) : (NV == 1 ? 3 : (NV == 2 ? 2 : 1)));
};
constexpr int N_PART = 1 + KGE_MAX_TABLES;   // loss + norm^2 per table

struct StepParams {
  ModelDev md;
  LossDev loss;
  SamplerDev sd;
  const void* triples;
  int id64;
  int64_t B;
  int k_eff;
  int wpp;                       // warps per positive: 1, 2, 4 or 8
  int nwarps;                    // warps per CTA (= STEP_WARPS except where StepBounds::WIDE lets the host pick)
  const int64_t* neg_override;   // nullable
  int64_t* neg_ids;              // [B*k_eff] negatives used (always written)
  float* rec[KGE_MAX_TABLES];    // gradient records per table slot
  uint32_t* keys;                // [2B + B*k_eff + B]
  uint32_t* vals;                // iota over the same range
  float* part;                   // [N_PART][B]
  int write_keys;                // 0: sort keys / negatives are produced by k_keys on the side stream
  int tail;                      // 1: the last warp to finish reduces the partials (no k_reduce launch)
  unsigned* done;                // ticket counter for the tail reduction (self-resetting)
  ReduceParams rp;
  // sharded entity tables (kge_shard): an entity-table record that is the ONLY record of its row is the
  // row's summed gradient, so the step kernel stores it straight into the owner's value inbox over
  // NVLink (the outbound stores overlap the inbound row gathers) instead of into the local record
  // buffer.  rec_dest[record] = owner << 29 | inbox row, or REC_LOCAL; inbox[s][owner] = this rank's
  // section of owner's value inbox of entity slot s (row stride = the record stride).
  const uint32_t* rec_dest;      // nullable
  float* inbox[2][KGE_MAX_PEERS];
  // Factored records (DistMult fast loop, dense single-GPU / data-parallel step): the gradient of a
  // negative's entity row is dneg * q with q = r o t or h o r fixed per (positive, corrupted side), so the
  // step kernel publishes q once (fac_q[(i * n_side + side) * rld]) and ONE float per negative
  // (fac_coef[i * k + c] = dneg) instead of a full record row; the update kernels rebuild
  // fl(dneg * q) element by element (update_kernels.cuh, RecSrc) — the bits the stored row would have had.
  float* fac_coef;               // nullable = records are materialised
  float* fac_q;
  // Negative rows staged through shared memory by TMA bulk copies (DistMult loop): every warp owns a ring of
  // ring_slots rows (multiple of 4; 0 = rows are loaded straight into registers) behind the step kernel's
  // other shared memory, at byte offset ring_off; one mbarrier per group of four rows.
  int ring_slots;
  uint32_t ring_off;
};
// one row global -> shared by the TMA unit (cp.async.bulk), completion counted on `bar`
__device__ __forceinline__ void bulk_row_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   tc::smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(tc::smem_u32(bar)) : "memory");
}
constexpr uint32_t REC_LOCAL = 0xffffffffu;
// where entity record `rec` goes: the local record buffers (row = rec) or this rank's section of the
// owner's value inbox
struct RecTarget { float* b0; float* b1; int64_t row; };
__device__ __forceinline__ RecTarget rec_target(const StepParams& p, int64_t rec) {
  RecTarget t;
  t.b0 = p.rec[KGE_T_ENT]; t.b1 = p.rec[KGE_T_ENT_PROJ]; t.row = rec;
  if (p.rec_dest != nullptr) {
    const uint32_t d = __ldg(p.rec_dest + rec);
    if (d != REC_LOCAL) {
      const int o = (int) (d >> 29);
      t.b0 = p.inbox[0][o]; t.b1 = p.inbox[1][o];
      t.row = (int64_t) (d & 0x1fffffffu);
    }
  }
  return t;
}

// ticket handshake: the caller's stores are fenced; returns true in the last warp to arrive
__device__ __forceinline__ bool last_warp_done(unsigned* done, unsigned total, int lane) {
  __threadfence();
  unsigned t = 0;
  if (lane == 0) t = atomicAdd(done, 1u);
  t = __shfl_sync(FULL, t, 0);
  if (t != total - 1) return false;
  if (lane == 0) *done = 0u;
  __threadfence();
  return true;
}

__device__ __forceinline__ void group_barrier(int grp, int wpp) {
  if (wpp == 1) { __syncwarp(); return; }
  asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(wpp * WARP) : "memory");
}

template <class M, int NV, bool TRAIN>
__global__ void __launch_bounds__(StepBounds<M, NV>::THREADS, StepBounds<M, NV>::MIN_CTAS) k_step(const StepParams p) {
  using Ent = typename M::template Ent<NV>;
  using Rel = typename M::template Rel<NV>;
  extern __shared__ float4 smem4[];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nw = StepBounds<M, NV>::WIDE ? (int) (blockDim.x >> 5) : STEP_WARPS;
  const int wpp = p.wpp, ppc = nw / wpp;
  const int grp = warp / wpp, wig = warp % wpp;
  const int64_t i = (int64_t) blockIdx.x * ppc + grp;
  if (i >= p.B) return;                       // whole group leaves together
  const ModelDev& md = p.md;
  const int k = p.k_eff;

  // shared memory carve-up (per CTA): scores [ppc][k], scalars [8][8], accumulators
  constexpr int ACC_F4 = (2 * Ent::NROWS + Rel::NROWS) * NV;   // float4 per lane
  float* s_neg = reinterpret_cast<float*>(smem4) + (size_t) grp * k;
  float* s_red = reinterpret_cast<float*>(smem4) + (((size_t) ppc * k + 3) & ~(size_t) 3) + warp * 8;
  float4* s_acc = smem4 + (((size_t) ppc * k + 3) / 4) + (nw * 8) / 4 + (size_t) warp * ACC_F4 * WARP;

  const int64_t h = load_id(p.triples, 3 * i + 0, p.id64);
  const int64_t r = load_id(p.triples, 3 * i + 1, p.id64);
  const int64_t t = load_id(p.triples, 3 * i + 2, p.id64);

  // Negative ids: lane l draws the id of this warp's (it0 + l)-th negative, so the ~100
  // instruction Philox runs once per 32 negatives per lane instead of once per negative per
  // warp; the next negative's rows are requested before the current one is processed.
  const int n_it = (k - wig + wpp - 1) / wpp;
  int64_t my_e = 0;
  auto draw = [&](int it0) {
    const int c = wig + (it0 + lane) * wpp;
    if (c < k) my_e = p.neg_override ? p.neg_override[i * k + c] : sample_negative(p.sd, i, c, h, t);
  };
  // DistMult: the negatives' rows travel global -> shared by TMA bulk copies into this warp's ring, requested
  // before anything else so that they are in flight while the positive is loaded and scored.  Rows
  // j0 .. j0+3 (j0 % 4 == 0) share one mbarrier; the lane that holds a row's id issues its copy.
  const int RS = (M::KIND == KGE_DISTMULT && md.shard_rows == 0) ? p.ring_slots : 0;
  const int RSM = RS > 0 ? RS : 4, RSG = RSM >> 2;       // ring depth is a power of two: modulo = mask
  const int rs_log = 31 - __clz(RSM);
  const uint32_t row_bytes = (uint32_t) md.rld[KGE_T_ENT] * 4u;
  uint64_t* ring_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(smem4) + p.ring_off) + (size_t) warp * (RS >> 2);
  float* ring = reinterpret_cast<float*>(reinterpret_cast<char*>(smem4) + p.ring_off + (((size_t) nw * (RS >> 2) * 8 + 15) & ~(size_t) 15)) +
                (size_t) warp * RS * md.rld[KGE_T_ENT];
  auto ring_issue = [&](int j0) {
    const int l0 = j0 & 31, nrows = (n_it - j0 < 4) ? n_it - j0 : 4;
    uint64_t* bar = ring_bar + ((j0 >> 2) & (RSG - 1));
    if (lane == l0) tc::mbar_expect_tx(bar, (uint32_t) nrows * row_bytes);
    const int x = lane - l0;
    if (x >= 0 && x < nrows) {
      bulk_row_g2s(ring + (size_t) ((j0 + x) & (RSM - 1)) * md.rld[KGE_T_ENT], md.t[KGE_T_ENT].w + my_e * (int64_t) md.t[KGE_T_ENT].ld,
                   row_bytes, bar);
      if (p.write_keys) {
        const int64_t slot = i * k + wig + (int64_t) (j0 + x) * wpp;
        p.neg_ids[slot] = my_e;
        if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) my_e; p.vals[2 * p.B + slot] = (uint32_t)(2 * p.B + slot); }
      }
    }
  };
  if constexpr (M::KIND == KGE_DISTMULT) {
    if (n_it > 0) draw(0);
    if (RS > 0) {
      if (lane < (RS >> 2)) tc::mbar_init(ring_bar + lane, 1);
      tc::mbar_fence_init();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      for (int j0 = 0; j0 < RS && j0 < n_it; j0 += 4) ring_issue(j0);
    }
  }
  Ent H, T;
  Rel R;
  H.load(md, h, lane);
  T.load(md, t, lane);
  R.load(md, r, lane);

  Row<NV> x, y;
  FwdSaved<NV> sv;
  const float sp = M::template forward<NV>(md, H, R, T, lane, x, y, sv);

  // RotatE with LpDistance(1) (its default, RotatE.py:46): |h o r - N| for tail corruption and
  // |N o r - t| = |N - t o conj(r)| for head corruption (|r| = 1) are both "distance of the
  // negative's row to a vector X fixed per positive".  One X instead of H, T, R and their
  // gradient rows in registers: 1000-float rows no longer spill.
  bool rot_fast = false;
  Row<NV> X;
  if constexpr (M::KIND == KGE_ROTATE) {
    rot_fast = md.sc.kind == SC_L1 && !md.sc.pow2 && (p.sd.side_mode != KGE_SIDE_HT || (wpp & 1) == 0);
    if (rot_fast) {
      const bool head = neg_is_head(p.sd, wig);
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const float4 cs = R.a.v[v];                         // (cos0, cos1, sin0, sin1)
        if (head) {                                          // t o conj(r)
          const float4 e = T.e.v[v];
          X.v[v] = f4(e.x * cs.x + e.y * cs.z, -e.x * cs.z + e.y * cs.x, e.z * cs.y + e.w * cs.w, -e.z * cs.w + e.w * cs.y);
        } else {                                             // h o r
          const float4 e = H.e.v[v];
          X.v[v] = f4(e.x * cs.x - e.y * cs.z, e.x * cs.z + e.y * cs.x, e.z * cs.y - e.w * cs.w, e.z * cs.w + e.w * cs.y);
        }
      }
    }
  }
  // sum of moduli of X - N over the lane's elements (all lanes then reduce)
  auto rot_dist = [&](const Row<NV>& Nr) {
    float a = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const float4 u = X.v[v] - Nr.v[v];
      a += sqrtf(u.x * u.x + u.y * u.y) + sqrtf(u.z * u.z + u.w * u.w);
    }
    return warp_sum(a);
  };

  // self-adversarial weights need all k scores of this positive first (loss.py:176-177)
  float sm_max = 0.f, sm_inv = 1.f;
  const bool softmax = (p.loss.kind == KGE_LOSS_SELF_ADV);
  if (softmax) {
    for (int c = wig; c < k; c += wpp) {
      const bool head = neg_is_head(p.sd, c);
      const int64_t e = p.neg_override ? p.neg_override[i * k + c] : sample_negative(p.sd, i, c, h, t);
      Ent N;
      N.load(md, e, lane);
      float sn;
      if (rot_fast) sn = -fmaxf(rot_dist(N.e), 1e-9f);
      else sn = M::template forward<NV>(md, head ? N : H, R, head ? T : N, lane, x, y, sv);
      if (lane == 0) s_neg[c] = sn;
    }
    group_barrier(grp, wpp);
    float m = -INFINITY;
    for (int c = lane; c < k; c += WARP) m = fmaxf(m, p.loss.temperature * s_neg[c]);
    m = warp_max(m);
    float z = 0.f;
    for (int c = lane; c < k; c += WARP) z += expf(p.loss.temperature * s_neg[c] - m);
    z = warp_sum(z);
    sm_max = m;
    sm_inv = 1.0f / z;
  }

  Ent GH, GT;
  Rel GR;
  GH.zero(); GT.zero(); GR.zero();
  float nrm[KGE_MAX_TABLES] = {0.f, 0.f, 0.f, 0.f};
  float lsum = 0.f, gpos = 0.f;
  const int64_t nE = 2 * p.B + p.B * (int64_t) k;

  constexpr bool PREFETCH = (NV <= 2);       // wider rows: the second row in flight would spill
  // TransE / UM fast loop (below) loads its negatives in groups of GROUP rows instead
  constexpr int GROUP = 4;
  Ent Nnext;
  if (n_it > 0) {
    if constexpr (M::KIND != KGE_DISTMULT) draw(0);
    if (PREFETCH && RS == 0) Nnext.load(md, (int64_t) __shfl_sync(FULL, (uint32_t) my_e, 0), lane);      // in flight while the positive is scored
  }
  // DistMult fast path.  s = sum(e_h * r * e_t) is linear in the corrupted entity, so with
  // q = r*t (head corrupted) or h*r (tail corrupted) fixed per positive:  s_neg = N . q,
  // d s/d N = q, and the shared rows' gradients are (sum_neg dneg * N) * {r, kept entity}: one
  // accumulator A instead of three gradient rows per negative.  Per-occurrence norm^2 (what
  // tf.clip_by_norm sums over the un-merged IndexedSlices) stays exact per negative:
  // dneg^2 * (|q|^2 + sum N^2 r^2) for the entity table, dneg^2 * sum N^2 kept^2 for the relation.
  // Needs the corrupted side to be fixed per warp: true unless 'h+t' meets an odd warp count.
  bool fast_done = false;
  if constexpr (M::KIND == KGE_ROTATE) {
    if (rot_fast) {
      fast_done = true;
      const bool head = neg_is_head(p.sd, wig);
      Row<NV> A;                     // sum over negatives of coef * unit(X - N)
      Row<NV> Gp;                    // phase gradient, (g0, g1, 0, 0) per chunk like RelRot's gradient row
      A.zero(); Gp.zero();
      for (int it = 0; it < n_it; ++it) {
        const int c = wig + it * wpp;
        const int64_t slot = i * k + c;
        const int64_t e = (int64_t) __shfl_sync(FULL, (uint32_t) my_e, it & 31);
        Ent N;
        N.load(md, e, lane);
        if (it + 1 < n_it && ((it + 1) & 31) == 0) draw(it + 1);
        if (lane == 0 && p.write_keys) {
          p.neg_ids[slot] = e;
          if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) e; p.vals[2 * p.B + slot] = (uint32_t)(2 * p.B + slot); }
        }
        // unit vectors of X - N in place of N, sum of moduli, count of non-zero moduli
        float acc = 0.f, cnt = 0.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          const float4 u = X.v[v] - N.e.v[v];
          const float m0 = sqrtf(u.x * u.x + u.y * u.y), m1 = sqrtf(u.z * u.z + u.w * u.w);
          const float i0 = m0 > 0.f ? 1.0f / m0 : 0.f, i1 = m1 > 0.f ? 1.0f / m1 : 0.f;
          acc += m0 + m1;
          cnt += (m0 > 0.f ? 1.f : 0.f) + (m1 > 0.f ? 1.f : 0.f);
          N.e.v[v] = f4(u.x * i0, u.y * i0, u.z * i1, u.w * i1);
        }
        const float S = warp_sum(acc);
        const float sn = -fmaxf(S, 1e-9f);
        const float w = softmax ? expf(p.loss.temperature * sn - sm_max) * sm_inv : 1.f;
        float term, dneg, dpp;
        loss_neg(p.loss, sp, sn, w, term, dneg, dpp);
        lsum += term;
        gpos += dpp;
        if (TRAIN) {
          // d score / d X = -unit(X - N) inside the clip's pass band (score.py:61)
          const float coef = (S >= 1e-9f) ? -dneg : 0.f;
          const float cps = (head ? -coef : coef) * md.rotate_scale;
          float pr = 0.f;
#pragma unroll
          for (int v = 0; v < NV; ++v) {
            const float4 uh = N.e.v[v], xv = X.v[v];
            A.v[v] += coef * uh;
            // phase gradient: scale * Im(conj(X) * coef * unit)
            // (head corruption: unit(N o r - t) = -r o unit(X - N), which flips the sign)
            const float g0 = (xv.x * uh.y - xv.y * uh.x) * cps;
            const float g1 = (xv.z * uh.w - xv.w * uh.z) * cps;
            Gp.v[v].x += g0; Gp.v[v].y += g1;
            pr += g0 * g0 + g1 * g1;
            N.e.v[v] = (-coef) * uh;            // d score / d N, up to the rotation for head corruption (same norm)
          }
          nrm[KGE_T_ENT] += 2.0f * coef * coef * cnt;      // the negative's row and the kept entity's occurrence
          nrm[KGE_T_REL_A] += pr;
          { const RecTarget rt = rec_target(p, 2 * p.B + slot);
            store_row(N.e, rt.b0, rt.row, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane); }
        }
      }
      // the positive's rows were dead during the loop (1000-float rows: 96 registers): fetched again (L1 / L2 hits)
      H.load(md, h, lane);
      T.load(md, t, lane);
      R.load(md, r, lane);
      if (TRAIN) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          const float4 q = A.v[v], cs = R.a.v[v];
          if (head) {       // kept entity is t: d/dt = -(d/dx) with x = N o r, i.e. -rot(A)... A is in the X = t o conj(r) frame
            GT.e.v[v] += f4(q.x * cs.x - q.y * cs.z, q.x * cs.z + q.y * cs.x, q.z * cs.y - q.w * cs.w, q.z * cs.w + q.w * cs.y);
          } else {          // kept entity is h: d/dh = conj(r) o (d/dX)
            GH.e.v[v] += f4(q.x * cs.x + q.y * cs.z, -q.x * cs.z + q.y * cs.x, q.z * cs.y + q.w * cs.w, -q.z * cs.w + q.w * cs.y);
          }
          GR.a.v[v] += Gp.v[v];
        }
      }
    }
  }
  // TransE / UM: the score depends on the negative through X - N only, X = h + r (tail corrupted)
  // or t - r (head corrupted; |N + r - t| = |X - N|), for every LpDistance / LpDistancePow / Dot
  // variant score.cuh knows.  One gradient row per negative (d/dX) feeds the record (= -d/dX), the
  // kept entity and the relation (+-d/dX, same norm each), instead of three rows and three norms.
  if constexpr (M::KIND == KGE_TRANSE || M::KIND == KGE_UM) {
    if (md.sc.kind != SC_DOT && (p.sd.side_mode != KGE_SIDE_HT || (wpp & 1) == 0)) {
      fast_done = true;
      const bool head = neg_is_head(p.sd, wig);
      Row<NV> Xq, A;
      A.zero();
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        if constexpr (M::KIND == KGE_UM) Xq.v[v] = head ? T.e.v[v] : H.e.v[v];
        else Xq.v[v] = head ? T.e.v[v] - R.a.v[v] : H.e.v[v] + R.a.v[v];
      }
      auto body = [&](int it, int64_t e, const Ent& N) {
        const int c = wig + it * wpp;
        const int64_t slot = i * k + c;
        if (lane == 0 && p.write_keys) {
          p.neg_ids[slot] = e;
          if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) e; p.vals[2 * p.B + slot] = (uint32_t)(2 * p.B + slot); }
        }
        const float sn = score_fwd<NV, false>(md.sc, Xq, N.e, sv.sc);
        const float w = softmax ? expf(p.loss.temperature * sn - sm_max) * sm_inv : 1.f;
        float term, dneg, dpp;
        loss_neg(p.loss, sp, sn, w, term, dneg, dpp);
        lsum += term;
        gpos += dpp;
        if (TRAIN) {
          Row<NV> gx, gy;
          score_bwd<NV, false>(md.sc, Xq, N.e, sv.sc, dneg, gx, gy);
          const float ss = row_dot_partial(gx, gx);
          nrm[KGE_T_ENT] += 2.0f * ss;                     // the negative's row and the kept entity's occurrence
          if constexpr (M::KIND == KGE_TRANSE) nrm[KGE_T_REL_A] += ss;
          row_add(A, gx);
          { const RecTarget rt = rec_target(p, 2 * p.B + slot);
            store_row(gy, rt.b0, rt.row, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane); }
        }
      };
      if constexpr (NV == 1) {
        // GROUP rows are requested before the first is consumed: a warp keeps GROUP row loads in
        // flight (HBM latency; several times that when the row lives in a peer GPU's memory,
        // kge_shard).  The per-negative arithmetic and its order are unchanged.
        for (int it0 = 0; it0 < n_it; it0 += GROUP) {
          Ent Nb[GROUP];
          int64_t eb[GROUP];
          if (it0 > 0 && (it0 & 31) == 0) draw(it0);        // GROUP divides 32: a group never straddles a draw
#pragma unroll
          for (int x = 0; x < GROUP; ++x)
            if (it0 + x < n_it) {
              eb[x] = (int64_t) __shfl_sync(FULL, (uint32_t) my_e, (it0 + x) & 31);
              if (it0 + x == 0) Nb[x] = Nnext; else Nb[x].load(md, eb[x], lane);
            }
#pragma unroll
          for (int x = 0; x < GROUP; ++x)
            if (it0 + x < n_it) body(it0 + x, eb[x], Nb[x]);
        }
      } else {
        for (int it = 0; it < n_it; ++it) {
          const int64_t e = (int64_t) __shfl_sync(FULL, (uint32_t) my_e, it & 31);
          Ent N;
          if (PREFETCH) N = Nnext; else N.load(md, e, lane);
          if (it + 1 < n_it) {
            if (((it + 1) & 31) == 0) draw(it + 1);
            if (PREFETCH) Nnext.load(md, (int64_t) __shfl_sync(FULL, (uint32_t) my_e, (it + 1) & 31), lane);
          }
          body(it, e, N);
        }
      }
      if (TRAIN) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          if (head) GT.e.v[v] += A.v[v]; else GH.e.v[v] += A.v[v];
          if constexpr (M::KIND == KGE_TRANSE) GR.a.v[v] += head ? -A.v[v] : A.v[v];
        }
      }
    }
  }
  if constexpr (M::KIND == KGE_DISTMULT) {
    if (p.sd.side_mode != KGE_SIDE_HT || (wpp & 1) == 0) {
      fast_done = true;
      const bool head = neg_is_head(p.sd, wig);
      Row<NV> q, r2, o2, A;
      A.zero();
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const float4 o = head ? T.e.v[v] : H.e.v[v];
        q.v[v] = R.a.v[v] * o;
        r2.v[v] = R.a.v[v] * R.a.v[v];
        o2.v[v] = o * o;
      }
      const float q2 = row_dot_partial(q, q);
      const bool ht = p.sd.side_mode == KGE_SIDE_HT;
      if (TRAIN && p.fac_coef != nullptr && wig < (ht ? 2 : 1))      // one writer per (positive, side)
        store_row(q, p.fac_q, ht ? 2 * i + (wig & 1) : i, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane);
      // Four negatives per pass: their rows are requested together, the four partial dot products are
      // reduced "transposed" (xor 16 / 8 hand half of the values to the partner lane, xor 4 / 2 / 1
      // finish one value per lane: 6 shuffles for 4 sums, the same addition tree as warp_sum, so the
      // scores keep their bits) and lanes 8x .. 8x+7 then hold the score of negative x: the loss term
      // and its derivative are evaluated ONCE per pass for four negatives in different lane groups,
      // not once per negative by every lane.  lsum / gpos are kept per lane group and combined after
      // the loop.
      const int lx = lane >> 3;
      constexpr int G4 = 4;
      // loop-invariant pieces of the index arithmetic
      const int ld4 = md.rld[KGE_T_ENT] >> 2;                         // float4 per staged row
      const float4* ring4 = reinterpret_cast<const float4*>(ring) + lane;
      bool cok[NV];
#pragma unroll
      for (int v = 0; v < NV; ++v) cok[v] = 4 * (lane + v * WARP) < md.t[KGE_T_ENT].cols;
      float2* coef_w = p.fac_coef != nullptr ? reinterpret_cast<float2*>(p.fac_coef) + (i * k + wig) : nullptr;
      const float qrow = __uint_as_float((uint32_t) (ht ? 2 * i + (wig & 1) : i));       // index of this warp's q row (wig parity = side)
      for (int it0 = 0; it0 < n_it; it0 += G4) {
        Ent Nb[G4];
        float d[G4];
        const int nval = n_it - it0;                                   // negatives of this pass: x < nval
        if (RS > 0) {
          tc::mbar_wait(ring_bar + ((it0 >> 2) & (RSG - 1)), (uint32_t) ((it0 >> rs_log) & 1));
          const float4* src = ring4 + (it0 & (RSM - 1)) * ld4;        // a pass never wraps: it0 and the depth are multiples of 4
#pragma unroll
          for (int x = 0; x < G4; ++x) {
#pragma unroll
            for (int v = 0; v < NV; ++v) Nb[x].e.v[v] = (x < nval && cok[v]) ? src[x * ld4 + v * WARP] : f4zero();
          }
        } else {
          if (it0 > 0 && (it0 & 31) == 0) draw(it0);          // G4 divides 32: a pass never straddles a draw
#pragma unroll
          for (int x = 0; x < G4; ++x) {
            if (it0 + x < n_it) {
              const int64_t e = (int64_t) __shfl_sync(FULL, (uint32_t) my_e, (it0 + x) & 31);
              if (it0 + x == 0 && PREFETCH) Nb[x] = Nnext; else Nb[x].load(md, e, lane);
              if (lane == 0 && p.write_keys) {
                const int64_t slot = i * k + wig + (int64_t) (it0 + x) * wpp;
                p.neg_ids[slot] = e;
                if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) e; p.vals[2 * p.B + slot] = (uint32_t)(2 * p.B + slot); }
              }
            } else Nb[x].zero();
          }
        }
#pragma unroll
        for (int x = 0; x < G4; ++x) d[x] = row_dot_partial(Nb[x].e, q);
        const bool b16 = (lane & 16) != 0, b8 = (lane & 8) != 0;
        float k0 = b16 ? d[2] : d[0], k1 = b16 ? d[3] : d[1];
        const float g0 = b16 ? d[0] : d[2], g1 = b16 ? d[1] : d[3];
        k0 += __shfl_xor_sync(FULL, g0, 16);
        k1 += __shfl_xor_sync(FULL, g1, 16);
        float sn = b8 ? k1 : k0;
        sn += __shfl_xor_sync(FULL, b8 ? k0 : k1, 8);
        sn += __shfl_xor_sync(FULL, sn, 4);
        sn += __shfl_xor_sync(FULL, sn, 2);
        sn += __shfl_xor_sync(FULL, sn, 1);
        const float w = softmax ? expf(p.loss.temperature * sn - sm_max) * sm_inv : 1.f;
        float term, dneg, dpp;
        loss_neg(p.loss, sp, sn, w, term, dneg, dpp);
        if (lx >= nval) { term = 0.f; dneg = 0.f; dpp = 0.f; }
        lsum += term;
        gpos += dpp;
        if (TRAIN) {
#pragma unroll
          for (int x = 0; x < G4; ++x) {
            if (x >= nval) break;
            const float dn = __shfl_sync(FULL, dneg, 8 * x);
            float pe = 0.f, pr = 0.f;
#pragma unroll
            for (int v = 0; v < NV; ++v) {
              A.v[v] += dn * Nb[x].e.v[v];
              const float4 n2 = Nb[x].e.v[v] * Nb[x].e.v[v];
              pe += dot4(n2, r2.v[v]);
              pr += dot4(n2, o2.v[v]);
            }
            const float d2 = dn * dn;
            nrm[KGE_T_ENT] += d2 * (q2 + pe);
            nrm[KGE_T_REL_A] += d2 * pr;
            if (p.fac_coef == nullptr) {
              Row<NV> gN;
#pragma unroll
              for (int v = 0; v < NV; ++v) gN.v[v] = dn * q.v[v];
              const RecTarget rt = rec_target(p, 2 * p.B + i * k + wig + (int64_t) (it0 + x) * wpp);
              store_row(gN, rt.b0, rt.row, md.rld[KGE_T_ENT], md.t[KGE_T_ENT].cols, lane);
            }
          }
          if (coef_w != nullptr && (lane & 7) == 0 && lx < nval) coef_w[(it0 + lx) * wpp] = make_float2(dneg, qrow);
        }
        if (RS > 0 && it0 + RS < n_it) {                   // the pass's four slots are free: request rows it0 + RS ..
          __syncwarp();
          if (((it0 + RS) & 31) == 0) draw(it0 + RS);      // every row of the previous 32 has been requested
          ring_issue(it0 + RS);
        }
      }
      // lane groups -> warp-uniform sums (fixed order)
      {
        float ls = 0.f, gs = 0.f;
#pragma unroll
        for (int x = 0; x < G4; ++x) { ls += __shfl_sync(FULL, lsum, 8 * x); gs += __shfl_sync(FULL, gpos, 8 * x); }
        lsum = ls; gpos = gs;
      }
      // the positive's rows were dead during the loop (registers): fetched again (L1 / L2 hits)
      H.load(md, h, lane);
      T.load(md, t, lane);
      R.load(md, r, lane);
      if (TRAIN) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          const float4 ar = A.v[v] * R.a.v[v];
          if (head) { GT.e.v[v] += ar; GR.a.v[v] += A.v[v] * T.e.v[v]; }
          else      { GH.e.v[v] += ar; GR.a.v[v] += A.v[v] * H.e.v[v]; }
        }
      }
    }
  }
  for (int it = 0; !fast_done && it < n_it; ++it) {
    const int c = wig + it * wpp;
    const bool head = neg_is_head(p.sd, c);
    const int64_t slot = i * k + c;
    const int64_t e = (int64_t) __shfl_sync(FULL, (uint32_t) my_e, it & 31);
    Ent N;
    if (PREFETCH) N = Nnext; else N.load(md, e, lane);
    if (it + 1 < n_it) {
      if (((it + 1) & 31) == 0) draw(it + 1);
      if (PREFETCH) Nnext.load(md, (int64_t) __shfl_sync(FULL, (uint32_t) my_e, (it + 1) & 31), lane);
    }
    if (lane == 0 && p.write_keys) {
      p.neg_ids[slot] = e;
      if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) e; p.vals[2 * p.B + slot] = (uint32_t)(2 * p.B + slot); }
    }
    const Ent& A = head ? N : H;
    const Ent& Bq = head ? T : N;
    const float sn = M::template forward<NV>(md, A, R, Bq, lane, x, y, sv);
    const float w = softmax ? expf(p.loss.temperature * sn - sm_max) * sm_inv : 1.f;
    float term, dneg, dpp;
    loss_neg(p.loss, sp, sn, w, term, dneg, dpp);
    lsum += term;
    gpos += dpp;
    if (TRAIN) {
      Ent gA, gB;
      Rel gR;
      M::template backward<NV>(md, A, R, Bq, x, y, sv, dneg, lane, gA, gR, gB);
      gA.sumsq(nrm); gB.sumsq(nrm); gR.sumsq(nrm);   // each lookup row is one IndexedSlices occurrence
      const RecTarget rt = rec_target(p, 2 * p.B + slot);
      if (head) { gA.store_to(md, rt.b0, rt.b1, rt.row, lane); GT.add(gB); }
      else      { gB.store_to(md, rt.b0, rt.b1, rt.row, lane); GH.add(gA); }
      GR.add(gR);
    }
  }

  // d loss / d pos needs every warp's share
  if (lane == 0) s_red[0] = gpos;
  group_barrier(grp, wpp);
  if (wig == 0) {
    float gp = 0.f;
    for (int w2 = 0; w2 < wpp; ++w2) gp += (s_red - warp * 8 + (grp * wpp + w2) * 8)[0];
    float term, dpos;
    loss_pos(p.loss, sp, term, dpos);
    lsum += term;
    gp += dpos;
    if constexpr (M::KIND == KGE_DISTMULT) {
      if (md.constraint) {
        // DistMult.py:164-165: cw * mean_i sum(rel_inter[r_i]^2) — its own lookup occurrence
        const float rr = warp_sum(row_dot_partial(R.a, R.a));
        lsum += md.cw * rr * p.loss.inv_n;
        if (TRAIN) {
          Rel gReg;
          gReg.zero();
          row_axpy(gReg.a, 2.0f * md.cw * p.loss.inv_n, R.a);
          gReg.sumsq(nrm);
          GR.add(gReg);
        }
      }
    }
    if (TRAIN) {
      M::template forward<NV>(md, H, R, T, lane, x, y, sv);
      Ent gA, gB;
      Rel gR;
      M::template backward<NV>(md, H, R, T, x, y, sv, gp, lane, gA, gR, gB);
      gA.sumsq(nrm); gB.sumsq(nrm); gR.sumsq(nrm);
      GH.add(gA); GT.add(gB); GR.add(gR);
    }
  }

  // ---- cross-warp reduction of the per-positive rows and partials (fixed order) ----
  float red[N_PART];
  red[0] = lsum;
#pragma unroll
  for (int q = 0; q < KGE_MAX_TABLES; ++q) red[1 + q] = TRAIN ? warp_sum(nrm[q]) : 0.f;
  if (wpp > 1) {
    group_barrier(grp, wpp);                // s_red[0] reads are done
    if (lane == 0) {
#pragma unroll
      for (int q = 0; q < N_PART; ++q) s_red[q] = red[q];
    }
    if (TRAIN) {
      GH.each([&](Row<NV>& rw, int q) {
#pragma unroll
        for (int v = 0; v < NV; ++v) s_acc[((q) * NV + v) * WARP + lane] = rw.v[v];
      });
      GT.each([&](Row<NV>& rw, int q) {
#pragma unroll
        for (int v = 0; v < NV; ++v) s_acc[((Ent::NROWS + q) * NV + v) * WARP + lane] = rw.v[v];
      });
      GR.each([&](Row<NV>& rw, int q) {
#pragma unroll
        for (int v = 0; v < NV; ++v) s_acc[((2 * Ent::NROWS + q) * NV + v) * WARP + lane] = rw.v[v];
      });
    }
    group_barrier(grp, wpp);
    if (wig != 0) return;
#pragma unroll
    for (int q = 0; q < N_PART; ++q) red[q] = 0.f;
    for (int w2 = 0; w2 < wpp; ++w2) {
      const float* sr = s_red + w2 * 8;
#pragma unroll
      for (int q = 0; q < N_PART; ++q) red[q] += sr[q];
    }
    if (TRAIN) {
      GH.zero(); GT.zero(); GR.zero();
      for (int w2 = 0; w2 < wpp; ++w2) {
        const float4* sa = s_acc + (size_t) w2 * ACC_F4 * WARP;
        GH.each([&](Row<NV>& rw, int q) {
#pragma unroll
          for (int v = 0; v < NV; ++v) rw.v[v] += sa[((q) * NV + v) * WARP + lane];
        });
        GT.each([&](Row<NV>& rw, int q) {
#pragma unroll
          for (int v = 0; v < NV; ++v) rw.v[v] += sa[((Ent::NROWS + q) * NV + v) * WARP + lane];
        });
        GR.each([&](Row<NV>& rw, int q) {
#pragma unroll
          for (int v = 0; v < NV; ++v) rw.v[v] += sa[((2 * Ent::NROWS + q) * NV + v) * WARP + lane];
        });
      }
    }
  }
  if (p.tail) {
    // CTA level: the leader warps of the CTA's positives meet on their own named barrier (the other warps
    // have left), the first one sums their partials in positive order and publishes one value per quantity
    const int left = (int) (p.B - (int64_t) blockIdx.x * ppc), nvalid = left < ppc ? left : ppc;
    float* s_red0 = s_red - warp * 8;
    if (lane == 0) {
#pragma unroll
      for (int q = 0; q < N_PART; ++q) s_red[q] = red[q];
    }
    if (nvalid > 1) asm volatile("bar.sync 15, %0;" ::"r"(nvalid * WARP) : "memory");
    else __syncwarp();
    if (grp == 0) {
      if (lane < N_PART) {
        float a = 0.f;
        for (int g2 = 0; g2 < nvalid; ++g2) a += s_red0[g2 * wpp * 8 + lane];
        p.part[(size_t) lane * gridDim.x + blockIdx.x] = a;
      }
      if (last_warp_done(p.done, gridDim.x, lane)) tail_reduce_ctas(p.rp, (int) gridDim.x, lane);
    }
  } else if (lane == 0) {
#pragma unroll
    for (int q = 0; q < N_PART; ++q) p.part[(size_t) q * p.B + i] = red[q];
  }
  if (TRAIN) {
    { const RecTarget rt = rec_target(p, i); GH.store_to(md, rt.b0, rt.b1, rt.row, lane); }
    { const RecTarget rt = rec_target(p, p.B + i); GT.store_to(md, rt.b0, rt.b1, rt.row, lane); }
    GR.store(md, p.rec, i, lane);
    if (lane == 0 && p.write_keys) {
      p.keys[i] = (uint32_t) h;           p.vals[i] = (uint32_t) i;
      p.keys[p.B + i] = (uint32_t) t;     p.vals[p.B + i] = (uint32_t)(p.B + i);
      p.keys[nE + i] = (uint32_t)(md.t[KGE_T_ENT].rows + r);
      p.vals[nE + i] = (uint32_t)(nE + i);
    }
  }
}

}  // namespace kge
