// step_rot_kernels.cuh — training step of RotatE with LpDistance(1) (its default, RotatE.py:46,160-166) for WIDE
// rows (d >= 256 complex elements: C3 is d = 500, 4000-byte rows).
//
// k_step's RotatE path keeps whole rows in registers: the per-positive vector X, the negative's row, the
// accumulator of the kept entity's gradient and the phase gradient, beside H, T, R and the generic positive
// backward — 255 registers and spills at 8 chunks per lane, one 8-warp CTA per SM, and C3's 256 CTAs then run as
// two waves (1.27 ms per step).  Here everything that is elementwise per complex element is STREAMED chunk by
// chunk, the accumulators live in shared memory, and only X stays in registers:
//   * |h o r - N| (tail corrupted) and |N o r - t| = |N - t o conj(r)| (head corrupted) are both the distance of
//     the negative's row to a vector X fixed per (positive, side)            (as k_step's rot_fast path)
//   * per negative: one pass over its row for the sum of moduli S (skipped when the self-adversarial pre-pass
//     already has it), the loss term, then a second pass (L1 hits) that writes the gradient record
//     -coef * unit(X - N) and adds coef * unit to the shared-memory accumulator A and the phase gradient to G
//   * the positive's own backward and the fixed-order cross-warp sums are one more streamed pass by the
//     positive's first warp, straight into the records.
//   * the negatives' rows travel global -> shared memory by TMA bulk copies (cp.async.bulk + mbarrier, a private
//     two-slot ring per warp): row it + 1 is in flight while row it is processed, both passes over a row read
//     shared memory, and a warp no longer sits out one L2 round trip per pass and negative (measured: that
//     wait, not occupancy, was the 9 us a warp spent per negative)
//   * CTAs of up to 14 warps sized so that the batch is ONE balanced wave (C3: 1024 positives x 2 warps = 147 CTAs)
// Scores, loss terms and every summation order are k_step's (losses agree bitwise); the negatives' unit vectors
// u / |u| use MUFU.RSQ instead of an IEEE square root and division, so tables agree to rounding (<= 2 ulp per element)
// (tests/test_gpu_step.py::test_rotate_streamed_kernel_equals_register_kernel).
#pragma once
#include "step_kernels.cuh"

namespace kge {

constexpr int ROT_RS = 2;            // ring slots (rows) per warp
inline size_t step_rot_smem_bytes(int nv, int wpp, int k, int nwarps, int ld_floats) {
  const int ppc = nwarps / wpp;
  size_t f4 = ((size_t) ppc * k + 3) / 4 + ((size_t) nwarps * 8) / 4;
  f4 += (size_t) nwarps * nv * WARP;                // A: float4 per lane and chunk
  f4 += (size_t) nwarps * nv * WARP / 2;            // G: float2 per lane and chunk
  f4 += ((size_t) nwarps * ROT_RS * 8 + 15) / 16;   // mbarriers
  f4 += (size_t) nwarps * ROT_RS * ld_floats / 4;   // ring rows (ld is a multiple of 4 floats)
  return f4 * sizeof(float4);
}

template <int NV, bool TRAIN>
__global__ void __launch_bounds__(STEP_MAX_WARPS * WARP, 1) k_step_rot(const StepParams p) {
  extern __shared__ float4 smem4[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nw = (int) (blockDim.x >> 5);
  const int wpp = p.wpp, ppc = nw / wpp;
  const int grp = warp / wpp, wig = warp % wpp;
  const int64_t i = (int64_t) blockIdx.x * ppc + grp;
  if (i >= p.B) return;                       // whole group leaves together
  const ModelDev& md = p.md;
  const int k = p.k_eff;

  float* s_neg = reinterpret_cast<float*>(smem4) + (size_t) grp * k;          // S (unclipped) of every negative: softmax pre-pass
  float* s_red0 = reinterpret_cast<float*>(smem4) + (((size_t) ppc * k + 3) & ~(size_t) 3);
  float* s_red = s_red0 + warp * 8;                                             // [0 .. N_PART) partials, [7] d loss / d pos share
  float4* s_Abase = smem4 + (((size_t) ppc * k + 3) / 4) + (nw * 8) / 4;
  float4* s_A = s_Abase + (size_t) warp * NV * WARP;
  float2* s_Gbase = reinterpret_cast<float2*>(s_Abase + (size_t) nw * NV * WARP);
  float2* s_G = s_Gbase + (size_t) warp * NV * WARP;
  uint64_t* ring_bar = reinterpret_cast<uint64_t*>(s_Gbase + (size_t) nw * NV * WARP) + warp * ROT_RS;

  const int64_t h = load_id(p.triples, 3 * i + 0, p.id64);
  const int64_t r = load_id(p.triples, 3 * i + 1, p.id64);
  const int64_t t = load_id(p.triples, 3 * i + 2, p.id64);

  const int n_it = (k - wig + wpp - 1) / wpp;
  int64_t my_e = 0;
  auto draw = [&](int it0) {
    const int c = wig + (it0 + lane) * wpp;
    if (c < k) my_e = p.neg_override ? p.neg_override[i * k + c] : sample_negative(p.sd, i, c, h, t);
  };

  const float* Ew = md.t[KGE_T_ENT].w;
  const int64_t ldE = md.t[KGE_T_ENT].ld;
  const int nch = (md.t[KGE_T_ENT].cols + 3) >> 2;
  const int ld4 = (int) (ldE >> 2);
  float4* ring = smem4 + (((size_t) ppc * k + 3) / 4) + ((size_t) nw * 8) / 4 + (size_t) nw * NV * WARP + (size_t) nw * NV * WARP / 2 +
                 ((size_t) nw * ROT_RS * 8 + 15) / 16 + (size_t) warp * ROT_RS * ld4;
  const uint32_t row_bytes = (uint32_t) ldE * 4u;
  // ring: row g of this warp's sequence (pre-pass rows first, then the main pass) lives in slot g % ROT_RS; the lane
  // that holds the row's id issues the copy (and files the id, main pass only)
  if (lane < ROT_RS) tc::mbar_init(ring_bar + lane, 1);
  tc::mbar_fence_init();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  auto ring_issue = [&](int g, int it, bool file) {        // it = index into this warp's negatives; my_e of lane it & 31 is its id
    if (lane == (it & 31)) {
      uint64_t* bar = ring_bar + (g % ROT_RS);
      tc::mbar_expect_tx(bar, row_bytes);
      bulk_row_g2s(ring + (size_t) (g % ROT_RS) * ld4, Ew + my_e * ldE, row_bytes, bar);
      if (file && p.write_keys) {
        const int64_t slot = i * k + wig + (int64_t) it * wpp;
        p.neg_ids[slot] = my_e;
        if (TRAIN) { p.keys[2 * p.B + slot] = (uint32_t) my_e; p.vals[2 * p.B + slot] = (uint32_t) (2 * p.B + slot); }
      }
    }
  };
  auto ring_wait = [&](int g) { tc::mbar_wait(ring_bar + (g % ROT_RS), (uint32_t) ((g / ROT_RS) & 1)); };
  const float4* Hp = reinterpret_cast<const float4*>(Ew + h * ldE);
  const float4* Tp = reinterpret_cast<const float4*>(Ew + t * ldE);
  const float* Rb = md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld;
  const bool head = neg_is_head(p.sd, wig);
  // (cos0, cos1, sin0, sin1) of the two complex elements of chunk c (RelRot::load)
  auto rel_cs = [&](int c) -> float4 {
    const float p0 = (2 * c < md.d_ent) ? __ldg(Rb + 2 * c) : 0.f;
    const float p1 = (2 * c + 1 < md.d_ent) ? __ldg(Rb + 2 * c + 1) : 0.f;
    float s0, c0, s1, c1;
    sincosf(p0 * md.rotate_scale, &s0, &c0);
    sincosf(p1 * md.rotate_scale, &s1, &c1);
    return f4(c0, c1, s0, s1);
  };
  auto ldrow = [&](const float4* rowp, int c) -> float4 { return c < nch ? __ldg(rowp + c) : f4zero(); };

  // X of this warp's side and the positive's score (MRotatE::forward + score_fwd, chunk by chunk)
  Row<NV> X;
  float spa = 0.f;
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const int c = lane + 32 * v;
    const float4 hh = ldrow(Hp, c), tt = ldrow(Tp, c), cs = rel_cs(c);
    const float4 xr = f4(hh.x * cs.x - hh.y * cs.z, hh.x * cs.z + hh.y * cs.x, hh.z * cs.y - hh.w * cs.w, hh.z * cs.w + hh.w * cs.y);
    const float4 u = xr - tt;
    spa = spa + (cmod(u.x, u.y) + cmod(u.z, u.w));
    X.v[v] = head ? f4(tt.x * cs.x + tt.y * cs.z, -tt.x * cs.z + tt.y * cs.x, tt.z * cs.y + tt.w * cs.w, -tt.z * cs.w + tt.w * cs.y) : xr;
  }
  const float Spos = warp_sum(spa);
  const float sp = -fmaxf(Spos, 1e-9f);

  // sum of moduli of X - N over the row
  auto row_dist = [&](const float4* Ns) {             // Ns: the row in shared memory
    float a = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int c = lane + 32 * v;
      const float4 u = X.v[v] - (c < nch ? Ns[c] : f4zero());
      a += sqrtf(u.x * u.x + u.y * u.y) + sqrtf(u.z * u.z + u.w * u.w);
    }
    return warp_sum(a);
  };
  // one pass of this warp over its negatives: rows g0 .. g0 + n_it of the ring sequence
  auto pass_begin = [&](int g0, bool file) {
    if (n_it > 0) draw(0);
    for (int j = 0; j < ROT_RS && j < n_it; ++j) ring_issue(g0 + j, j, file);
  };
  auto pass_next = [&](int g0, int it, bool file) {     // slot of row `it` is free: request row it + ROT_RS
    __syncwarp();
    const int nx = it + ROT_RS;
    if (nx < n_it) {
      if ((nx & 31) == 0) draw(nx);                      // every row of the previous 32 has been requested
      ring_issue(g0 + nx, nx, file);
    }
  };

  // self-adversarial weights need all k scores of this positive first (loss.py:176-177)
  float sm_max = 0.f, sm_inv = 1.f;
  const bool softmax = (p.loss.kind == KGE_LOSS_SELF_ADV);
  int g_main = 0;
  if (softmax) {
    pass_begin(0, false);
    for (int it = 0; it < n_it; ++it) {
      ring_wait(it);
      const float S = row_dist(ring + (size_t) (it % ROT_RS) * ld4);
      if (lane == 0) s_neg[wig + it * wpp] = S;
      pass_next(0, it, false);
    }
    g_main = n_it;
    group_barrier(grp, wpp);
    float m = -INFINITY;
    for (int c = lane; c < k; c += WARP) m = fmaxf(m, p.loss.temperature * -fmaxf(s_neg[c], 1e-9f));
    m = warp_max(m);
    float z = 0.f;
    for (int c = lane; c < k; c += WARP) z += expf(p.loss.temperature * -fmaxf(s_neg[c], 1e-9f) - m);
    z = warp_sum(z);
    sm_max = m;
    sm_inv = 1.0f / z;
  }

  if (TRAIN) {
#pragma unroll
    for (int v = 0; v < NV; ++v) { s_A[v * WARP + lane] = f4zero(); s_G[v * WARP + lane] = make_float2(0.f, 0.f); }
  }
  float nrmE = 0.f, nrmR = 0.f, lsum = 0.f, gpos = 0.f;
  const int64_t nE = 2 * p.B + p.B * (int64_t) k;
  const int rld = md.rld[KGE_T_ENT];

  pass_begin(g_main, true);
  for (int it = 0; it < n_it; ++it) {
    const int c = wig + it * wpp;
    const int64_t slot = i * k + c;
    ring_wait(g_main + it);
    const float4* Np = ring + (size_t) ((g_main + it) % ROT_RS) * ld4;
    const float S = softmax ? s_neg[c] : row_dist(Np);
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
      float4* rec = reinterpret_cast<float4*>(p.rec[KGE_T_ENT] + (2 * p.B + slot) * (int64_t) rld);
      float pr = 0.f, cnt = 0.f;
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const int c4 = lane + 32 * v;
        if (c4 < nch) {
          const float4 xv = X.v[v];
          const float4 u = xv - Np[c4];
          // 1 / |u| by MUFU.RSQ (<= 2 ulp) where k_step divides by an IEEE square root: the SCORE above is exact,
          // only the gradient's unit vectors carry the 2 ulp (36 instructions and two slow-path branches less per chunk)
          const float q0 = u.x * u.x + u.y * u.y, q1 = u.z * u.z + u.w * u.w;
          const float i0 = q0 > 0.f ? rsqrtf(q0) : 0.f, i1 = q1 > 0.f ? rsqrtf(q1) : 0.f;
          cnt += (q0 > 0.f ? 1.f : 0.f) + (q1 > 0.f ? 1.f : 0.f);
          const float4 uh = f4(u.x * i0, u.y * i0, u.z * i1, u.w * i1);
          float4 a = s_A[v * WARP + lane];
          a += coef * uh;
          s_A[v * WARP + lane] = a;
          // phase gradient: scale * Im(conj(X) * coef * unit); head corruption flips the sign (k_step)
          const float g0 = (xv.x * uh.y - xv.y * uh.x) * cps;
          const float g1 = (xv.z * uh.w - xv.w * uh.z) * cps;
          float2 g = s_G[v * WARP + lane];
          g.x += g0; g.y += g1;
          s_G[v * WARP + lane] = g;
          pr += g0 * g0 + g1 * g1;
          rec[c4] = (-coef) * uh;            // d score / d N
        }
      }
      nrmE += 2.0f * coef * coef * cnt;      // the negative's row and the kept entity's occurrence
      nrmR += pr;
    }
    pass_next(g_main, it, true);
  }

  // this warp's share of the kept entity's gradient: A back through the rotation, in place
  if (TRAIN) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int c = lane + 32 * v;
      const float4 q = s_A[v * WARP + lane], cs = rel_cs(c);
      // kept entity is t (head corrupted): rot(A);  kept entity is h: conj(r) o A
      const float4 o = head ? f4(q.x * cs.x - q.y * cs.z, q.x * cs.z + q.y * cs.x, q.z * cs.y - q.w * cs.w, q.z * cs.w + q.w * cs.y)
                            : f4(q.x * cs.x + q.y * cs.z, -q.x * cs.z + q.y * cs.x, q.z * cs.y + q.w * cs.w, -q.z * cs.w + q.w * cs.y);
      s_A[v * WARP + lane] = f4zero() + o;     // k_step adds to a zeroed row
    }
  }
  if (wig != 0) {
    const float rE = TRAIN ? warp_sum(nrmE) : 0.f, rR = TRAIN ? warp_sum(nrmR) : 0.f;
    if (lane == 0) { s_red[0] = lsum; s_red[1 + KGE_T_ENT] = rE; s_red[1 + KGE_T_ENT_PROJ] = 0.f; s_red[1 + KGE_T_REL_A] = rR; s_red[1 + KGE_T_REL_B] = 0.f; }
  }
  if (lane == 0) s_red[7] = gpos;
  group_barrier(grp, wpp);
  if (wig != 0) return;

  // ---- the positive's first warp: d loss / d pos, the positive's own backward, fixed-order sums, records ----
  float gp = 0.f;
  for (int w2 = 0; w2 < wpp; ++w2) gp += s_red0[(grp * wpp + w2) * 8 + 7];
  {
    float term, dpos;
    loss_pos(p.loss, sp, term, dpos);
    lsum += term;
    gp += dpos;
  }
  if (TRAIN) {
    const float coefp = (Spos >= 1e-9f) ? -gp : 0.f;          // score_bwd, SC_L1
    float4* recH = reinterpret_cast<float4*>(p.rec[KGE_T_ENT] + i * (int64_t) rld);
    float4* recT = reinterpret_cast<float4*>(p.rec[KGE_T_ENT] + (p.B + i) * (int64_t) rld);
    float* recR = p.rec[KGE_T_REL_A] + i * (int64_t) md.rld[KGE_T_REL_A];
    float sA = 0.f, sB = 0.f, sR = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int c = lane + 32 * v;
      const float4 hh = ldrow(Hp, c), tt = ldrow(Tp, c), cs = rel_cs(c);
      const float4 xr = f4(hh.x * cs.x - hh.y * cs.z, hh.x * cs.z + hh.y * cs.x, hh.z * cs.y - hh.w * cs.w, hh.z * cs.w + hh.w * cs.y);
      const float4 u = xr - tt;
      const float m0 = cmod(u.x, u.y), m1 = cmod(u.z, u.w);
      const float i0 = m0 > 0.f ? 1.0f / m0 : 0.f, i1 = m1 > 0.f ? 1.0f / m1 : 0.f;
      const float4 d = f4(1.f * u.x * i0, 1.f * u.y * i0, 1.f * u.z * i1, 1.f * u.w * i1);
      const float4 q = coefp * d;                                        // gx; gy = -gx
      const float4 gB = -q;
      const float4 gA = f4(q.x * cs.x + q.y * cs.z, -q.x * cs.z + q.y * cs.x, q.z * cs.y + q.w * cs.w, -q.z * cs.w + q.w * cs.y);
      const float gr0 = (-q.x * xr.y + q.y * xr.x) * md.rotate_scale, gr1 = (-q.z * xr.w + q.w * xr.z) * md.rotate_scale;
      sA += dot4(gA, gA); sB += dot4(gB, gB); sR += dot4(f4(gr0, gr1, 0.f, 0.f), f4(gr0, gr1, 0.f, 0.f));
      // warp 0 of the positive: its own rows first (k_step: GH.add(gA) ...), then the other warps in order
      const float4 own = s_A[v * WARP + lane];
      float4 gh = f4zero(), gt = f4zero();
      if (head) { gt += own; } else { gh += own; }
      gh += gA; gt += gB;
      float2 g = s_G[v * WARP + lane];
      float4 gr = f4zero() + f4(g.x, g.y, 0.f, 0.f);
      gr += f4(gr0, gr1, 0.f, 0.f);
      if (wpp > 1) {
        float4 th = f4zero(), tt2 = f4zero(), tr = f4zero();
        th += gh; tt2 += gt; tr += gr;
        for (int w2 = 1; w2 < wpp; ++w2) {
          const float4 o = s_Abase[((size_t) (grp * wpp + w2) * NV + v) * WARP + lane];
          const float2 g2 = s_Gbase[((size_t) (grp * wpp + w2) * NV + v) * WARP + lane];
          if (neg_is_head(p.sd, w2)) { tt2 += o; th += f4zero(); } else { th += o; tt2 += f4zero(); }
          tr += f4(g2.x, g2.y, 0.f, 0.f);
        }
        gh = th; gt = tt2; gr = tr;
      }
      if (c < nch) { recH[c] = gh; recT[c] = gt; }
      if (2 * c < md.d_ent) recR[2 * c] = gr.x;
      if (2 * c + 1 < md.d_ent) recR[2 * c + 1] = gr.y;
    }
    nrmE += sA; nrmE += sB; nrmR += sR;
  }
  float red[N_PART];
  red[0] = lsum;
#pragma unroll
  for (int q = 0; q < KGE_MAX_TABLES; ++q) red[1 + q] = 0.f;
  red[1 + KGE_T_ENT] = TRAIN ? warp_sum(nrmE) : 0.f;
  red[1 + KGE_T_REL_A] = TRAIN ? warp_sum(nrmR) : 0.f;
  if (wpp > 1) {
    float tot[N_PART];
#pragma unroll
    for (int q = 0; q < N_PART; ++q) tot[q] = 0.f + red[q];
    for (int w2 = 1; w2 < wpp; ++w2) {
      const float* sr = s_red0 + (grp * wpp + w2) * 8;
#pragma unroll
      for (int q = 0; q < N_PART; ++q) tot[q] += sr[q];
    }
#pragma unroll
    for (int q = 0; q < N_PART; ++q) red[q] = tot[q];
  }
  if (p.tail) {
    const int left = (int) (p.B - (int64_t) blockIdx.x * ppc), nvalid = left < ppc ? left : ppc;
    __syncwarp();
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
  if (TRAIN && lane == 0 && p.write_keys) {
    p.keys[i] = (uint32_t) h;           p.vals[i] = (uint32_t) i;
    p.keys[p.B + i] = (uint32_t) t;     p.vals[p.B + i] = (uint32_t) (p.B + i);
    p.keys[nE + i] = (uint32_t) (md.t[KGE_T_ENT].rows + r);
    p.vals[nE + i] = (uint32_t) (nE + i);
  }
}

template <bool TRAIN>
inline int launch_step_rot_nv(const StepParams& p, const StepLaunch& L) {
  auto go = [&](auto kern) -> int {
    if (L.smem > 48 * 1024) KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) L.smem));
    kern<<<(unsigned) L.grid, p.nwarps * WARP, L.smem, L.stream>>>(p);
    KGE_CUDA_OK(cudaGetLastError());
    return KGE_OK;
  };
  if (L.nv == 4) return go(k_step_rot<4, TRAIN>);
  if (L.nv == 8) return go(k_step_rot<8, TRAIN>);
  set_error("k_step_rot: rows of 4 or 8 chunks per lane (nv=%d)", L.nv);
  return KGE_EUNSUPPORTED;
}

}  // namespace kge
