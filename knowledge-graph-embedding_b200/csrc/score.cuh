// score.cuh — KGE/score.py:49-89 on warp-distributed rows, forward and analytic
// backward.  One warp owns one (x, y) pair.
#pragma once
#include "common.cuh"

namespace kge {

enum { SC_L1 = 0, SC_L2 = 1, SC_LINF = 2, SC_LPGEN = 3, SC_DOT = 4 };

struct ScoreDev {
  int kind;    // SC_*
  int pow2;    // LpDistancePow: s' = -(s^2)   (score.py:76)
  float p;     // for SC_LPGEN
};

// scalars saved by the forward for the backward
struct ScoreSaved {
  float S;     // reduced sum (pre-clip) or max
  float s;     // LpDistance value (before pow2)
  float ties;  // Linf: number of arg-max elements
};

__device__ __forceinline__ float cmod(float re, float im) { return sqrtf(re * re + im * im); }

// per-element forward contribution f(u) (real) / f(re, im) (complex)
template <bool CPLX>
__device__ __forceinline__ float chunk_fwd(int kind, float p, float4 u) {
  if (!CPLX) {
    switch (kind) {
      case SC_L1: return (fabsf(u.x) + fabsf(u.y)) + (fabsf(u.z) + fabsf(u.w));
      case SC_L2: return dot4(u, u);
      case SC_LINF: return fmaxf(fmaxf(fabsf(u.x), fabsf(u.y)), fmaxf(fabsf(u.z), fabsf(u.w)));
      default: return (powf(fabsf(u.x), p) + powf(fabsf(u.y), p)) + (powf(fabsf(u.z), p) + powf(fabsf(u.w), p));
    }
  } else {
    float m0 = cmod(u.x, u.y), m1 = cmod(u.z, u.w);
    switch (kind) {
      case SC_L1: return m0 + m1;
      case SC_L2: return dot4(u, u);
      case SC_LINF: return fmaxf(m0, m1);
      default: return powf(m0, p) + powf(m1, p);
    }
  }
}

// Forward: all lanes return the same score.  score.py:59-63 (Lp), :76 (pow), :89 (Dot).
template <int NV, bool CPLX, int W = 32>
__device__ __forceinline__ float score_fwd(const ScoreDev& sc, const Row<NV>& x, const Row<NV>& y,
                                           ScoreSaved& sv) {
  if (sc.kind == SC_DOT) {
    float s = warp_sum_w<W>(row_dot_partial(x, y));
    sv.S = s; sv.s = s; sv.ties = 1.f;
    return s;
  }
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float c = chunk_fwd<CPLX>(sc.kind, sc.p, x.v[i] - y.v[i]);
    acc = (sc.kind == SC_LINF) ? fmaxf(acc, c) : acc + c;
  }
  float S, s;
  if (sc.kind == SC_LINF) {
    S = warp_max_w<W>(acc);
    s = -S;
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float4 u = x.v[i] - y.v[i];
      if (!CPLX) {
        t += (fabsf(u.x) == S) + (fabsf(u.y) == S) + (fabsf(u.z) == S) + (fabsf(u.w) == S);
      } else {
        t += (cmod(u.x, u.y) == S) + (cmod(u.z, u.w) == S);
      }
    }
    sv.ties = warp_sum_w<W>(t);
  } else {
    S = warp_sum_w<W>(acc);
    float Sc = fmaxf(S, 1e-9f);                       // clip_by_value(., 1e-9, inf)
    s = (sc.kind == SC_L1) ? -Sc : (sc.kind == SC_L2 ? -sqrtf(Sc) : -powf(Sc, 1.0f / sc.p));
    sv.ties = 1.f;
  }
  sv.S = S; sv.s = s;
  return sc.pow2 ? -(s * s) : s;
}

// Backward: gx = g * ds/dx, gy = g * ds/dy.
template <int NV, bool CPLX>
__device__ __forceinline__ void score_bwd(const ScoreDev& sc, const Row<NV>& x, const Row<NV>& y,
                                          const ScoreSaved& sv, float g, Row<NV>& gx, Row<NV>& gy) {
  if (sc.kind == SC_DOT) {
#pragma unroll
    for (int i = 0; i < NV; ++i) { gx.v[i] = g * y.v[i]; gy.v[i] = g * x.v[i]; }
    return;
  }
  // d(-s^2)/ds = -2 s
  float gs = sc.pow2 ? g * (-2.0f * sv.s) : g;
  // s = -h(S): coef = gs * ds/dS (0 outside the clip's pass band: S < 1e-9)
  float coef;
  if (sc.kind == SC_L1) coef = (sv.S >= 1e-9f) ? -gs : 0.f;
  else if (sc.kind == SC_L2) coef = (sv.S >= 1e-9f) ? -gs * 0.5f / sqrtf(sv.S) : 0.f;
  else if (sc.kind == SC_LINF) coef = -gs / sv.ties;
  else coef = (sv.S >= 1e-9f) ? -gs * (1.0f / sc.p) * powf(sv.S, 1.0f / sc.p - 1.0f) : 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float4 u = x.v[i] - y.v[i];
    float4 d;
    if (!CPLX) {
      switch (sc.kind) {
        case SC_L1: d = f4(sgn(u.x), sgn(u.y), sgn(u.z), sgn(u.w)); break;
        case SC_L2: d = 2.0f * u; break;
        case SC_LINF:
          d = f4(fabsf(u.x) == sv.S ? sgn(u.x) : 0.f, fabsf(u.y) == sv.S ? sgn(u.y) : 0.f,
                 fabsf(u.z) == sv.S ? sgn(u.z) : 0.f, fabsf(u.w) == sv.S ? sgn(u.w) : 0.f);
          break;
        default: {
          float pm = sc.p - 1.0f;
          d = f4(u.x != 0.f ? sc.p * powf(fabsf(u.x), pm) * sgn(u.x) : 0.f,
                 u.y != 0.f ? sc.p * powf(fabsf(u.y), pm) * sgn(u.y) : 0.f,
                 u.z != 0.f ? sc.p * powf(fabsf(u.z), pm) * sgn(u.z) : 0.f,
                 u.w != 0.f ? sc.p * powf(fabsf(u.w), pm) * sgn(u.w) : 0.f);
        }
      }
    } else {
      float m0 = cmod(u.x, u.y), m1 = cmod(u.z, u.w);
      float i0 = m0 > 0.f ? 1.0f / m0 : 0.f, i1 = m1 > 0.f ? 1.0f / m1 : 0.f;
      float w0, w1;   // d f(|u|)/d|u|
      switch (sc.kind) {
        case SC_L1: w0 = 1.f; w1 = 1.f; break;
        case SC_L2: w0 = 2.f * m0; w1 = 2.f * m1; break;
        case SC_LINF: w0 = (m0 == sv.S) ? 1.f : 0.f; w1 = (m1 == sv.S) ? 1.f : 0.f; break;
        default:
          w0 = m0 > 0.f ? sc.p * powf(m0, sc.p - 1.0f) : 0.f;
          w1 = m1 > 0.f ? sc.p * powf(m1, sc.p - 1.0f) : 0.f;
      }
      d = f4(w0 * u.x * i0, w0 * u.y * i0, w1 * u.z * i1, w1 * u.w * i1);
    }
    gx.v[i] = coef * d;
    gy.v[i] = -gx.v[i];
  }
}

}  // namespace kge
