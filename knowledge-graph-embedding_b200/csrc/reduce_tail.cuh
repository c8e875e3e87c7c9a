// reduce_tail.cuh — parameters of the partial reduction and its in-kernel ("tail") form,
// shared by the step kernels (every model TU) and k_reduce (train_step.cu).
#pragma once
#include "common.cuh"

namespace kge {

// ---- fixed-order reduction of per-positive partials ---------------------------------
// scal[0..3] = clip scale per table, scal[4..7] = norm^2 per table, scal[8] = loss.
struct ReduceParams {
  const float* part;     // [n_part][B]
  int64_t B;
  int n_part;
  float clip_norm;
  float* scal;
  float* loss_out;
};

// The same reduction run by the LAST warp to finish the step kernel (saves a launch when B is
// small): partials are read with ld.cg after a __threadfence()/ticket handshake.
__device__ __forceinline__ void tail_reduce_warp(const ReduceParams& p, int lane) {
  // Every load of a batch is issued before the first add (a warp issues in order: an add waiting for its
  // operand would hold back the loads behind it, one L2 round trip per iteration).  16 values per lane
  // and quantity in flight; four fixed accumulation chains per quantity.
  constexpr int U = 8;
  float acc[5][4];
#pragma unroll
  for (int q = 0; q < 5; ++q) acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = 0.f;
  for (int64_t i0 = lane; i0 < p.B; i0 += (int64_t) U * WARP) {
    float v[5][U];
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + (int64_t) u * WARP;
        v[q][u] = (q < p.n_part && i < p.B) ? __ldcg(p.part + (size_t) q * p.B + i) : 0.f;
      }
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
      for (int u = 0; u < U; ++u) acc[q][u & 3] += v[q][u];
  }
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    const float a = warp_sum((acc[q][0] + acc[q][1]) + (acc[q][2] + acc[q][3]));
    if (lane == 0 && q < p.n_part) {
      if (q == 0) { p.scal[8] = a; *p.loss_out = a; }
      else { p.scal[4 + q - 1] = a; p.scal[q - 1] = p.clip_norm / fmaxf(sqrtf(a), p.clip_norm); }
    }
  }
}
constexpr int64_t TAIL_REDUCE_MAX_B = 4096;

}  // namespace kge
