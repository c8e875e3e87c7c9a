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
  for (int q = 0; q < p.n_part; ++q) {
    const float* src = p.part + (size_t) q * p.B;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;       // four independent chains: loads overlap
    int64_t i = lane;
    for (; i + 3 * WARP < p.B; i += 4 * WARP) {
      a0 += __ldcg(src + i); a1 += __ldcg(src + i + WARP); a2 += __ldcg(src + i + 2 * WARP); a3 += __ldcg(src + i + 3 * WARP);
    }
    for (; i < p.B; i += WARP) a0 += __ldcg(src + i);
    const float a = warp_sum((a0 + a1) + (a2 + a3));
    if (lane == 0) {
      if (q == 0) { p.scal[8] = a; *p.loss_out = a; }
      else { p.scal[4 + q - 1] = a; p.scal[q - 1] = p.clip_norm / fmaxf(sqrtf(a), p.clip_norm); }
    }
  }
}
constexpr int64_t TAIL_REDUCE_MAX_B = 4096;

}  // namespace kge
