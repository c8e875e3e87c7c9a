// reduce_tail.cuh — parameters of the partial reduction and its in-kernel ("tail") form,
// shared by the step kernels (every model TU) and k_reduce (train_step.cu).
#pragma once
#include "common.cuh"

namespace kge {

// ---- fixed-order reduction of per-positive partials ---------------------------------
// scal[0..3] = clip scale per table, scal[4..7] = norm^2 per table, scal[8] = loss.
constexpr int N_PART_MAX = 5;     // loss + norm^2 per table (1 + KGE_MAX_TABLES)
struct ReduceParams {
  const float* part;     // [n_part][B]
  int64_t B;
  int n_part;
  float clip_norm;
  float* scal;
  float* loss_out;
  float* stage;          // k_reduce: [N_PART_MAX][RED_CTAS] per-CTA sums (behind the partials in the workspace)
  unsigned* done;        // k_reduce: self-resetting ticket counter
  const DynScal* dyn;    // captured step graph: where the host wants the loss (DynScal.host_loss / host_seq), else NULL
};
// the loss is final: hand it to a host that waits for it (kge_step_graph_wait_loss) before the update kernels run
__device__ __forceinline__ void publish_loss_to_host(const DynScal* dyn, float loss) {
  if (dyn == nullptr) return;
  float* hl = dyn->host_loss;
  if (hl == nullptr) return;
  *reinterpret_cast<volatile float*>(hl) = loss;
  __threadfence_system();
  *reinterpret_cast<volatile unsigned long long*>(dyn->host_seq) = dyn->seq;
}
constexpr int RED_CTAS = 64;      // k_reduce: CTAs of the first level (one SM's L2 bandwidth was the single-CTA kernel's limit)

// The same reduction folded into the step kernel (saves a launch when B is small), in two levels so that
// the serial end of the kernel stays short: every CTA sums the partials of its own positives (shared memory)
// and publishes ONE value per quantity, part[q][cta]; the LAST CTA to finish (ticket counter) sums those —
// n_cta values per quantity, all loads of a batch issued before the first add (a warp issues in order: an add
// waiting for its operand would hold back the loads behind it, one L2 round trip per iteration).
__device__ __forceinline__ void tail_reduce_ctas(const ReduceParams& p, int n_cta, int lane) {
  constexpr int U = 8;
  float acc[5];
#pragma unroll
  for (int q = 0; q < 5; ++q) acc[q] = 0.f;
  for (int c0 = lane; c0 < n_cta; c0 += U * WARP) {
    float v[5][U];
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int c = c0 + u * WARP;
        v[q][u] = (q < p.n_part && c < n_cta) ? __ldcg(p.part + (size_t) q * n_cta + c) : 0.f;
      }
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
      for (int u = 0; u < U; ++u) acc[q] += v[q][u];
  }
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    const float a = warp_sum(acc[q]);
    if (lane == 0 && q < p.n_part) {
      if (q == 0) { p.scal[8] = a; *p.loss_out = a; publish_loss_to_host(p.dyn, a); }
      else { p.scal[4 + q - 1] = a; p.scal[q - 1] = p.clip_norm / fmaxf(sqrtf(a), p.clip_norm); }
    }
  }
}
constexpr int64_t TAIL_REDUCE_MAX_B = 4096;

}  // namespace kge
