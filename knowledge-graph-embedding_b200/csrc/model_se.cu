// model_se.cu — kernel instantiations for MSE (step_mat_kernels.cuh, rank_mat_kernels.cuh).
#include "step_mat_launch.cuh"
#include "rank_mat_kernels.cuh"
#include "rank_proj_tile.cuh"
namespace kge {
int launch_step_se(const MatStepParams& q, const StepLaunch& L) { return launch_step_mat_model<MSE>(q, L); }
int launch_score_se(const ScoreParams& p, const AuxLaunch& L, const MatAux& a, size_t smem) {
  return launch_score_mat_model<MSE>(p, L, a, smem);
}
int launch_rank_se(const RankParams& p, const AuxLaunch& L, const MatAux& a, size_t smem) {
  return launch_rank_mat_model<MSE>(p, L, a, smem);
}
int launch_rank_proj_se(const ProjParams& p, int nv, int kind, int64_t span, int sm_count, cudaStream_t st) {
  return launch_rank_proj_mat<MSE>(p, nv, kind, span, sm_count, st);
}
}  // namespace kge
