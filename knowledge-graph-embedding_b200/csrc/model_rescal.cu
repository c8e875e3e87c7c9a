// model_rescal.cu — kernel instantiations for MRESCAL (step_mat_kernels.cuh, rank_mat_kernels.cuh).
#include "step_mat_launch.cuh"
#include "rank_mat_kernels.cuh"
namespace kge {
int launch_step_rescal(const MatStepParams& q, const StepLaunch& L) { return launch_step_mat_model<MRESCAL>(q, L); }
int launch_score_rescal(const ScoreParams& p, const AuxLaunch& L, const MatAux& a, size_t smem) {
  return launch_score_mat_model<MRESCAL>(p, L, a, smem);
}
int launch_rank_rescal(const RankParams& p, const AuxLaunch& L, const MatAux& a, size_t smem) {
  return launch_rank_mat_model<MRESCAL>(p, L, a, smem);
}
}  // namespace kge
