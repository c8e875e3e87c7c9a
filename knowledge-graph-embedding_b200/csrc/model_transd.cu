// model_transd.cu — kernel instantiations for MTransD (step_kernels.cuh, rank_kernels.cuh);
// one translation unit per model so they compile in parallel.
#include "step_launch.cuh"
#include "rank_kernels.cuh"
#include "rank_proj_tile.cuh"
namespace kge {
int launch_step_transd(const StepParams& p, const StepLaunch& L) { return launch_step_model<MTransD>(p, L); }
int launch_score_transd(const ScoreParams& p, const AuxLaunch& L) { return launch_score_model<MTransD>(p, L); }
int launch_rank_transd(const RankParams& p, const AuxLaunch& L) { return launch_rank_model<MTransD>(p, L); }
int launch_rank_proj_transd(const ProjParams& p, int nv, int kind, int64_t span, int sm_count, cudaStream_t st) {
  return launch_rank_proj_vec<MTransD>(p, nv, kind, span, sm_count, st);
}
}  // namespace kge
