// model_transe.cu — kernel instantiations for MTransE (step_kernels.cuh, rank_kernels.cuh);
// one translation unit per model so they compile in parallel.
#include "step_launch.cuh"
#include "rank_kernels.cuh"
#include "step_narrow_kernels.cuh"
namespace kge {
int launch_step_transe(const StepParams& p, const StepLaunch& L) { return launch_step_model<MTransE>(p, L); }
int launch_score_transe(const ScoreParams& p, const AuxLaunch& L) { return launch_score_model<MTransE>(p, L); }
int launch_rank_transe(const RankParams& p, const AuxLaunch& L) { return launch_rank_model<MTransE>(p, L); }
int launch_step_narrow_transe(const StepParams& p, const StepLaunch& L) { return launch_step_narrow<MTransE>(p, L); }
}  // namespace kge
