// model_um.cu — kernel instantiations for MUM (step_kernels.cuh, rank_kernels.cuh);
// one translation unit per model so they compile in parallel.
#include "step_launch.cuh"
#include "rank_kernels.cuh"
#include "step_narrow_kernels.cuh"
namespace kge {
int launch_step_um(const StepParams& p, const StepLaunch& L) { return launch_step_model<MUM>(p, L); }
int launch_score_um(const ScoreParams& p, const AuxLaunch& L) { return launch_score_model<MUM>(p, L); }
int launch_rank_um(const RankParams& p, const AuxLaunch& L) { return launch_rank_model<MUM>(p, L); }
int launch_step_narrow_um(const StepParams& p, const StepLaunch& L) { return launch_step_narrow<MUM>(p, L); }
}  // namespace kge
