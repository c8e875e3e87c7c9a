// model_distmult.cu — kernel instantiations for MDistMult (step_kernels.cuh, rank_kernels.cuh);
// one translation unit per model so they compile in parallel.
#include "step_launch.cuh"
#include "rank_kernels.cuh"
namespace kge {
int launch_step_distmult(const StepParams& p, const StepLaunch& L) { return launch_step_model<MDistMult>(p, L); }
int launch_score_distmult(const ScoreParams& p, const AuxLaunch& L) { return launch_score_model<MDistMult>(p, L); }
int launch_rank_distmult(const RankParams& p, const AuxLaunch& L) { return launch_rank_model<MDistMult>(p, L); }
}  // namespace kge
