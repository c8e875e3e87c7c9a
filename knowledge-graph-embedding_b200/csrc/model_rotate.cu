// model_rotate.cu — kernel instantiations for MRotatE (step_kernels.cuh, rank_kernels.cuh);
// one translation unit per model so they compile in parallel.
#include "step_launch.cuh"
#include "rank_kernels.cuh"
#include "step_rot_kernels.cuh"
namespace kge {
int launch_step_rotate(const StepParams& p, const StepLaunch& L) { return launch_step_model<MRotatE>(p, L); }
int launch_step_rot(const StepParams& p, const StepLaunch& L) { return L.train ? launch_step_rot_nv<true>(p, L) : launch_step_rot_nv<false>(p, L); }
size_t step_rot_smem_bytes_host(int nv, int wpp, int k, int nwarps, int ld) { return step_rot_smem_bytes(nv, wpp, k, nwarps, ld); }
int launch_score_rotate(const ScoreParams& p, const AuxLaunch& L) { return launch_score_model<MRotatE>(p, L); }
int launch_rank_rotate(const RankParams& p, const AuxLaunch& L) { return launch_rank_model<MRotatE>(p, L); }
}  // namespace kge
