// step_mat_launch.cuh — launcher of k_step_mat; one translation unit per matrix model.
#pragma once
#include "step_launch.cuh"
#include "step_mat_kernels.cuh"
#include "step_proj_kernels.cuh"

namespace kge {

// shared memory (bytes) k_step_mat needs
inline size_t step_mat_smem_bytes(int nv, int k, int dpad, bool train, bool stage, int n_mats, int mrows, int mcols) {
  size_t f = (size_t) STEP_WARPS * dpad + 2 * (size_t) dpad + STEP_WARPS * 8 + 8 + (size_t) ((k + 3) & ~3);
  if (train) f += (size_t) STEP_WARPS * MAT_ACC_ROWS * nv * WARP * 4;
  if (stage) f += (size_t) n_mats * mrows * (mcols | 1);
  return f * sizeof(float);
}

template <class M, int NV, bool TRAIN>
int launch_step_mat_one(const MatStepParams& q, const StepLaunch& L) {
  if constexpr (!M::BILINEAR) {
    if (q.proj) {
      auto kp = k_step_proj<M, NV, TRAIN>;
      if (L.smem > 48 * 1024)
        KGE_CUDA_OK(cudaFuncSetAttribute(kp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) L.smem));
      kp<<<(unsigned) L.grid, STEP_THREADS, L.smem, L.stream>>>(q);
      KGE_CUDA_OK(cudaGetLastError());
      return KGE_OK;
    }
  }
  auto kern = k_step_mat<M, NV, TRAIN>;
  if (L.smem > 48 * 1024)
    KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) L.smem));
  kern<<<(unsigned) L.grid, STEP_THREADS, L.smem, L.stream>>>(q);
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}
template <class M>
int launch_step_mat_model(const MatStepParams& q, const StepLaunch& L) {
  if (L.nv == 1) return L.train ? launch_step_mat_one<M, 1, true>(q, L) : launch_step_mat_one<M, 1, false>(q, L);
  if (L.nv == 2) return L.train ? launch_step_mat_one<M, 2, true>(q, L) : launch_step_mat_one<M, 2, false>(q, L);
  set_error("matrix models support embedding sizes up to 256 (nv=%d)", L.nv);
  return KGE_EUNSUPPORTED;
}
int launch_step_se(const MatStepParams&, const StepLaunch&);
int launch_step_transr(const MatStepParams&, const StepLaunch&);
int launch_step_rescal(const MatStepParams&, const StepLaunch&);

}  // namespace kge
