// step_launch.cuh — per-model launcher for k_step; each step_<model>.cu instantiates
// one model so the translation units compile in parallel.
#pragma once
#include "step_kernels.cuh"

namespace kge {

struct StepLaunch {
  int nv;          // float4 chunks per lane: 1, 2, 4, 8
  int train;
  int64_t grid;
  size_t smem;
  cudaStream_t stream;
};

// shared memory k_step needs for (nv, wpp, k) and a model with n_ent/n_rel row bundles
inline size_t step_smem_bytes(int nv, int wpp, int k, int ent_rows, int rel_rows, int nwarps = STEP_WARPS) {
  const int ppc = nwarps / wpp;
  size_t f4 = ((size_t) ppc * k + 3) / 4 + (nwarps * 8) / 4;
  if (wpp > 1) f4 += (size_t) nwarps * (2 * ent_rows + rel_rows) * nv * WARP;
  return f4 * sizeof(float4);
}

template <class M, int NV, bool TRAIN>
int launch_step_one(const StepParams& p, const StepLaunch& L) {
  auto kern = k_step<M, NV, TRAIN>;
  if (L.smem > 48 * 1024)
    KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) L.smem));
  const int threads = StepBounds<M, NV>::WIDE ? p.nwarps * WARP : STEP_THREADS;
  kern<<<(unsigned) L.grid, threads, L.smem, L.stream>>>(p);
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}

template <class M>
int launch_step_model(const StepParams& p, const StepLaunch& L) {
#define KGE_NV_CASE(N)                                                    \
  case N:                                                                 \
    return L.train ? launch_step_one<M, N, true>(p, L) : launch_step_one<M, N, false>(p, L);
  switch (L.nv) {
    KGE_NV_CASE(1)
    KGE_NV_CASE(2)
    KGE_NV_CASE(4)
    KGE_NV_CASE(8)
  }
#undef KGE_NV_CASE
  set_error("unsupported row width (nv=%d)", L.nv);
  return KGE_EUNSUPPORTED;
}

// one per step_<model>.cu
int launch_step_um(const StepParams&, const StepLaunch&);
int launch_step_transe(const StepParams&, const StepLaunch&);
int launch_step_transh(const StepParams&, const StepLaunch&);
int launch_step_transd(const StepParams&, const StepLaunch&);
int launch_step_rotate(const StepParams&, const StepLaunch&);
int launch_step_narrow_transe(const StepParams&, const StepLaunch&);   // step_narrow_kernels.cuh
int launch_step_narrow_um(const StepParams&, const StepLaunch&);
int launch_step_rot(const StepParams&, const StepLaunch&);      // step_rot_kernels.cuh (model_rotate.cu)
size_t step_rot_smem_bytes_host(int nv, int wpp, int k, int nwarps, int ld);
int launch_step_distmult(const StepParams&, const StepLaunch&);

}  // namespace kge
