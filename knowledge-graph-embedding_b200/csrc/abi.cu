// abi.cu — context, versioning and error string of the C-ABI (include/kge_abi.h).
#include <stdarg.h>
#include <stdio.h>
#include "common.cuh"

namespace kge {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace kge
using namespace kge;

extern "C" int kge_abi_version(void) { return KGE_ABI_VERSION; }
extern "C" const char* kge_last_error(void) { return g_err; }

extern "C" int kge_ctx_create(int device, kge_ctx** out) {
  KGE_REQUIRE(out, KGE_EINVAL, "out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  KGE_REQUIRE(e == cudaSuccess && n > 0, KGE_ECUDA,
              "no CUDA device (%s): libkge_b200 has no CPU fallback", cudaGetErrorString(e));
  KGE_REQUIRE(device >= 0 && device < n, KGE_EINVAL, "device %d out of range [0, %d)", device, n);
  cudaDeviceProp prop;
  KGE_CUDA_OK(cudaGetDeviceProperties(&prop, device));
  KGE_REQUIRE(prop.major == 10, KGE_EUNSUPPORTED, "device is sm_%d%d; this library is built for sm_100a only",
              prop.major, prop.minor);
  KGE_CUDA_OK(cudaSetDevice(device));
  auto* c = new kge_ctx_impl;
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->smem_optin = (int) prop.sharedMemPerBlockOptin;
  int prio_least = 0, prio_greatest = 0;     // side stream first: its small kernels fill SMs the step kernel frees
  KGE_CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
  KGE_CUDA_OK(cudaStreamCreateWithPriority(&c->aux, cudaStreamNonBlocking, prio_greatest));
  KGE_CUDA_OK(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  KGE_CUDA_OK(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  KGE_CUDA_OK(cudaEventCreateWithFlags(&c->ev_mid, cudaEventDisableTiming));
  KGE_CUDA_OK(cudaEventCreateWithFlags(&c->ev_heavy, cudaEventDisableTiming));
  // the look-ahead grouping has a whole step of slack: lowest priority, it fills what the step and update kernels leave
  KGE_CUDA_OK(cudaStreamCreateWithPriority(&c->pre_stream, cudaStreamNonBlocking, prio_least));
  KGE_CUDA_OK(cudaEventCreateWithFlags(&c->ev_pre_fork, cudaEventDisableTiming));
  KGE_CUDA_OK(cudaEventCreateWithFlags(&c->ev_pre_done, cudaEventDisableTiming));
  KGE_CUDA_OK(cudaMalloc(&c->dyn, 64));
  KGE_CUDA_OK(cudaMemset(c->dyn, 0, 64));
  KGE_CUDA_OK(cudaMalloc(&c->ticket, 64 * sizeof(unsigned)));
  KGE_CUDA_OK(cudaMemset(c->ticket, 0, 64 * sizeof(unsigned)));
  *out = reinterpret_cast<kge_ctx*>(c);
  return KGE_OK;
}
static void free_events(kge_ctx_impl* c) {
  if (c->ev) {
    for (int i = 0; i < c->prof_cap * 2 * KGE_N_PHASES; ++i) cudaEventDestroy(c->ev[i]);
    delete[] c->ev; delete[] c->ev_mask;
    c->ev = nullptr; c->ev_mask = nullptr; c->prof_cap = c->prof_n = 0;
  }
}
extern "C" int kge_ctx_destroy(kge_ctx* ctx) {
  auto* c = reinterpret_cast<kge_ctx_impl*>(ctx);
  if (c) {
    free_events(c);
    if (c->aux) { cudaStreamSynchronize(c->aux); cudaStreamDestroy(c->aux); }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->ev_mid) cudaEventDestroy(c->ev_mid);
    if (c->ev_heavy) cudaEventDestroy(c->ev_heavy);
    if (c->pre_stream) { cudaStreamSynchronize(c->pre_stream); cudaStreamDestroy(c->pre_stream); }
    if (c->ev_pre_fork) cudaEventDestroy(c->ev_pre_fork);
    if (c->ev_pre_done) cudaEventDestroy(c->ev_pre_done);
    if (c->dyn) cudaFree(c->dyn);
    if (c->ticket) cudaFree(c->ticket);
  }
  delete c;
  return KGE_OK;
}
extern "C" int kge_set_option(kge_ctx* ctx, int option, int value) {
  auto* c = reinterpret_cast<kge_ctx_impl*>(ctx);
  KGE_REQUIRE(c, KGE_EINVAL, "ctx is NULL");
  switch (option) {
    case KGE_OPT_RADIX_SORT: c->opt_radix_sort = value; return KGE_OK;
    case KGE_OPT_NO_SHADOW: c->opt_no_shadow = value; return KGE_OK;
    case KGE_OPT_NO_FUSED_PUSH: c->opt_no_fused_push = value; return KGE_OK;
    case KGE_OPT_RING_SLOTS: c->opt_ring_max = value; return KGE_OK;
    case KGE_OPT_GRAPH_PROFILE: c->opt_graph_profile = value; return KGE_OK;
    case KGE_OPT_RANK_TC_PASSES: c->opt_tc_passes = value; return KGE_OK;
    case KGE_OPT_CLUSTER_GROUP: c->opt_cluster_group = value; return KGE_OK;
    case KGE_OPT_PRE_EARLY: c->opt_pre_early = value; return KGE_OK;
    default: set_error("unknown option %d", option); return KGE_EINVAL;
  }
}
extern "C" int kge_profile_begin(kge_ctx* ctx, int max_calls) {
  auto* c = reinterpret_cast<kge_ctx_impl*>(ctx);
  KGE_REQUIRE(c && max_calls > 0 && max_calls <= (1 << 16), KGE_EINVAL, "bad profile_begin arguments");
  free_events(c);
  c->ev = new cudaEvent_t[(size_t) max_calls * 2 * KGE_N_PHASES];
  c->ev_mask = new unsigned[max_calls];
  for (int i = 0; i < max_calls * 2 * KGE_N_PHASES; ++i) KGE_CUDA_OK(cudaEventCreate(&c->ev[i]));
  c->prof_cap = max_calls;
  c->prof_n = 0;
  return KGE_OK;
}
extern "C" int kge_profile_end(kge_ctx* ctx, float* phase_ms, int* n_calls) {
  auto* c = reinterpret_cast<kge_ctx_impl*>(ctx);
  KGE_REQUIRE(c && phase_ms && n_calls, KGE_EINVAL, "bad profile_end arguments");
  KGE_REQUIRE(c->ev, KGE_EINVAL, "kge_profile_begin was not called");
  KGE_CUDA_OK(cudaDeviceSynchronize());
  for (int p = 0; p < KGE_N_PHASES; ++p) phase_ms[p] = 0.f;
  for (int s = 0; s < c->prof_n; ++s)
    for (int p = 0; p < KGE_N_PHASES; ++p)
      if (c->ev_mask[s] & (1u << p)) {
        float ms = 0.f;
        KGE_CUDA_OK(cudaEventElapsedTime(&ms, c->ev[s * 2 * KGE_N_PHASES + 2 * p], c->ev[s * 2 * KGE_N_PHASES + 2 * p + 1]));
        phase_ms[p] += ms;
      }
  *n_calls = c->prof_n;
  free_events(c);
  return KGE_OK;
}
