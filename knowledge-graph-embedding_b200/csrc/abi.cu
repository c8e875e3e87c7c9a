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
  *out = reinterpret_cast<kge_ctx*>(c);
  return KGE_OK;
}
extern "C" int kge_ctx_destroy(kge_ctx* ctx) {
  delete reinterpret_cast<kge_ctx_impl*>(ctx);
  return KGE_OK;
}
