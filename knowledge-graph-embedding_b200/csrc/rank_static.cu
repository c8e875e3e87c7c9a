// rank_static.cu — kge_rank_lp_static: the CUDA-core tiled Lp ranking (rank_lp_tile.cuh) for callers that
// cannot allocate or pass scratch memory (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
//
// The kernels are exactly those of kge_rank's Lp-tile path; only where the scratch lives differs: a
// file-scope `static __device__` array of KGE_RANK_STATIC_BYTES (every GPU that loads the library holds
// it until exit).  In HOST code the array's name is a host-side stand-in, not a device address, so the
// device address is resolved once per device with cudaGetSymbolAddress (that device current) and that
// pointer is what the kernels receive.  The entry point calls no cudaMalloc / cudaFree / allocating
// library routine.  All calls share the one array, which keeps the previous call's contents: the part
// a call uses is cleared first, and calls that overlap on the GPU (two streams, two host threads) race.
#include "host_common.cuh"
#include "rank_lp_tile.cuh"

namespace kge {

static __device__ __align__(256) unsigned char g_rank_static[KGE_RANK_STATIC_BYTES];

constexpr int MAX_DEVICES = 64;
static void* g_rank_static_ptr[MAX_DEVICES] = {};      // device address of g_rank_static on device i (0 = not resolved yet)

// once per device, with that device current
static int rank_static_ptr(int device, void** out) {
  KGE_REQUIRE(device >= 0 && device < MAX_DEVICES, KGE_EUNSUPPORTED, "device %d beyond the %d the static scratch tracks",
              device, MAX_DEVICES);
  if (!g_rank_static_ptr[device]) {
    int cur = -1;
    KGE_CUDA_OK(cudaGetDevice(&cur));
    if (cur != device) KGE_CUDA_OK(cudaSetDevice(device));
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_rank_static);
    if (cur != device) cudaSetDevice(cur);
    KGE_REQUIRE(e == cudaSuccess && ptr, KGE_ECUDA, "cudaGetSymbolAddress(g_rank_static) failed: %s", cudaGetErrorString(e));
    KGE_REQUIRE(((uintptr_t) ptr & 255) == 0, KGE_ECUDA, "static scratch is not 256-byte aligned");
    g_rank_static_ptr[device] = ptr;
  }
  *out = g_rank_static_ptr[device];
  return KGE_OK;
}

}  // namespace kge
using namespace kge;

extern "C" int kge_rank_lp_static(kge_ctx* ctx_, const kge_cfg* cfg, const kge_tables* tables, const void* queries,
                                  int64_t Q, int side, const kge_filter* filter, int64_t ent_lo, int64_t ent_hi,
                                  int32_t* count_out, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  if (int rc = check_tables(cfg, tables, false)) return rc;
  KGE_REQUIRE(queries && count_out && Q > 0, KGE_EINVAL, "queries / count_out / Q invalid");
  KGE_REQUIRE(side == KGE_SIDE_H || side == KGE_SIDE_T, KGE_EINVAL, "side must be 'h' or 't' (BaseModel.py:639-644)");
  KGE_REQUIRE(0 <= ent_lo && ent_lo < ent_hi && ent_hi <= cfg->n_ent, KGE_EINVAL, "bad entity shard [%lld, %lld)",
              (long long) ent_lo, (long long) ent_hi);
  KGE_REQUIRE(!filter || !filter->off || filter->ids, KGE_EINVAL, "filter.off given without filter.ids");
  KGE_REQUIRE(rank_lp_eligible(cfg), KGE_EUNSUPPORTED,
              "kge_rank_lp_static covers the Lp-tile path only (UM / TransE / RotatE with LpDistance / LpDistancePow)");
  LpWs L;
  layout_lp_ws(cfg, Q, L);
  KGE_REQUIRE(L.total <= (size_t) KGE_RANK_STATIC_BYTES, KGE_EWORKSPACE,
              "kge_rank_lp_static: %lld queries of %d floats need %zu bytes of scratch, the static array holds %d "
              "(split the queries, or use kge_rank with a caller workspace)",
              (long long) Q, L.ldq, L.total, (int) KGE_RANK_STATIC_BYTES);
  void* ws = nullptr;
  if (int rc = rank_static_ptr(ctx->device, &ws)) return rc;
  // the array keeps the last call's contents: clear what this call uses (no allocation: a memset on the device pointer)
  KGE_CUDA_OK(cudaMemsetAsync(ws, 0, L.total, (cudaStream_t) stream));
  return rank_lp_tiled(cfg, tables, queries, Q, side, filter, ent_lo, ent_hi, count_out, ws, L.total, (cudaStream_t) stream);
}
