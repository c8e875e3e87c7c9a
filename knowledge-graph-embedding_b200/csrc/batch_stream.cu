// batch_stream.cu — the batch pipeline on the device.
// Replaces the reference's tf.data pipeline (KGE/data_utils.py:176-196: from_tensor_slices / csv readers ->
// shuffle -> repeat -> batch -> one (B, 3) tensor per step, copied host -> device every step) for graphs that fit
// the GPU (every bundled one does: FB15k is 5.8 MB of int32 ids):
//   kge_epoch_perm   : the order of one pass over the n training triples — a permutation drawn from the library's
//                      counter-based stream: key_i = Philox4x32-10(seed; counter = (i, epoch, bit 1)) (the sampler's
//                      generator, oracle/philox.py), perm = STABLE argsort of the keys (cub radix sort of
//                      (key, i) pairs): reproducible, no state, any epoch can be regenerated (resume)
//   kge_batch_gather : batch = triples[perm[offset .. offset + B)]  (perm NULL = identity: shuffle off), ids of
//                      either width in, the caller's width out
// Nothing touches the host: a training loop holds the triples and the permutation on the device and asks for one
// (B, 3) tensor per step; with the look-ahead step it asks one batch early.
#include <cub/device/device_radix_sort.cuh>
#include "host_common.cuh"

namespace kge {
struct PermWs { size_t off_kin, off_kout, off_vin, off_cub, cub_bytes, total; };
static int layout_perm_ws(int64_t n, PermWs& L) {
  size_t b = 0;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, b, (const uint32_t*) nullptr, (uint32_t*) nullptr, (const uint32_t*) nullptr,
                                                  (uint32_t*) nullptr, n, 0, 32, (cudaStream_t) 0);
  KGE_REQUIRE(e == cudaSuccess, KGE_ECUDA, "cub sort temp-size query failed: %s", cudaGetErrorString(e));
  L.cub_bytes = b;
  L.off_kin = 0;
  L.off_kout = L.off_kin + align_up(sizeof(uint32_t) * (size_t) n);
  L.off_vin = L.off_kout + align_up(sizeof(uint32_t) * (size_t) n);
  L.off_cub = L.off_vin + align_up(sizeof(uint32_t) * (size_t) n);
  L.total = L.off_cub + align_up(b ? b : 1);
  return KGE_OK;
}
__global__ void __launch_bounds__(256) k_perm_keys(int64_t n, uint64_t seed, uint64_t epoch, uint32_t* __restrict__ keys,
                                                   uint32_t* __restrict__ vals) {
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
    keys[i] = philox_word0(seed, epoch, 1u, (uint64_t) i);
    vals[i] = (uint32_t) i;
  }
}
__global__ void __launch_bounds__(256) k_batch_gather(const void* __restrict__ triples, int in64, const uint32_t* __restrict__ perm,
                                                      int64_t offset, int64_t B, void* __restrict__ out, int out64) {
  const int64_t x = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;     // one thread per id
  if (x >= 3 * B) return;
  const int64_t j = x / 3, c = x - 3 * j;
  const int64_t src = perm ? (int64_t) __ldg(perm + offset + j) : offset + j;
  const int64_t v = load_id(triples, 3 * src + c, in64);
  if (out64) reinterpret_cast<int64_t*>(out)[x] = v;
  else reinterpret_cast<int32_t*>(out)[x] = (int32_t) v;
}
}  // namespace kge
using namespace kge;

extern "C" int kge_epoch_perm_bytes(int64_t n, size_t* ws_bytes) {
  KGE_REQUIRE(n > 0 && n < 0xffffffffll && ws_bytes, KGE_EINVAL, "n / ws_bytes invalid");
  PermWs L;
  if (int rc = layout_perm_ws(n, L)) return rc;
  *ws_bytes = L.total;
  return KGE_OK;
}

extern "C" int kge_epoch_perm(kge_ctx* ctx_, int64_t n, uint64_t seed, uint64_t epoch, uint32_t* perm_out, void* ws,
                              size_t ws_bytes, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  KGE_REQUIRE(n > 0 && n < 0xffffffffll && perm_out, KGE_EINVAL, "n / perm_out invalid");
  PermWs L;
  if (int rc = layout_perm_ws(n, L)) return rc;
  KGE_REQUIRE(ws && ws_bytes >= L.total, KGE_EWORKSPACE, "permutation workspace %zu < required %zu bytes", ws_bytes, L.total);
  KGE_REQUIRE(((uintptr_t) ws & 255) == 0, KGE_EINVAL, "workspace must be 256-byte aligned");
  char* w = reinterpret_cast<char*>(ws);
  uint32_t* kin = reinterpret_cast<uint32_t*>(w + L.off_kin);
  uint32_t* kout = reinterpret_cast<uint32_t*>(w + L.off_kout);
  uint32_t* vin = reinterpret_cast<uint32_t*>(w + L.off_vin);
  cudaStream_t st = (cudaStream_t) stream;
  int64_t blocks = (n + 255) / 256;
  if (blocks > (int64_t) ctx->sm_count * 16) blocks = (int64_t) ctx->sm_count * 16;
  k_perm_keys<<<(unsigned) blocks, 256, 0, st>>>(n, seed, epoch, kin, vin);
  KGE_CUDA_OK(cudaGetLastError());
  size_t cb = L.cub_bytes;
  KGE_CUDA_OK(cub::DeviceRadixSort::SortPairs(w + L.off_cub, cb, kin, kout, vin, perm_out, n, 0, 32, st));   // stable: ties keep index order
  return KGE_OK;
}

extern "C" int kge_batch_gather(kge_ctx* ctx_, int in_id_dtype, const void* triples, int64_t n, const uint32_t* perm, int64_t offset,
                                int64_t B, int out_id_dtype, void* out, void* stream) {
  KGE_REQUIRE(ctx_, KGE_EINVAL, "ctx is NULL");
  KGE_REQUIRE(triples && out && n > 0 && B > 0, KGE_EINVAL, "triples / out / n / B invalid");
  KGE_REQUIRE(offset >= 0 && offset + B <= n, KGE_EINVAL, "batch [%lld, %lld) leaves the pass of %lld triples", (long long) offset,
              (long long) (offset + B), (long long) n);
  KGE_REQUIRE((in_id_dtype == 0 || in_id_dtype == 1) && (out_id_dtype == 0 || out_id_dtype == 1), KGE_EINVAL, "id dtype must be 0 (int32) or 1 (int64)");
  k_batch_gather<<<(unsigned) ((3 * B + 255) / 256), 256, 0, (cudaStream_t) stream>>>(triples, in_id_dtype, perm, offset, B, out, out_id_dtype);
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}
