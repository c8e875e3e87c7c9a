// filter.cu — known-positive filter index for ranking, built on the device.
// Replaces the two full scans of positive_X the reference does PER QUERY (BaseModel.py:646-650:
// r_mask, e_mask, positive_X[r_mask & e_mask, corrupt_side]) by
//   kge_filter_index_build : ONE sort of positive_X per (positive_X, side) —
//       composite key (r * N + kept entity) * N + candidate entity  ->  cub radix sort (64-bit keys)
//       -> cub unique (a triple listed twice must be filtered once)  ->  split into key[] | ent[]
//   kge_filter_lookup      : per query two binary searches over key[] -> [begin, end) into ent[]
// kge_rank then reads ent[begin[q] .. end[q]) in place (kge_filter.end): nothing is gathered or copied,
// and no step needs the host (the number of unique triples stays in the index header on the device).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include "host_common.cuh"

namespace kge {

constexpr size_t FIDX_HDR = 256;                  // bytes: int64 n_unique | n_pos | side | n_ent  (device-resident)
struct FilterIdxLayout { size_t off_key, off_ent, total; };
inline void layout_filter_index(int64_t n_pos, FilterIdxLayout& L) {
  L.off_key = FIDX_HDR;
  L.off_ent = L.off_key + align_up(sizeof(int64_t) * (size_t) n_pos);
  L.total = L.off_ent + align_up(sizeof(int64_t) * (size_t) n_pos);
}
struct FilterWsLayout { size_t off_a, off_b, off_cub, cub_bytes, total; };

static int comp_bits(const kge_cfg* c) {          // bits of the largest composite key
  const long double mx = (long double) c->n_rel * (long double) c->n_ent * (long double) c->n_ent;
  int b = 1;
  while (b < 64 && ldexpl(1.0L, b) < mx) ++b;
  return b;
}

static int layout_filter_ws(const kge_cfg* c, int64_t n_pos, FilterWsLayout& L) {
  size_t sort_b = 0, uniq_b = 0;
  cudaError_t e = cub::DeviceRadixSort::SortKeys(nullptr, sort_b, (const uint64_t*) nullptr, (uint64_t*) nullptr, n_pos, 0,
                                                 comp_bits(c) + 1, (cudaStream_t) 0);
  KGE_REQUIRE(e == cudaSuccess, KGE_ECUDA, "cub sort temp-size query failed: %s", cudaGetErrorString(e));
  e = cub::DeviceSelect::Unique(nullptr, uniq_b, (const uint64_t*) nullptr, (uint64_t*) nullptr, (int64_t*) nullptr, n_pos,
                                (cudaStream_t) 0);
  KGE_REQUIRE(e == cudaSuccess, KGE_ECUDA, "cub unique temp-size query failed: %s", cudaGetErrorString(e));
  L.cub_bytes = sort_b > uniq_b ? sort_b : uniq_b;
  L.off_a = 0;
  L.off_b = L.off_a + align_up(sizeof(uint64_t) * (size_t) n_pos);
  L.off_cub = L.off_b + align_up(sizeof(uint64_t) * (size_t) n_pos);
  L.total = L.off_cub + align_up(L.cub_bytes ? L.cub_bytes : 1);
  return KGE_OK;
}

// composite key of every positive triple; ids outside the tables get the key 2^bits (`invalid`: above
// every real key, so it sorts last, survives `unique` once and is dropped by k_filter_split)
__global__ void __launch_bounds__(256) k_filter_keys(const void* __restrict__ pos, int id64, int64_t n, int side,
                                                     uint64_t N, uint64_t R, uint64_t invalid,
                                                     uint64_t* __restrict__ out) {
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
    const uint64_t h = (uint64_t) load_id(pos, 3 * i, id64), r = (uint64_t) load_id(pos, 3 * i + 1, id64),
                   t = (uint64_t) load_id(pos, 3 * i + 2, id64);
    const uint64_t keep = (side == KGE_SIDE_H) ? t : h, cand = (side == KGE_SIDE_H) ? h : t;
    out[i] = (h < N && t < N && r < R) ? (r * N + keep) * N + cand : invalid;
  }
}

// unique composites -> (pair key, candidate entity); the tail [n_unique, n) is left as it is (never read)
__global__ void __launch_bounds__(256) k_filter_split(const uint64_t* __restrict__ comp, int64_t* __restrict__ hdr,
                                                      int64_t n, uint64_t N, uint64_t invalid,
                                                      int64_t* __restrict__ key, int64_t* __restrict__ ent) {
  int64_t nu = hdr[0];
  // out-of-table triples left ONE `invalid` key at the very end: drop it from the index
  if (nu > 0 && comp[nu - 1] == invalid) nu -= 1;
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nu; i += (int64_t) gridDim.x * blockDim.x) {
    const uint64_t c = comp[i];
    key[i] = (int64_t) (c / N);
    ent[i] = (int64_t) (c % N);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) hdr[4] = nu;          // what the look-ups search
}

// [begin, end) of every query's pair key (r, kept entity) in the sorted key array
__global__ void __launch_bounds__(256) k_filter_lookup(const int64_t* __restrict__ hdr, const int64_t* __restrict__ key,
                                                       const void* __restrict__ queries, int id64, int64_t Q, int side,
                                                       int64_t N, int64_t* __restrict__ begin, int64_t* __restrict__ end) {
  const int64_t q = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  const int64_t n = hdr[4];
  const int64_t r = load_id(queries, 3 * q + 1, id64);
  const int64_t keep = load_id(queries, 3 * q + (side == KGE_SIDE_H ? 2 : 0), id64);
  const int64_t k = r * N + keep;
  int64_t lo = 0, hi = n;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (__ldg(key + mid) < k) lo = mid + 1; else hi = mid; }
  const int64_t b = lo;
  // upper bound: filter lists are short, so gallop from b instead of a second full search
  int64_t step = 1, e = b;
  while (e + step < n && __ldg(key + e + step - 1) <= k) { e += step; step <<= 1; }
  int64_t l2 = e, h2 = (e + step < n) ? e + step : n;
  while (l2 < h2) { const int64_t mid = (l2 + h2) >> 1; if (__ldg(key + mid) <= k) l2 = mid + 1; else h2 = mid; }
  begin[q] = b;
  end[q] = l2;
}

}  // namespace kge
using namespace kge;

extern "C" int kge_filter_index_bytes(const kge_cfg* cfg, int64_t n_pos, size_t* index_bytes, size_t* ws_bytes) {
  if (int rc = check_cfg(cfg)) return rc;
  KGE_REQUIRE(n_pos > 0 && index_bytes && ws_bytes, KGE_EINVAL, "n_pos / index_bytes / ws_bytes invalid");
  KGE_REQUIRE(comp_bits(cfg) <= 63, KGE_EUNSUPPORTED, "n_rel * n_ent^2 must stay below 2^63 (composite sort key)");
  FilterIdxLayout I;
  layout_filter_index(n_pos, I);
  FilterWsLayout W;
  if (int rc = layout_filter_ws(cfg, n_pos, W)) return rc;
  *index_bytes = I.total;
  *ws_bytes = W.total;
  return KGE_OK;
}

extern "C" int kge_filter_index_build(kge_ctx* ctx_, const kge_cfg* cfg, const void* positive, int64_t n_pos, int side,
                                      void* index, size_t index_bytes, void* ws, size_t ws_bytes, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  cudaStream_t st = (cudaStream_t) stream;
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  KGE_REQUIRE(positive && n_pos > 0 && index && ws, KGE_EINVAL, "positive / n_pos / index / ws invalid");
  KGE_REQUIRE(side == KGE_SIDE_H || side == KGE_SIDE_T, KGE_EINVAL, "side must be 'h' or 't'");
  KGE_REQUIRE(comp_bits(cfg) <= 63, KGE_EUNSUPPORTED, "n_rel * n_ent^2 must stay below 2^63 (composite sort key)");
  FilterIdxLayout I;
  layout_filter_index(n_pos, I);
  FilterWsLayout W;
  if (int rc = layout_filter_ws(cfg, n_pos, W)) return rc;
  KGE_REQUIRE(index_bytes >= I.total, KGE_EWORKSPACE, "filter index %zu < required %zu bytes", index_bytes, I.total);
  KGE_REQUIRE(ws_bytes >= W.total, KGE_EWORKSPACE, "filter workspace %zu < required %zu bytes", ws_bytes, W.total);
  KGE_REQUIRE((((uintptr_t) index | (uintptr_t) ws) & 255) == 0, KGE_EINVAL, "index / workspace must be 256-byte aligned");
  char* ib = reinterpret_cast<char*>(index);
  char* wb = reinterpret_cast<char*>(ws);
  int64_t* hdr = reinterpret_cast<int64_t*>(ib);
  uint64_t* a = reinterpret_cast<uint64_t*>(wb + W.off_a);
  uint64_t* b = reinterpret_cast<uint64_t*>(wb + W.off_b);
  int64_t blocks = (n_pos + 255) / 256;
  if (blocks > (int64_t) ctx->sm_count * 16) blocks = (int64_t) ctx->sm_count * 16;
  const int bits = comp_bits(cfg);
  const uint64_t invalid = 1ull << bits;
  KGE_CUDA_OK(cudaMemsetAsync(hdr, 0, FIDX_HDR, st));
  k_filter_keys<<<(unsigned) blocks, 256, 0, st>>>(positive, cfg->id_dtype, n_pos, side, (uint64_t) cfg->n_ent,
                                                   (uint64_t) cfg->n_rel, invalid, a);
  KGE_CUDA_OK(cudaGetLastError());
  size_t tb = W.cub_bytes;
  KGE_CUDA_OK(cub::DeviceRadixSort::SortKeys(wb + W.off_cub, tb, (const uint64_t*) a, b, n_pos, 0, bits + 1, st));
  tb = W.cub_bytes;
  KGE_CUDA_OK(cub::DeviceSelect::Unique(wb + W.off_cub, tb, (const uint64_t*) b, a, hdr, n_pos, st));
  k_filter_split<<<(unsigned) blocks, 256, 0, st>>>(a, hdr, n_pos, (uint64_t) cfg->n_ent, invalid,
                                                    reinterpret_cast<int64_t*>(ib + I.off_key),
                                                    reinterpret_cast<int64_t*>(ib + I.off_ent));
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}

extern "C" int kge_filter_lookup(kge_ctx* ctx_, const kge_cfg* cfg, const void* index, int64_t n_pos, const void* queries,
                                 int64_t Q, int side, int64_t* begin_out, int64_t* end_out, kge_filter* filter_out,
                                 void* stream) {
  KGE_REQUIRE(ctx_, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  KGE_REQUIRE(index && n_pos > 0 && queries && Q > 0 && begin_out && end_out, KGE_EINVAL,
              "index / n_pos / queries / Q / begin_out / end_out invalid");
  KGE_REQUIRE(side == KGE_SIDE_H || side == KGE_SIDE_T, KGE_EINVAL, "side must be 'h' or 't'");
  FilterIdxLayout I;
  layout_filter_index(n_pos, I);
  const char* ib = reinterpret_cast<const char*>(index);
  k_filter_lookup<<<(unsigned) ((Q + 255) / 256), 256, 0, (cudaStream_t) stream>>>(
      reinterpret_cast<const int64_t*>(ib), reinterpret_cast<const int64_t*>(ib + I.off_key), queries, cfg->id_dtype, Q, side,
      cfg->n_ent, begin_out, end_out);
  KGE_CUDA_OK(cudaGetLastError());
  if (filter_out) {
    filter_out->off = begin_out;
    filter_out->end = end_out;
    filter_out->ids = reinterpret_cast<const int64_t*>(ib + I.off_ent);
  }
  return KGE_OK;
}
