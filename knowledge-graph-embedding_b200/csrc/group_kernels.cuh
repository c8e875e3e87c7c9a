// group_kernels.cuh — grouping of the step's gradient records by (table, row) key.
//
// What the update kernels need is, per key, the list of its record ids in a CANONICAL order
// (ascending record id = the order of the lookups in the reference's IndexedSlices, which is what
// a stable sort by key yields).  Keys are dense small integers (< n_ent + n_rel), so a counting
// sort does it in four small launches instead of a multi-pass radix sort:
//
//   k_keys      (train_step.cu)  key of every record slot + histogram        cnt[key]++
//   k_scan      exclusive scan of cnt -> off[0..K]   (seg_lo = off, seg_hi = off + 1)
//   k_scatter   vals_tmp[off[key] + --cnt[key]] = slot     arrival order: NOT deterministic;
//               leaves cnt all zero again
//   k_canon     per key: sort the segment's record ids ascending -> vals     deterministic
//
// k_canon: segments of <= 32 ids are sorted by a warp bitonic network; longer ones by bitmap
// windows in shared memory (ids are unique, so a set bit per id and an ordered walk over the
// words is a sort; only the windows between the segment's min and max id are visited).
// The result is bit-for-bit what cub::DeviceRadixSort::SortPairs (stable) produces
// (tests/test_gpu_step.py::test_grouping_equals_stable_radix_sort).
#pragma once
#include "common.cuh"

namespace kge {

constexpr int SCAN_THREADS = 1024;
constexpr int64_t SCAN_SMALL_MAX = 1 << 20;       // single-CTA scan up to this many keys

// exclusive scan of cnt[0..K) into off[0..K]; one CTA (K is the number of table rows, small)
__global__ void __launch_bounds__(SCAN_THREADS) k_scan(const uint32_t* __restrict__ cnt, uint32_t* __restrict__ off, int64_t K) {
  __shared__ uint32_t s_w[32];
  __shared__ uint32_t s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  constexpr int PER = 4;                                   // elements per thread per round
  for (int64_t base = 0; base < K; base += SCAN_THREADS * PER) {
    const int64_t i0 = base + (int64_t) threadIdx.x * PER;
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) { v[j] = (i0 + j < K) ? cnt[i0 + j] : 0u; sum += v[j]; }
    uint32_t inc = sum;                                    // inclusive scan over the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = s_w[lane], winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, winc, o); if (lane >= o) winc += t; }
      s_w[lane] = winc - w;                                // exclusive over warps
    }
    __syncthreads();
    uint32_t run = s_carry + s_w[warp] + (inc - sum);
#pragma unroll
    for (int j = 0; j < PER; ++j) { if (i0 + j < K) off[i0 + j] = run; run += v[j]; }
    __syncthreads();
    if (threadIdx.x == SCAN_THREADS - 1) s_carry = run;
    __syncthreads();
  }
  if (threadIdx.x == 0) off[K] = s_carry;
}

// large K: cub::DeviceScan writes off[0..K); this closes the array
__global__ void k_scan_close(const uint32_t* cnt, uint32_t* off, int64_t K) { off[K] = off[K - 1] + cnt[K - 1]; }

__global__ void __launch_bounds__(256) k_scatter(const uint32_t* __restrict__ keys, int64_t n, const uint32_t* __restrict__ off,
                                                 uint32_t* cnt, uint32_t* __restrict__ vals_tmp) {
  for (int64_t s = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (int64_t) gridDim.x * blockDim.x) {
    const uint32_t key = keys[s];
    const uint32_t slot = atomicSub(cnt + key, 1u) - 1u;   // counts return to zero for the next step
    vals_tmp[off[key] + slot] = (uint32_t) s;
  }
}

__device__ __forceinline__ uint32_t warp_bitonic_sort(uint32_t v, int lane) {
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      const uint32_t o = __shfl_xor_sync(FULL, v, j);
      const bool up = ((lane & k) == 0);                   // ascending block
      const bool lower = ((lane & j) == 0);
      v = (lower == up) ? min(v, o) : max(v, o);
    }
  }
  return v;
}

constexpr int CANON_WORDS = 1024;                          // 32768-bit window per warp
constexpr int CANON_WARPS = 8;

// one warp per key
__global__ void __launch_bounds__(CANON_WARPS * 32) k_canon(const uint32_t* __restrict__ off, int64_t K,
                                                            const uint32_t* __restrict__ vals_tmp, uint32_t* __restrict__ vals,
                                                            uint32_t* __restrict__ hdr, uint32_t* __restrict__ heavy_keys,
                                                            uint32_t* n_heavy) {
  __shared__ uint32_t s_bits[CANON_WARPS][CANON_WORDS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int64_t key = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (; key < K; key += nw) {
    const uint32_t lo = off[key], hi = off[key + 1], L = hi - lo;
    // header of the key for k_update_fast: [count, first position, first six ids, ascending]
    if (lane == 0) {
      hdr[key * 8] = L; hdr[key * 8 + 1] = lo;
      if (L > (uint32_t) UPD_HEAVY) heavy_keys[atomicAdd(n_heavy, 1u)] = (uint32_t) key;   // list order is irrelevant
    }
    if (L == 0) continue;
    if (L <= 32) {
      uint32_t v = (lane < L) ? vals_tmp[lo + lane] : 0xffffffffu;
      if (L > 1) v = warp_bitonic_sort(v, lane);
      if (lane < L) vals[lo + lane] = v;
      if (lane < 6 && lane < L) hdr[key * 8 + 2 + lane] = v;
      continue;
    }
    // long segment: min / max id, then bitmap windows
    uint32_t mn = 0xffffffffu, mx = 0u;
    for (uint32_t i = lane; i < L; i += 32) { const uint32_t v = vals_tmp[lo + i]; mn = min(mn, v); mx = max(mx, v); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mn = min(mn, __shfl_xor_sync(FULL, mn, o)); mx = max(mx, __shfl_xor_sync(FULL, mx, o)); }
    uint32_t* bits = s_bits[warp];
    uint32_t written = 0;
    for (uint64_t w0 = mn; w0 <= mx; w0 += (uint64_t) CANON_WORDS * 32) {
      for (int j = lane; j < CANON_WORDS; j += 32) bits[j] = 0u;
      __syncwarp();
      for (uint32_t i = lane; i < L; i += 32) {
        const uint64_t v = vals_tmp[lo + i];
        if (v >= w0 && v < w0 + (uint64_t) CANON_WORDS * 32) atomicOr(bits + ((v - w0) >> 5), 1u << ((v - w0) & 31));
      }
      __syncwarp();
      // lane owns words [lane*32, lane*32+32): ids in lane order are ascending
      uint32_t mine = 0;
      for (int j = 0; j < CANON_WORDS / 32; ++j) mine += __popc(bits[lane * (CANON_WORDS / 32) + j]);
      uint32_t inc = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
      uint32_t at = lo + written + (inc - mine);
      for (int j = 0; j < CANON_WORDS / 32; ++j) {
        const int wj = lane * (CANON_WORDS / 32) + j;
        uint32_t b = bits[wj];
        while (b) {
          const int bit = __ffs(b) - 1;
          b &= b - 1;
          vals[at++] = (uint32_t) (w0 + (uint64_t) wj * 32 + bit);
        }
      }
      written += __shfl_sync(FULL, inc, 31);
      __syncwarp();
    }
    if (lane < 6) hdr[key * 8 + 2 + lane] = vals[lo + lane];       // written by this warp, ordered by __syncwarp
  }
}


// ---- the same grouping in ONE kernel on one thread-block cluster (small key spaces) ----------------------
// k_keys -> k_scan -> k_scatter -> k_canon are four dependent launches of little work each: 36 us of mostly
// launch and tail latency for 70 k records, the longer branch of a C1 / C2 step.  With K = n_ent + n_rel keys
// small enough for the counters to live in the shared memory of one cluster (8 CTAs x 1024 threads,
// K <= GC_CTAS * GC_MAX_KC), the phases are separated by cluster barriers instead of kernel boundaries and the
// histogram / scatter counters are distributed-shared-memory atomics:
//   0  zero this CTA's slice of the counters                                   (CTA c owns keys [c Kc, (c+1) Kc))
//   1  key of every record (GroupKeyFn: ids, Philox negatives) -> keys[], remote atomicAdd on the owner's counter
//   2  exclusive scan: per-CTA slice scan, slice totals exchanged through DSMEM -> off[] (shared + global)
//   3  scatter: position = off[key] + (atomicSub(cnt[key]) - 1) on the owner's slice -> vals_tmp[]
//   4  canonical order of every segment of <= 32 records (warp bitonic network) + the 32-byte headers; longer
//      segments (hot rows) are queued for k_canon_long, the bitmap-window pass of k_canon
// The results (off, vals, headers, heavy list) are bit for bit those of the four-kernel chain.
namespace cgx {
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `p` (a shared-memory address of THIS CTA) in CTA `rank` of the cluster, as a shared::cluster address
__device__ __forceinline__ uint32_t map_rank(const void* p, uint32_t rank) {
  uint32_t a = (uint32_t) __cvta_generic_to_shared(p), r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ uint32_t atom_add_cluster(uint32_t addr, uint32_t v) {
  uint32_t old;
  asm volatile("atom.shared::cluster.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(addr), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t ld_cluster(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_cluster(uint32_t addr, uint32_t v) {
  asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
}  // namespace cgx

constexpr int GC_CTAS = 8, GC_THREADS = 1024;
constexpr int GC_MAX_KC = 24576;                   // keys per CTA slice: 2 x 96 KB of shared memory
struct GroupOut {
  uint32_t* keys; uint32_t* vals_tmp; uint32_t* vals; uint32_t* off; uint32_t* hdr;
  uint32_t* heavy_keys; uint32_t* n_heavy; uint32_t* long_keys; uint32_t* n_long;
  int64_t K, n_keys; int Kc;
};
inline size_t group_cluster_smem(int Kc) { return sizeof(uint32_t) * (2 * (size_t) Kc + 1 + GC_CTAS + 64); }

template <class KeyFn>
__global__ void __cluster_dims__(GC_CTAS, 1, 1) __launch_bounds__(GC_THREADS, 1) k_group_cluster(const KeyFn kf, const GroupOut g) {
  extern __shared__ uint32_t gsm[];
  uint32_t* cnt = gsm;                        // [Kc]
  uint32_t* off = gsm + g.Kc;                 // [Kc + 1]
  uint32_t* tot = off + g.Kc + 1;             // [GC_CTAS]   slice totals of every CTA
  uint32_t* wsum = tot + GC_CTAS;             // [32] + carry
  const uint32_t rank = cgx::cluster_ctarank();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t key0 = (int64_t) rank * g.Kc;
  int64_t nk = g.K - key0;                    // keys of this slice
  nk = nk < 0 ? 0 : (nk > g.Kc ? g.Kc : nk);
  for (int i = tid; i < g.Kc; i += GC_THREADS) cnt[i] = 0u;
  if (rank == 0 && tid == 0) { *g.n_heavy = 0u; *g.n_long = 0u; }
  cgx::cluster_sync();

  // 1. keys + histogram
  const int64_t gt = (int64_t) rank * GC_THREADS + tid, gn = (int64_t) GC_CTAS * GC_THREADS;
  for (int64_t s = gt; s < g.n_keys; s += gn) {
    const uint32_t key = kf(s);
    g.keys[s] = key;
    const uint32_t owner = key / (uint32_t) g.Kc, idx = key - owner * (uint32_t) g.Kc;
    cgx::atom_add_cluster(cgx::map_rank(cnt + idx, owner), 1u);
  }
  cgx::cluster_sync();

  // 2. exclusive scan of this slice (per-thread runs -> warp scan -> CTA scan), then the slices' bases
  const int per = (g.Kc + GC_THREADS - 1) / GC_THREADS;
  const int i0 = tid * per;
  uint32_t sum = 0;
  for (int j = 0; j < per; ++j) sum += (i0 + j < g.Kc) ? cnt[i0 + j] : 0u;
  uint32_t inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = wsum[lane];
    uint32_t winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, winc, o); if (lane >= o) winc += t; }
    wsum[lane] = winc - w;
    if (lane == 31) wsum[32] = winc;          // slice total
  }
  __syncthreads();
  uint32_t run = wsum[warp] + (inc - sum);    // exclusive prefix inside the slice
  for (int j = 0; j < per; ++j)
    if (i0 + j < g.Kc) { off[i0 + j] = run; run += cnt[i0 + j]; }
  if (tid < GC_CTAS) cgx::st_cluster(cgx::map_rank(tot + rank, (uint32_t) tid), wsum[32]);   // my total to every CTA
  cgx::cluster_sync();
  uint32_t base = 0;
  for (uint32_t c = 0; c < rank; ++c) base += tot[c];
  for (int i = tid; i < g.Kc; i += GC_THREADS) {
    const uint32_t o = off[i] + base;
    off[i] = o;
    if (i < nk) g.off[key0 + i] = o;
  }
  if (rank == GC_CTAS - 1 && tid == 0) g.off[g.K] = base + tot[rank];
  cgx::cluster_sync();

  // 3. scatter (arrival order inside a segment is arbitrary; phase 4 / k_canon_long order it)
  for (int64_t s = gt; s < g.n_keys; s += gn) {
    const uint32_t key = g.keys[s];
    const uint32_t owner = key / (uint32_t) g.Kc, idx = key - owner * (uint32_t) g.Kc;
    const uint32_t slot = cgx::atom_add_cluster(cgx::map_rank(cnt + idx, owner), 0xffffffffu) - 1u;
    g.vals_tmp[cgx::ld_cluster(cgx::map_rank(off + idx, owner)) + slot] = (uint32_t) s;
  }
  __threadfence();
  cgx::cluster_sync();

  // 4. canonical order + headers of this CTA's keys (one warp per key)
  for (int64_t i = warp; i < nk; i += GC_THREADS / 32) {
    const int64_t key = key0 + i;
    const uint32_t lo = off[i];
    const uint32_t hi = (i + 1 < nk) ? off[i + 1] : base + tot[rank];
    const uint32_t L = hi - lo;
    if (lane == 0) {
      g.hdr[key * 8] = L; g.hdr[key * 8 + 1] = lo;
      if (L > (uint32_t) UPD_HEAVY) g.heavy_keys[atomicAdd(g.n_heavy, 1u)] = (uint32_t) key;
      if (L > 32u) g.long_keys[atomicAdd(g.n_long, 1u)] = (uint32_t) key;
    }
    if (L == 0 || L > 32u) continue;
    uint32_t v = (lane < L) ? __ldcg(g.vals_tmp + lo + lane) : 0xffffffffu;
    if (L > 1) v = warp_bitonic_sort(v, lane);
    if (lane < L) g.vals[lo + lane] = v;
    if (lane < 6 && lane < L) g.hdr[key * 8 + 2 + lane] = v;
  }
}

// segments of more than 32 records left by k_group_cluster: the bitmap-window pass of k_canon, one warp per key
__global__ void __launch_bounds__(CANON_WARPS * 32) k_canon_long(const uint32_t* __restrict__ off, const uint32_t* __restrict__ long_keys,
                                                                 const uint32_t* __restrict__ n_long, const uint32_t* __restrict__ vals_tmp,
                                                                 uint32_t* __restrict__ vals, uint32_t* __restrict__ hdr) {
  __shared__ uint32_t s_bits[CANON_WARPS][CANON_WORDS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n = *n_long;
  for (uint32_t item = blockIdx.x * CANON_WARPS + warp; item < n; item += gridDim.x * CANON_WARPS) {
    const int64_t key = long_keys[item];
    const uint32_t lo = off[key], L = off[key + 1] - lo;
    uint32_t mn = 0xffffffffu, mx = 0u;
    for (uint32_t i = lane; i < L; i += 32) { const uint32_t v = vals_tmp[lo + i]; mn = min(mn, v); mx = max(mx, v); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mn = min(mn, __shfl_xor_sync(FULL, mn, o)); mx = max(mx, __shfl_xor_sync(FULL, mx, o)); }
    uint32_t* bits = s_bits[warp];
    uint32_t written = 0;
    for (uint64_t w0 = mn; w0 <= mx; w0 += (uint64_t) CANON_WORDS * 32) {
      for (int j = lane; j < CANON_WORDS; j += 32) bits[j] = 0u;
      __syncwarp();
      for (uint32_t i = lane; i < L; i += 32) {
        const uint64_t v = vals_tmp[lo + i];
        if (v >= w0 && v < w0 + (uint64_t) CANON_WORDS * 32) atomicOr(bits + ((v - w0) >> 5), 1u << ((v - w0) & 31));
      }
      __syncwarp();
      uint32_t mine = 0;
      for (int j = 0; j < CANON_WORDS / 32; ++j) mine += __popc(bits[lane * (CANON_WORDS / 32) + j]);
      uint32_t inc = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
      uint32_t at = lo + written + (inc - mine);
      for (int j = 0; j < CANON_WORDS / 32; ++j) {
        const int wj = lane * (CANON_WORDS / 32) + j;
        uint32_t b = bits[wj];
        while (b) {
          const int bit = __ffs(b) - 1;
          b &= b - 1;
          vals[at++] = (uint32_t) (w0 + (uint64_t) wj * 32 + bit);
        }
      }
      written += __shfl_sync(FULL, inc, 31);
      __syncwarp();
    }
    if (lane < 6) hdr[key * 8 + 2 + lane] = vals[lo + lane];
  }
}

}  // namespace kge
