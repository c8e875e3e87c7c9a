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

}  // namespace kge
