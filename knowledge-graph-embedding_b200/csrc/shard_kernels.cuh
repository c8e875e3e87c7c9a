// shard_kernels.cuh — training step with the ENTITY TABLES SHARDED over the GPUs of one
// NVLink / NVSwitch node (configuration C5).  Rank g owns entity rows
// [g S, (g+1) S): their w, m, v and `last` live in its HBM only; w is also mapped into every other
// rank (peer memory), relation tables are replicated.  One step, every rank in lock-step:
//
//   A  lookup  (rank)   keys -> sort -> k_heads (lazy_kernels.cuh) -> k_route: the row ids this
//                       rank's positives and negatives look up are STORED INTO THE OWNER'S INBOX
//                       over NVLink                                              [barrier]
//   B  pre     (owner)  k_claim / k_contrib: de-duplicate the ids of all sources into a work list;
//                       k_lazy_pre: replay skipped steps + start-of-batch constraint   [barrier]
//   C  grad    (rank)   k_step gathers rows straight from the owners' HBM (128-bit peer loads),
//                       k_shard_push stores one summed gradient row per looked-up row into the
//                       owner's value inbox; relation gradients / norms / loss go through the
//                       data-parallel all-reduce                                 [all-reduce]
//   D  apply   (owner)  k_shard_update: contributions summed in RANK ORDER (deterministic), clip,
//                       m / v replay, Adam, last = t
//
// Traffic per rank and step over NVLink: 4 B per looked-up row (ids), one row read and one row
// written per looked-up remote row; nothing is all-gathered.  The update work (6 row moves per
// touched row) is divided by the number of ranks.
#pragma once
#include "lazy_kernels.cuh"

namespace kge {

struct ShardDev {
  int world, rank;
  uint32_t S;                           // rows per shard
  uint32_t cap;                         // inbox rows per (owner, source)
  uint32_t* ids_peer[KGE_MAX_PEERS];    // owner j's id inbox      [world][cap]
  uint32_t* cnt_peer[KGE_MAX_PEERS];    // owner j's count inbox   [world]
  float* val_peer[KGE_MAX_PEERS][2];    // owner j's value inbox of entity slot s  [world][cap][ld]
  uint32_t *stamp, *idx_of, *work, *contrib;   // owner-local
};

// ---- A: route the looked-up row ids to their owners -----------------------------------------
// uniq = k_heads' work list (global row ids).  Slot inside the (owner, this rank) inbox section by
// a warp-aggregated atomic on a local counter; item_slot remembers it for k_shard_push.
// rows with a single gradient record: rec_dest[record] = owner << 29 | slot, so that k_step stores that
// record straight into the owner's inbox (rec_dest is preset to REC_LOCAL = 0xffffffff).
__global__ void __launch_bounds__(256) k_route(const uint32_t* __restrict__ uniq, ShardDev sd, uint32_t* __restrict__ rcnt,
                                               uint32_t* __restrict__ item_slot, const uint32_t* __restrict__ vals,
                                               uint32_t* __restrict__ rec_dest) {
  const uint32_t n = __ldg(uniq);
  const int lane = threadIdx.x & 31;
  for (int64_t base = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) & ~31ll; base < n; base += (int64_t) gridDim.x * blockDim.x) {
    const int64_t it = base + lane;
    const bool ok = it < n;
    uint32_t row = 0, o = 0xffffffffu;
    if (ok) { row = __ldg(uniq + 4 + 3 * it); o = row / sd.S; }
    const unsigned peers = __match_any_sync(FULL, o);
    if (ok) {
      const int leader = __ffs(peers) - 1;
      uint32_t b = 0;
      if (lane == leader) b = atomicAdd(rcnt + o, (uint32_t) __popc(peers));
      b = __shfl_sync(peers, b, leader);
      const uint32_t slot = b + __popc(peers & ((1u << lane) - 1));
      sd.ids_peer[o][(size_t) sd.rank * sd.cap + slot] = row - o * sd.S;
      item_slot[it] = slot;
      const uint32_t lo = __ldg(uniq + 4 + 3 * it + 1), hi = __ldg(uniq + 4 + 3 * it + 2);
      if (hi - lo == 1u) rec_dest[__ldg(vals + lo)] = (o << 29) | slot;
    }
  }
}
__global__ void k_route_cnt(ShardDev sd, const uint32_t* __restrict__ rcnt) {
  const int o = threadIdx.x;
  if (o < sd.world) sd.cnt_peer[o][sd.rank] = rcnt[o];
}

// ---- B: owner de-duplicates ------------------------------------------------------------------
// blockIdx.y = source rank.  stamp[row] = last step that looked the row up (steps only grow, so the
// array is never cleared); the first source to raise it appends the row to the work list.
__global__ void __launch_bounds__(256) k_claim(ShardDev sd, uint32_t t) {
  const int r = blockIdx.y, lane = threadIdx.x & 31;
  const uint32_t n = sd.cnt_peer[sd.rank][r];
  const uint32_t* ids = sd.ids_peer[sd.rank] + (size_t) r * sd.cap;
  for (int64_t base = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) & ~31ll; base < n; base += (int64_t) gridDim.x * blockDim.x) {
    const int64_t j = base + lane;
    bool first = false;
    uint32_t row = 0;
    if (j < n) { row = ids[j]; first = atomicMax(sd.stamp + row, t) < t; }
    const unsigned bal = __ballot_sync(FULL, first);
    if (bal) {
      const int leader = __ffs(bal) - 1;
      uint32_t b = 0;
      if (lane == leader) b = atomicAdd(sd.work, (uint32_t) __popc(bal));
      b = __shfl_sync(FULL, b, leader);
      if (first) {
        const uint32_t u = b + __popc(bal & ((1u << lane) - 1));
        sd.work[4 + 3 * (size_t) u] = row;
        sd.idx_of[row] = u;
      }
    }
  }
}
// contrib[u][r] = 1 + position of row work[u] in source r's inbox section (0 = source r did not look it up)
__global__ void __launch_bounds__(256) k_contrib(ShardDev sd) {
  const int r = blockIdx.y;
  const uint32_t n = sd.cnt_peer[sd.rank][r];
  const uint32_t* ids = sd.ids_peer[sd.rank] + (size_t) r * sd.cap;
  for (int64_t j = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t) gridDim.x * blockDim.x)
    sd.contrib[(size_t) sd.idx_of[ids[j]] * sd.world + r] = (uint32_t) j + 1;
}

// ---- C: one summed gradient row per looked-up row -> the owner's value inbox ------------------
// Record sum in the dense kernels' order (k_lazy_update).  lp.tab[].rec = this rank's records.
template <int CPL>
__global__ void __launch_bounds__(256) k_shard_push(const LazyParams lp, ShardDev sd, const uint32_t* __restrict__ item_slot,
                                                    int skip_single) {
  const int lane = threadIdx.x & 31;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  const uint32_t n = __ldg(lp.uniq);
  for (int64_t it = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; it < n; it += nw) {
    const uint32_t* item = lp.uniq + 4 + 3 * it;
    const uint32_t row = __ldg(item);
    const int64_t lo = __ldg(item + 1), L = (int64_t) __ldg(item + 2) - lo;
    if (L == 1 && skip_single) continue;        // k_step delivered the row's only record itself
    const uint32_t o = row / sd.S, slot = __ldg(item_slot + it);
    for (int ti = 0; ti < lp.n_tabs; ++ti) {
      const LazyTable& T = lp.tab[ti];
      const int nch = (T.cols + 3) >> 2;
      float4 g[CPL];
#pragma unroll
      for (int u = 0; u < CPL; ++u) g[u] = f4zero();
      if (L <= UPD_HEAVY) sum_records<CPL>(lp.vals, T.rec, T.rec_ld, 0u, lo, lo + L, nch, lane, g);
      else
        for (int w3 = 0; w3 < 8; ++w3) {
          float4 part[CPL];
#pragma unroll
          for (int u = 0; u < CPL; ++u) part[u] = f4zero();
          sum_records<CPL, RecDepth<CPL>::HEAVY>(lp.vals, T.rec, T.rec_ld, 0u, lo + L * w3 / 8, lo + L * (w3 + 1) / 8, nch, lane, part);
#pragma unroll
          for (int u = 0; u < CPL; ++u) g[u] += part[u];
        }
      store_row<CPL>(sd.val_peer[o][T.slot] + ((size_t) sd.rank * sd.cap + slot) * T.ld, nch, lane, g);
    }
  }
}

// ---- D: owner sums the sources in rank order and applies the step -----------------------------
template <int CPL>
__global__ void __launch_bounds__(256) k_shard_update(const LazyParams lp, ShardDev sd) {
  const int lane = threadIdx.x & 31;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  const uint32_t n = __ldg(lp.uniq);
  const float lr_t = lp.opt == KGE_OPTIM_ADAM ? adam_lr(lp.lr, lp.beta1, lp.beta2, lp.t) : lp.lr;
  for (int64_t it = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; it < n; it += nw) {
    const int64_t row = __ldg(lp.uniq + 4 + 3 * it);
    uint32_t* cb = sd.contrib + (size_t) it * sd.world;
    uint32_t mine = (lane < sd.world) ? cb[lane] : 0u;
    for (int ti = 0; ti < lp.n_tabs; ++ti) {
      const LazyTable& T = lp.tab[ti];
      const int nch = (T.cols + 3) >> 2;
      float4 w[CPL], m[CPL], v[CPL], g[CPL];
      load_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);
      load_row<CPL>(T.m + row * (int64_t) T.ld, nch, lane, m);
      load_row<CPL>(T.v + row * (int64_t) T.ld, nch, lane, v);
      const uint32_t s = __ldg(T.last + row);
#pragma unroll
      for (int u = 0; u < CPL; ++u) g[u] = f4zero();
      const float* val = sd.val_peer[sd.rank][T.slot];
      for (int r = 0; r < sd.world; ++r) {
        const uint32_t c = __shfl_sync(FULL, mine, r);
        if (c == 0) continue;                                  // warp-uniform
        float4 x[CPL];
        load_row<CPL>(val + ((size_t) r * sd.cap + (c - 1)) * T.ld, nch, lane, x);
#pragma unroll
        for (int u = 0; u < CPL; ++u) g[u] += x[u];
      }
      if (lp.opt == KGE_OPTIM_ADAM && s + 1 < lp.t && !all_zero<CPL>(m, v)) decay_mv<CPL>(m, v, lp.t - 1 - s, lp.beta1, lp.beta2);
      const float scale = lp.scal[T.slot];
#pragma unroll
      for (int u = 0; u < CPL; ++u) {
        const int c = lane + 32 * u;
        if (c >= nch) continue;
        adam_chunk(w[u], m[u], v[u], scale4(scale, g[u]), lp.beta1, lp.beta2, lr_t, lp.eps, lp.opt);
        mask_tail(w[u], m[u], v[u], c, T.cols);
      }
      store_row<CPL>(T.m + row * (int64_t) T.ld, nch, lane, m);
      store_row<CPL>(T.v + row * (int64_t) T.ld, nch, lane, v);
      store_row<CPL>(T.w + row * (int64_t) T.ld, nch, lane, w);
      if (lane == 0) T.last[row] = lp.t;
    }
    if (lane < sd.world) cb[lane] = 0u;                        // self-cleaning: ready for the next step
  }
}

template <int WHICH>   // 0 push, 1 update
inline int launch_shard_rows(const LazyParams& lp, const ShardDev& sd, const uint32_t* item_slot, int sm_count, cudaStream_t st,
                             int skip_single = 0) {
  int max_ch = 1;
  for (int j = 0; j < lp.n_tabs; ++j) max_ch = ((lp.tab[j].cols + 3) / 4 > max_ch) ? (lp.tab[j].cols + 3) / 4 : max_ch;
  const int cpl = max_ch <= 32 ? 1 : (max_ch <= 64 ? 2 : (max_ch <= 128 ? 4 : 8));
  const unsigned grid = (unsigned) sm_count * (cpl <= 2 ? 8 : 4);
  if (WHICH == 0) {
    switch (cpl) {
      case 1: k_shard_push<1><<<grid, 256, 0, st>>>(lp, sd, item_slot, skip_single); break;
      case 2: k_shard_push<2><<<grid, 256, 0, st>>>(lp, sd, item_slot, skip_single); break;
      case 4: k_shard_push<4><<<grid, 256, 0, st>>>(lp, sd, item_slot, skip_single); break;
      default: k_shard_push<8><<<grid, 256, 0, st>>>(lp, sd, item_slot, skip_single); break;
    }
  } else {
    switch (cpl) {
      case 1: k_shard_update<1><<<grid, 256, 0, st>>>(lp, sd); break;
      case 2: k_shard_update<2><<<grid, 256, 0, st>>>(lp, sd); break;
      case 4: k_shard_update<4><<<grid, 256, 0, st>>>(lp, sd); break;
      default: k_shard_update<8><<<grid, 256, 0, st>>>(lp, sd); break;
    }
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("shard row kernel launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
  return KGE_OK;
}

}  // namespace kge
