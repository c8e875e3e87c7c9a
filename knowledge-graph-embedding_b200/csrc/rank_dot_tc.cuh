// rank_dot_tc.cuh — filtered ranking for dot-product scores (DistMult, RESCAL, any model
// with KGE.score.Dot whose candidate side is a plain entity row) as a tcgen05 GEMM.
//
//   scores[Q x N] = Qm[Q x d] . E^T[d x N]  (+ bias_q),    rank count = #{e : score > pos_q}
//
// The score matrix is never written.  Tensor cores run kind::tf32 straight off the fp32
// tables (TMA -> 128B-swizzled shared memory -> tcgen05.mma -> TMEM); TF32 drops 13 mantissa
// bits, so the GEMM only DECIDES candidates that are clear of the true score by more than a
// rigorous error bound delta_q = eps * ||q|| * max||e||:
//      s_tc >  pos + delta            -> counted by the epilogue            (definite win)
//      s_tc <  pos - delta            -> dropped                            (definite loss)
//      otherwise                      -> one bit in an "uncertain" mask
// and k_rank_fix re-scores the uncertain candidates in fp32 with the very function that
// produced pos_q (warp_dot_exact), so the final integer counts equal an all-fp32 ranking
// (SURVEY.md H3).  Known positives are then subtracted by k_rank_filter (CSR subtraction,
// SURVEY.md App. F).  All counters are integers: results do not depend on scheduling.
//
// Kernel roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2-5 = epilogue (one TMEM lane quarter each).  4-stage smem ring (A 128x32 + B 256x32
// fp32 per stage), two 256-column TMEM accumulators so the epilogue of tile i overlaps the
// MMAs of tile i+1.
#pragma once
#include "tc_ptx.cuh"
#include "rank_kernels.cuh"

namespace kge {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 32, TC_STAGES = 4, TC_THREADS = 192;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4, TC_B_BYTES = TC_BN * TC_BK * 4;
constexpr int TC_SMEM_BYTES = TC_STAGES * (TC_A_BYTES + TC_B_BYTES) + 1024 /*align*/ + 256 /*barriers*/;

// fp32 dot of two rows with a fixed evaluation order: THE reference arithmetic of this path
// (pos_q, the re-scored uncertain candidates and the filtered candidates all come from it).
__device__ __forceinline__ float warp_dot_exact(const float* __restrict__ a, const float* __restrict__ b, int nch, int lane) {
  float s = 0.f;
  for (int c = lane; c < nch; c += WARP) {
    // `a` may have been written by this kernel (k_rank_dot_prep): plain load, not the read-only path
    const float4 x = reinterpret_cast<const float4*>(a)[c], y = __ldg(reinterpret_cast<const float4*>(b) + c);
    s = fmaf(x.x, y.x, s); s = fmaf(x.y, y.y, s); s = fmaf(x.z, y.z, s); s = fmaf(x.w, y.w, s);
  }
  return warp_sum(s);
}

struct DotPrepParams {
  ModelDev md;
  const void* queries; int id64; int64_t Q; int side;
  float* Qm; int kpad;            // [Qpad x kpad], zero padded
  float* pos;                     // exact score of the true triple
  float* bias;                    // per-query additive term (TransE head side: r . t)
  float* thr_hi; float* thr_lo;   // thresholds on the raw tensor-core dot
  const unsigned* maxnorm_bits;   // max_e ||e|| as float bits
  float eps;
};

// max row norm of the candidate shard (non-negative floats order like their bit patterns)
__global__ void __launch_bounds__(256) k_rank_maxnorm(const float* w, int ld, int cols, int64_t lo, int64_t hi, unsigned* out) {
  const int lane = threadIdx.x & 31;
  int64_t row = lo + (((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  float best = 0.f;
  for (; row < hi; row += nw) {
    const float* r = w + row * (int64_t) ld;
    best = fmaxf(best, sqrtf(warp_dot_exact(r, r, (cols + 3) >> 2, lane)));
  }
  if (lane == 0 && best > 0.f) atomicMax(out, __float_as_uint(best));
}

// one warp per query: query-side vector, exact true score, thresholds
__global__ void __launch_bounds__(256) k_rank_dot_prep(const DotPrepParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t q = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (q >= p.Q) return;
  const ModelDev& md = p.md;
  const int64_t h = load_id(p.queries, 3 * q + 0, p.id64), r = load_id(p.queries, 3 * q + 1, p.id64),
                t = load_id(p.queries, 3 * q + 2, p.id64);
  const bool head = (p.side == KGE_SIDE_H);           // candidates replace the head
  const int d = md.d_ent;
  const float* E = md.t[KGE_T_ENT].w;
  const int ldE = md.t[KGE_T_ENT].ld;
  const float* keep = E + (head ? t : h) * (int64_t) ldE;         // the entity that stays
  const float* truth = E + (head ? h : t) * (int64_t) ldE;        // the true candidate
  float* qm = p.Qm + q * (int64_t) p.kpad;
  float bias = 0.f;
  if (md.model == KGE_RESCAL) {
    const float* Rm = md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld;
    for (int j = lane; j < d; j += WARP) {
      float a = 0.f;
      if (head) for (int i = 0; i < d; ++i) a = fmaf(Rm[j * d + i], keep[i], a);      // (R t)_j
      else for (int i = 0; i < d; ++i) a = fmaf(keep[i], Rm[i * d + j], a);            // (h^T R)_j
      qm[j] = a;
    }
  } else {
    const float* rel = (md.model == KGE_UM) ? nullptr : md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld;
    float bpart = 0.f;
    for (int j = lane; j < d; j += WARP) {
      float v = keep[j];
      if (md.model == KGE_DISTMULT) v = head ? rel[j] * v : v * rel[j];               // r o t | h o r
      else if (md.model == KGE_TRANSE) {
        if (head) bpart = fmaf(rel[j], v, bpart);                                      // (e + r).t = e.t + r.t
        else v = v + rel[j];                                                           // (h + r).e
      }
      qm[j] = v;
    }
    bias = warp_sum(bpart);
  }
  __syncwarp();
  const int nch = (d + 3) >> 2;
  const float nq = sqrtf(warp_dot_exact(qm, qm, nch, lane));
  const float pos = warp_dot_exact(qm, truth, nch, lane) + bias;
  if (lane == 0) {
    const float delta = p.eps * nq * __uint_as_float(*p.maxnorm_bits);
    p.pos[q] = pos; p.bias[q] = bias;
    p.thr_hi[q] = (pos - bias) + delta;
    p.thr_lo[q] = (pos - bias) - delta;
  }
}

struct DotTcParams {
  int n_kc;                       // K chunks of 32
  int64_t Q; int n_qblocks;
  int64_t ent_lo, ent_hi;
  int n_etiles, n_split;          // entity tiles in the shard; work unit = (query block, tile range)
  const float* thr_hi; const float* thr_lo;
  int32_t* count;
  uint32_t* mask; int words_per_q;
};

__global__ void __launch_bounds__(TC_THREADS, 1)
k_rank_dot_tc(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const DotTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + TC_STAGES * TC_A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TC_STAGES * (TC_A_BYTES + TC_B_BYTES));
  uint64_t* full = bars; uint64_t* empty = bars + TC_STAGES;
  uint64_t* tfull = bars + 2 * TC_STAGES; uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&map_a); tc::tma_prefetch_desc(&map_b);
    for (int s = 0; s < TC_STAGES; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
    for (int a = 0; a < 2; ++a) { tc::mbar_init(tfull + a, 1); tc::mbar_init(tempty + a, 4); }
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int n_units = p.n_qblocks * p.n_split;
  const int tiles_per_split = (p.n_etiles + p.n_split - 1) / p.n_split;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int qb = u / p.n_split, sp = u % p.n_split;
        const int t0 = sp * tiles_per_split, t1 = min(p.n_etiles, t0 + tiles_per_split);
        for (int et = t0; et < t1; ++et)
          for (int kc = 0; kc < p.n_kc; ++kc, ++it) {
            const int s = it % TC_STAGES; const uint32_t ph = (it / TC_STAGES) & 1;
            tc::mbar_wait(empty + s, ph ^ 1);
            tc::mbar_expect_tx(full + s, TC_A_BYTES + TC_B_BYTES);
            tc::tma_load_2d(sA + s * TC_A_BYTES, &map_a, kc * TC_BK, qb * TC_BM, full + s);
            tc::tma_load_2d(sB + s * TC_B_BYTES, &map_b, kc * TC_BK, (int) (p.ent_lo + (int64_t) et * TC_BN), full + s);
          }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = tc::idesc_tf32(TC_BM, TC_BN);
      uint32_t it = 0, tile = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int sp = u % p.n_split;
        const int t0 = sp * tiles_per_split, t1 = min(p.n_etiles, t0 + tiles_per_split);
        for (int et = t0; et < t1; ++et, ++tile) {
          const uint32_t acc = tile & 1, aph = (tile >> 1) & 1;
          tc::mbar_wait(tempty + acc, aph ^ 1);
          tc::tc_fence_after();
          for (int kc = 0; kc < p.n_kc; ++kc, ++it) {
            const int s = it % TC_STAGES; const uint32_t ph = (it / TC_STAGES) & 1;
            tc::mbar_wait(full + s, ph);
            tc::tc_fence_after();
            const uint64_t da = tc::smem_desc_k_sw128(tc::smem_u32(sA + s * TC_A_BYTES));
            const uint64_t db = tc::smem_desc_k_sw128(tc::smem_u32(sB + s * TC_B_BYTES));
#pragma unroll
            for (int k = 0; k < TC_BK / 8; ++k)      // UMMA K = 8 tf32 = 32 bytes = 2 descriptor units
              tc::mma_tf32(tmem_base + acc * TC_BN, da + 2 * k, db + 2 * k, idesc, (kc | k) ? 1u : 0u);
            tc::mma_commit(empty + s);               // frees the smem stage when these MMAs retire
          }
          tc::mma_commit(tfull + acc);
        }
      }
    }
  } else {
    const int quarter = warp & 3;                    // TMEM lanes [32*quarter, +32)
    uint32_t tile = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int qb = u / p.n_split, sp = u % p.n_split;
      const int t0 = sp * tiles_per_split, t1 = min(p.n_etiles, t0 + tiles_per_split);
      const int64_t q = (int64_t) qb * TC_BM + quarter * 32 + lane;
      const bool live = q < p.Q;
      const float hi = live ? p.thr_hi[q] : INFINITY, lo = live ? p.thr_lo[q] : INFINITY;
      int cnt = 0;
      for (int et = t0; et < t1; ++et, ++tile) {
        const uint32_t acc = tile & 1, aph = (tile >> 1) & 1;
        tc::mbar_wait(tfull + acc, aph);
        tc::tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < TC_BN / 32; ++c) {
          uint32_t v[32];
          tc::tmem_ld_32x32(tmem_base + ((uint32_t) (quarter * 32) << 16) + acc * TC_BN + c * 32, v);
          tc::tmem_ld_wait();
          const int64_t e0 = p.ent_lo + (int64_t) et * TC_BN + c * 32;
          uint32_t m = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float s = __uint_as_float(v[j]);
            const bool valid = e0 + j < p.ent_hi;
            cnt += (valid && s > hi) ? 1 : 0;
            m |= (valid && !(s > hi) && s >= lo) ? (1u << j) : 0u;
          }
          if (live) p.mask[q * (int64_t) p.words_per_q + et * (TC_BN / 32) + c] = m;
        }
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(tempty + acc);
      }
      if (live && cnt) atomicAdd(p.count + q, cnt);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 512);
}

struct DotFixParams {
  const float* Qm; int kpad;
  const float* E; int ldE, cols;
  const float* pos; const float* bias;
  const uint32_t* mask; int words_per_q;
  int64_t Q, ent_lo;
  int32_t* count;
  const int64_t* f_off; const int64_t* f_ids;     // filter (nullable), ids unique per query
  int64_t ent_hi;
};

// uncertain candidates -> exact fp32 score; one warp per (query, 32 mask words)
__global__ void __launch_bounds__(256) k_rank_dot_fix(const DotFixParams p) {
  const int lane = threadIdx.x & 31;
  const int groups = (p.words_per_q + 31) / 32;
  const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (gw >= p.Q * groups) return;
  const int64_t q = gw / groups;
  const int g = (int) (gw % groups);
  const int widx = g * 32 + lane;
  const uint32_t mine = (widx < p.words_per_q) ? __ldg(p.mask + q * (int64_t) p.words_per_q + widx) : 0u;
  unsigned any = __ballot_sync(FULL, mine != 0u);
  if (!any) return;
  const float* qm = p.Qm + q * (int64_t) p.kpad;
  const float pos = p.pos[q], bias = p.bias[q];
  const int nch = (p.cols + 3) >> 2;
  int cnt = 0;
  while (any) {
    const int src = __ffs(any) - 1;
    any &= any - 1;
    uint32_t w = __shfl_sync(FULL, mine, src);
    const int64_t base = p.ent_lo + (int64_t) (g * 32 + src) * 32;
    while (w) {
      const int b = __ffs(w) - 1;
      w &= w - 1;
      const float s = warp_dot_exact(qm, p.E + (base + b) * (int64_t) p.ldE, nch, lane) + bias;
      cnt += (s > pos) ? 1 : 0;
    }
  }
  if (lane == 0 && cnt) atomicAdd(p.count + q, cnt);
}

// known positives: subtract those that out-score the true triple (they were counted above)
__global__ void __launch_bounds__(256) k_rank_dot_filter(const DotFixParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t q = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (q >= p.Q) return;
  const float* qm = p.Qm + q * (int64_t) p.kpad;
  const float pos = p.pos[q], bias = p.bias[q];
  const int nch = (p.cols + 3) >> 2;
  int cnt = 0;
  for (int64_t f = p.f_off[q]; f < p.f_off[q + 1]; ++f) {
    const int64_t e = p.f_ids[f];
    if (e < p.ent_lo || e >= p.ent_hi) continue;
    const float s = warp_dot_exact(qm, p.E + e * (int64_t) p.ldE, nch, lane) + bias;
    cnt += (s > pos) ? 1 : 0;
  }
  if (lane == 0 && cnt) atomicSub(p.count + q, cnt);
}

}  // namespace kge
