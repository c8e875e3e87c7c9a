// rank_dot_tc.cuh — filtered ranking for dot-product scores (DistMult, RESCAL, any model
// with KGE.score.Dot whose candidate side is a plain entity row) as a tcgen05 GEMM.
//
//   scores[Q x N] = Qm[Q x d] . E^T[d x N]  (+ bias_q),    rank count = #{e : score > pos_q}
//
// The score matrix is never written.  Tensor cores run kind::tf32 (TMA -> 128B-swizzled
// shared memory -> tcgen05.mma -> TMEM).  TF32 keeps 11 significant bits, so both operands are
// split exactly, x = x_hi + x_lo with x_hi TF32-representable, and three MMAs per K step
// accumulate q_hi.e_hi + q_lo.e_hi + q_hi.e_lo ("3xTF32").  What is dropped (q_lo.e_lo, the
// truncation of the lo parts, fp32 accumulation) is bounded rigorously and the GEMM only
// DECIDES candidates clear of the true score by more than delta_q = eps * ||q|| * max||e||:
//      s_tc >  pos + delta            -> counted by the epilogue            (definite win)
//      s_tc <  pos - delta            -> dropped                            (definite loss)
//      otherwise                      -> one bit in an "uncertain" mask
// and k_rank_dot_fix re-scores the (few) uncertain candidates in fp32 with the very function
// that produced pos_q (warp_dot_exact), so the final integer counts equal an all-fp32 ranking
// (SURVEY.md H3).  Known positives are a per-query bitmask over the candidate shard
// (k_rank_filter_bits, built from the CSR filter) that the epilogue ANDs out of both the win
// and the uncertain masks.  All counters are integers: results do not depend on scheduling.
//
// Kernel roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2-5 = epilogue (one TMEM lane quarter each).  2-stage smem ring (A_hi, A_lo 128x32 and
// B_hi, B_lo 256x32 fp32 per stage = 96 KB), two 256-column TMEM accumulators so the epilogue
// of tile i overlaps the MMAs of tile i+1.
#pragma once
#include "tc_ptx.cuh"
#include "rank_kernels.cuh"

namespace kge {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 32, TC_STAGES = 2, TC_THREADS = 192;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4, TC_B_BYTES = TC_BN * TC_BK * 4;
constexpr int TC_STAGE_BYTES = 2 * TC_A_BYTES + 2 * TC_B_BYTES;          // hi and lo of both operands
constexpr int TC_SMEM_BYTES = TC_STAGES * TC_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;

// exact split x = hi + lo with hi representable in TF32 (low 13 mantissa bits clear)
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

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
  float* Qhi; float* Qlo;         // exact TF32 split of Qm
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
  for (int j = lane; j < d; j += WARP) {
    const float v = qm[j], hi = tf32_hi(v);
    p.Qhi[q * (int64_t) p.kpad + j] = hi;
    p.Qlo[q * (int64_t) p.kpad + j] = v - hi;
  }
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

// E_hi / E_lo of the candidate shard (row e - lo of the outputs)
__global__ void __launch_bounds__(256) k_rank_split_table(const float* w, int ld, int cols, int64_t lo, int64_t hi,
                                                           float* out_hi, float* out_lo) {
  const int nch = ld >> 2;
  const int64_t total = (hi - lo) * nch;
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(w + lo * (int64_t) ld) + i);
    const float4 h = f4(tf32_hi(v.x), tf32_hi(v.y), tf32_hi(v.z), tf32_hi(v.w));
    reinterpret_cast<float4*>(out_hi)[i] = h;
    reinterpret_cast<float4*>(out_lo)[i] = v - h;
  }
}

// CSR filter -> per-query bitmask over the candidate shard (bit = excluded)
__global__ void __launch_bounds__(256) k_rank_filter_bits(const int64_t* f_off, const int64_t* f_end, const int64_t* f_ids, int64_t Q,
                                                          int64_t ent_lo, int64_t ent_hi, uint32_t* bits, int words_per_q) {
  const int lane = threadIdx.x & 31;
  const int64_t q = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (q >= Q) return;
  for (int64_t f = f_off[q] + lane; f < f_end[q]; f += WARP) {
    const int64_t e = f_ids[f];
    if (e >= ent_lo && e < ent_hi) atomicOr(bits + q * (int64_t) words_per_q + ((e - ent_lo) >> 5), 1u << ((e - ent_lo) & 31));
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
  const uint32_t* filt;           // nullable: excluded-candidate bits, same indexing as mask
};

__global__ void __launch_bounds__(TC_THREADS, 1)
k_rank_dot_tc(const __grid_constant__ CUtensorMap map_ah, const __grid_constant__ CUtensorMap map_al,
              const __grid_constant__ CUtensorMap map_bh, const __grid_constant__ CUtensorMap map_bl, const DotTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
  // stage s: [A_hi | A_lo | B_hi | B_lo], every tile 1024-byte aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TC_STAGES * TC_STAGE_BYTES);
  uint64_t* full = bars; uint64_t* empty = bars + TC_STAGES;
  uint64_t* tfull = bars + 2 * TC_STAGES; uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&map_ah); tc::tma_prefetch_desc(&map_al);
    tc::tma_prefetch_desc(&map_bh); tc::tma_prefetch_desc(&map_bl);
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
            uint8_t* st = smem + s * TC_STAGE_BYTES;
            tc::mbar_expect_tx(full + s, TC_STAGE_BYTES);
            tc::tma_load_2d(st, &map_ah, kc * TC_BK, qb * TC_BM, full + s);
            tc::tma_load_2d(st + TC_A_BYTES, &map_al, kc * TC_BK, qb * TC_BM, full + s);
            tc::tma_load_2d(st + 2 * TC_A_BYTES, &map_bh, kc * TC_BK, et * TC_BN, full + s);        // shard-relative rows
            tc::tma_load_2d(st + 2 * TC_A_BYTES + TC_B_BYTES, &map_bl, kc * TC_BK, et * TC_BN, full + s);
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
            const uint32_t sbase = tc::smem_u32(smem + s * TC_STAGE_BYTES);
            const uint64_t dah = tc::smem_desc_k_sw128(sbase), dal = tc::smem_desc_k_sw128(sbase + TC_A_BYTES);
            const uint64_t dbh = tc::smem_desc_k_sw128(sbase + 2 * TC_A_BYTES);
            const uint64_t dbl = tc::smem_desc_k_sw128(sbase + 2 * TC_A_BYTES + TC_B_BYTES);
            const uint32_t td = tmem_base + acc * TC_BN;
#pragma unroll
            for (int k = 0; k < TC_BK / 8; ++k) {    // UMMA K = 8 tf32 = 32 bytes = 2 descriptor units
              tc::mma_tf32(td, dah + 2 * k, dbh + 2 * k, idesc, (kc | k) ? 1u : 0u);     // q_hi . e_hi
              tc::mma_tf32(td, dal + 2 * k, dbh + 2 * k, idesc, 1u);                     // q_lo . e_hi
              tc::mma_tf32(td, dah + 2 * k, dbl + 2 * k, idesc, 1u);                     // q_hi . e_lo
            }
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
          uint32_t win = 0, unc = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float s = __uint_as_float(v[j]);
            win |= (s > hi) ? (1u << j) : 0u;
            unc |= (s >= lo) ? (1u << j) : 0u;
          }
          const int64_t left = p.ent_hi - e0;                      // candidates left in the shard
          uint32_t ok = left >= 32 ? 0xFFFFFFFFu : (left > 0 ? ((1u << left) - 1u) : 0u);
          const int64_t widx = q * (int64_t) p.words_per_q + et * (TC_BN / 32) + c;
          if (live && p.filt) ok &= ~__ldg(p.filt + widx);         // known positives never count
          win &= ok;
          cnt += __popc(win);
          if (live) p.mask[widx] = unc & ~win & ok;
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

}  // namespace kge
