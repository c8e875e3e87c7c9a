// rank_dot_tc.cuh — filtered ranking for dot-product scores (DistMult, RESCAL, any model
// with KGE.score.Dot whose candidate side is a plain entity row) as a tcgen05 GEMM.
//
//   scores[Q x N] = Qm[Q x d] . E^T[d x N]  (+ bias_q),    rank count = #{e : score > pos_q}
//
// The score matrix is never written.  Tensor cores run kind::tf32 (TMA -> 128B-swizzled
// shared memory -> tcgen05.mma -> TMEM).  TF32 keeps 11 significant bits, so both operands are
// split exactly, x = x_hi + x_lo with x_hi = x rounded to TF32, and TWO MMAs per K step
// accumulate q_hi.e_hi + q_lo.e_hi = q.e_hi: the query side is exact, the candidate side is
// e_hi.  What is missing, q.e_lo, is bounded PER CANDIDATE by Cauchy-Schwarz with the
// precomputed ||e_lo|| (k_rank_split_table), the rest (truncation of q_lo to TF32 inside the
// MMA, fp32 accumulation) by eps * ||q|| * max||e||, and the GEMM only DECIDES candidates clear
// of the true score:
//      s_tc >  pos + delta_q + ||q|| ||e_lo||      -> counted by the epilogue     (definite win)
//      s_tc <  pos - delta_q - ||q|| ||e_lo||      -> dropped                     (definite loss)
//      otherwise                                   -> appended to the query's "uncertain" list
// and k_rank_dot_fix re-scores the (few) uncertain candidates in fp32 with the very function
// that produced pos_q (warp_dot_exact), so the final integer counts equal an all-fp32 ranking
// (SURVEY.md H3).  Round 1 ran three MMAs per K step ("3xTF32", q_hi.e_lo as well): the same
// exact counts for 1.5x the tensor work and the e_lo tiles in shared memory.  Known positives are
// a per-query bitmask over the candidate shard
// (k_rank_filter_bits, built from the CSR filter) that the epilogue ANDs out of both the win
// and the uncertain masks.  All counters are integers: results do not depend on scheduling.
//
// Kernel roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2-5 = epilogue (one TMEM lane quarter each).  3-stage smem ring (A_hi, A_lo 128x32 and
// B_hi 256x32 fp32 per stage = 64 KB), two 256-column TMEM accumulators so the epilogue
// of tile i overlaps the MMAs of tile i+1.  K runs over ceil(d / 8) UMMA steps, not the padded
// 32-wide chunks (d = 200: 25 steps, not 28).
#pragma once
#include "tc_ptx.cuh"
#include "rank_kernels.cuh"

namespace kge {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 32, TC_THREADS = 192;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4, TC_B_BYTES = TC_BN * TC_BK * 4;
// PASSES = MMAs per K step.  3 ("3xTF32", the default): q_hi.e_hi + q_lo.e_hi + q_hi.e_lo, stage = A_hi, A_lo, B_hi, B_lo
// (96 KB, 2 stages).  2: q.e_hi only, the missing q.e_lo bounded per candidate by ||q|| ||e_lo|| (Cauchy-Schwarz, loose
// by ~sqrt(d)): two thirds of the tensor work and a 3-stage ring of 64 KB, but 3-4x the uncertain candidates for
// the exact fix-up — it wins only where few candidates score close to the true triple (a trained model);
// kge_set_option(KGE_OPT_RANK_TC_PASSES, 2).  Measured on C2 after 600 steps: 104 vs 31 uncertain per query.
template <int PASSES> struct TcCfg {
  static constexpr int STAGES = PASSES == 3 ? 2 : 3;
  static constexpr int STAGE_BYTES = 2 * TC_A_BYTES + (PASSES == 3 ? 2 : 1) * TC_B_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};
constexpr int TC_USEG = 1024;     // entries an epilogue warp reserves at a time (>= the 32 x 32 pairs of one chunk)
constexpr int TC_UCAP = 512;      // uncertain (query, candidate) pairs kept: TC_UCAP x queries in ONE list (any skew between
                                  // queries is fine); beyond that the block of queries is re-ranked exactly as a whole

// exact split x = hi + lo with hi = x rounded to the nearest TF32 (low 13 mantissa bits clear): |lo| <= 2^-11 |x|
__device__ __forceinline__ float tf32_hi(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r & 0xFFFFE000u);
}

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
  float* nq;                      // ||q|| (slightly inflated): scales the per-candidate bound ||e_lo||
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
    p.nq[q] = nq * 1.0001f;
  }
}

// E_hi of the candidate shard (row e - lo of the output) and ||e_lo|| per candidate (upper bound: computed in
// fp32 from exactly representable residuals, inflated by 1e-4); one warp per row
__global__ void __launch_bounds__(256) k_rank_split_table(const float* w, int ld, int cols, int64_t lo, int64_t hi,
                                                           float* out_hi, float* out_lo, float* nlo) {
  const int lane = threadIdx.x & 31, nch = ld >> 2;
  int64_t row = lo + (((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (; row < hi; row += nw) {
    const float4* src = reinterpret_cast<const float4*>(w + row * (int64_t) ld);
    float4* dst = reinterpret_cast<float4*>(out_hi + (row - lo) * (int64_t) ld);
    float4* dlo = out_lo ? reinterpret_cast<float4*>(out_lo + (row - lo) * (int64_t) ld) : nullptr;
    float ss = 0.f;
    for (int c = lane; c < nch; c += WARP) {
      const float4 v = __ldg(src + c);
      const float4 h = f4(tf32_hi(v.x), tf32_hi(v.y), tf32_hi(v.z), tf32_hi(v.w));
      dst[c] = h;
      const float4 r = v - h;
      if (dlo) dlo[c] = r;
      ss += dot4(r, r);
    }
    ss = warp_sum(ss);
    if (lane == 0) nlo[row - lo] = sqrtf(ss) * 1.0001f;
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
  int n_k8;                       // UMMA K steps of 8 that hold data: ceil(d / 8)
  int64_t Q; int n_qblocks;
  int64_t ent_lo, ent_hi;
  int n_etiles, n_split;          // entity tiles in the shard; work unit = (query block, tile range)
  const float* thr_hi; const float* thr_lo;
  const float* nq;                // per query
  const float* nlo;               // per candidate of the shard (index e - ent_lo), padded to whole tiles
  int32_t* wins;                  // [Q] definite wins (+ the fix-up's hits; k_rank_dot_final adds them to the caller's counts)
  uint32_t* ucnt;                 // [1] uncertain pairs seen (may exceed ucap: overflow)
  uint2* ulist;                   // [ucap] (query, entity id)
  uint32_t ucap;
  int words_per_q;
  const uint32_t* filt;           // nullable: excluded-candidate bits, [Q][words_per_q]
};

template <int PASSES>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_rank_dot_tc(const __grid_constant__ CUtensorMap map_ah, const __grid_constant__ CUtensorMap map_al,
              const __grid_constant__ CUtensorMap map_bh, const __grid_constant__ CUtensorMap map_bl, const DotTcParams p) {
  constexpr int TC_STAGES = TcCfg<PASSES>::STAGES, TC_STAGE_BYTES = TcCfg<PASSES>::STAGE_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
  // stage s: [A_hi | A_lo | B_hi (| B_lo)], every tile 1024-byte aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TC_STAGES * TC_STAGE_BYTES);
  uint64_t* full = bars; uint64_t* empty = bars + TC_STAGES;
  uint64_t* tfull = bars + 2 * TC_STAGES; uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&map_ah); tc::tma_prefetch_desc(&map_al);
    tc::tma_prefetch_desc(&map_bh);
    if (PASSES == 3) tc::tma_prefetch_desc(&map_bl);
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
            if (PASSES == 3) tc::tma_load_2d(st + 2 * TC_A_BYTES + TC_B_BYTES, &map_bl, kc * TC_BK, et * TC_BN, full + s);
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
            const int ksteps = min(TC_BK / 8, p.n_k8 - kc * (TC_BK / 8));     // the last chunk may be partly padding
            for (int k = 0; k < ksteps; ++k) {       // UMMA K = 8 tf32 = 32 bytes = 2 descriptor units
              tc::mma_tf32(td, dah + 2 * k, dbh + 2 * k, idesc, (kc | k) ? 1u : 0u);     // q_hi . e_hi
              tc::mma_tf32(td, dal + 2 * k, dbh + 2 * k, idesc, 1u);                     // q_lo . e_hi
              if (PASSES == 3) tc::mma_tf32(td, dah + 2 * k, dbl + 2 * k, idesc, 1u);    // q_hi . e_lo
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
    uint32_t seg_at = 0, seg_left = 0;               // this warp's open segment of the uncertain-pair list (warp-uniform)
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int qb = u / p.n_split, sp = u % p.n_split;
      const int t0 = sp * tiles_per_split, t1 = min(p.n_etiles, t0 + tiles_per_split);
      const int64_t q = (int64_t) qb * TC_BM + quarter * 32 + lane;
      const bool live = q < p.Q;
      const float hi = live ? p.thr_hi[q] : INFINITY, lo = live ? p.thr_lo[q] : INFINITY, nq = live ? p.nq[q] : 0.f;
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
          const float* nl = p.nlo + (e0 - p.ent_lo);               // ||e_lo|| of the chunk's candidates (same for every lane)
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float s = __uint_as_float(v[j]);
            const float band = PASSES == 3 ? 0.f : nq * __ldg(nl + j);   // 2 passes: |q . e_lo| <= ||q|| ||e_lo||
            win |= (s > hi + band) ? (1u << j) : 0u;
            unc |= (s >= lo - band) ? (1u << j) : 0u;
          }
          const int64_t left = p.ent_hi - e0;                      // candidates left in the shard
          uint32_t ok = left >= 32 ? 0xFFFFFFFFu : (left > 0 ? ((1u << left) - 1u) : 0u);
          const int64_t widx = q * (int64_t) p.words_per_q + et * (TC_BN / 32) + c;
          if (live && p.filt) ok &= ~__ldg(p.filt + widx);         // known positives never count
          win &= ok;
          cnt += __popc(win);
          uint32_t u = live ? (unc & ~win & ok) : 0u;
          const int n = __popc(u);
          if (__any_sync(FULL, n != 0)) {
            // Append to the global list through a warp-private SEGMENT of TC_USEG entries: the list's counter is
            // touched once per segment, not once per chunk (an atomic round trip per chunk made the epilogue,
            // not the MMAs, the critical path).  Unused entries of a closed segment hold a sentinel query id.
            int incl = n;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
            const uint32_t total = (uint32_t) __shfl_sync(FULL, incl, 31);
            if (total > seg_left) {
              for (uint32_t i2 = lane; i2 < seg_left; i2 += WARP)
                if (seg_at + i2 < p.ucap) p.ulist[seg_at + i2] = make_uint2(0xffffffffu, 0u);
              uint32_t base = 0;
              if (lane == 0) base = atomicAdd(p.ucnt, (uint32_t) TC_USEG);
              seg_at = __shfl_sync(FULL, base, 0);
              seg_left = TC_USEG;
            }
            uint32_t at = seg_at + (uint32_t) (incl - n);
            while (u) {
              const int b = __ffs(u) - 1;
              u &= u - 1;
              if (at < p.ucap) p.ulist[at] = make_uint2((uint32_t) q, (uint32_t) (e0 + b));
              ++at;
            }
            seg_at += total; seg_left -= total;
          }
        }
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(tempty + acc);
      }
      if (live && cnt) atomicAdd(p.wins + q, cnt);
    }
    for (uint32_t i2 = lane; i2 < seg_left; i2 += WARP)             // close the last segment
      if (seg_at + i2 < p.ucap) p.ulist[seg_at + i2] = make_uint2(0xffffffffu, 0u);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 512);
}

struct DotFixParams {
  const float* Qm; int kpad;
  const float* E; int ldE, cols;
  const float* pos; const float* bias;
  int32_t* wins; const uint32_t* ucnt; const uint2* ulist; uint32_t ucap;
  const uint32_t* filt; int words_per_q;
  int64_t Q, ent_lo, ent_hi;
  int32_t* count;
};

// uncertain pairs -> exact fp32 score, hits join the query's wins.  A warp takes FIX_U consecutive pairs per round
// and keeps their row loads in flight together (one pair per round left the kernel waiting on two dependent L2 round
// trips per pair); each pair's score is warp_dot_exact's arithmetic — the same fmaf chain per lane in ascending
// chunk order, the same butterfly — so the counts are unchanged.
constexpr int FIX_U = 4;
__global__ void __launch_bounds__(256) k_rank_dot_fix(const DotFixParams p) {
  const int lane = threadIdx.x & 31;
  const uint32_t n = *p.ucnt;
  if (n > p.ucap) return;                                   // overflow: k_rank_dot_final re-ranks the block exactly
  const int nch = (p.cols + 3) >> 2;
  const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
  // (measured and dropped: a CTA walking a contiguous run of the list so that its segment's 32 query rows come from
  // L1 — 24.9 vs 25.2 M queries/s: the sentinel tails of the segments unbalance the CTAs)
  for (uint32_t i0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * FIX_U; i0 < n; i0 += nw * FIX_U) {
    uint32_t q[FIX_U];
    bool ok[FIX_U];
    const float4* a[FIX_U];
    const float4* b[FIX_U];
    float s[FIX_U];
#pragma unroll
    for (int u = 0; u < FIX_U; ++u) {
      const bool in = i0 + u < n;
      const uint2 qe = in ? p.ulist[i0 + u] : make_uint2(0xffffffffu, 0u);
      ok[u] = qe.x != 0xffffffffu;                           // unused tail of a segment / past the end
      q[u] = ok[u] ? qe.x : 0u;
      a[u] = reinterpret_cast<const float4*>(p.Qm + q[u] * (int64_t) p.kpad);
      b[u] = reinterpret_cast<const float4*>(p.E + (ok[u] ? (int64_t) qe.y : p.ent_lo) * (int64_t) p.ldE);
      s[u] = 0.f;
    }
    for (int c = lane; c < nch; c += WARP) {
      float4 x[FIX_U], y[FIX_U];
#pragma unroll
      for (int u = 0; u < FIX_U; ++u) { x[u] = a[u][c]; y[u] = __ldg(b[u] + c); }
#pragma unroll
      for (int u = 0; u < FIX_U; ++u) {
        s[u] = fmaf(x[u].x, y[u].x, s[u]); s[u] = fmaf(x[u].y, y[u].y, s[u]);
        s[u] = fmaf(x[u].z, y[u].z, s[u]); s[u] = fmaf(x[u].w, y[u].w, s[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < FIX_U; ++u) s[u] = warp_sum(s[u]);
    if (lane == 0) {
#pragma unroll
      for (int u = 0; u < FIX_U; ++u)
        if (ok[u] && s[u] + p.bias[q[u]] > p.pos[q[u]]) atomicAdd(p.wins + q[u], 1);
    }
  }
}
// count_out += wins; or, when the list overflowed (degenerate score distributions: e.g. an all-equal table), the
// exact count over the whole shard, one CTA per query — the tensor-core tallies are then ignored
constexpr int FIX_WARPS = 8;
__global__ void __launch_bounds__(FIX_WARPS * 32) k_rank_dot_final(const DotFixParams p) {
  __shared__ int s_cnt[FIX_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (*p.ucnt <= p.ucap) {
    const int64_t q = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (q < p.Q && p.wins[q]) atomicAdd(p.count + q, p.wins[q]);
    return;
  }
  for (int64_t q = blockIdx.x; q < p.Q; q += gridDim.x) {
    const float* qm = p.Qm + q * (int64_t) p.kpad;
    const float pos = p.pos[q], bias = p.bias[q];
    const int nch = (p.cols + 3) >> 2;
    int cnt = 0;
    for (int64_t e = p.ent_lo + warp; e < p.ent_hi; e += FIX_WARPS) {
      const int64_t off = e - p.ent_lo;
      if (p.filt && ((__ldg(p.filt + q * (int64_t) p.words_per_q + (off >> 5)) >> (off & 31)) & 1u)) continue;
      const float s = warp_dot_exact(qm, p.E + e * (int64_t) p.ldE, nch, lane) + bias;
      cnt += (s > pos) ? 1 : 0;
    }
    __syncthreads();
    if (lane == 0) s_cnt[warp] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
      for (int w2 = 0; w2 < FIX_WARPS; ++w2) tot += s_cnt[w2];
      if (tot) atomicAdd(p.count + q, tot);
    }
  }
}

}  // namespace kge
