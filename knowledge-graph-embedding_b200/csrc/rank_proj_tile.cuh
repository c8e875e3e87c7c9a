// rank_proj_tile.cuh — entity-tiled Lp ranking for the models whose candidate side depends on the RELATION:
// TransH (hyperplane projection, TransH.py:173-185), TransD (rank-1 projection + clip, TransD.py:194-222),
// TransR and SE (per-relation matrices, TransR.py:178-191, SE.py:155-161) with LpDistance / LpDistancePow.
//
// for a fixed relation r every one of these scores is  -g( sum_j |x_j - y_j|^p )  with
//      x = f_h(h, r) + r_vec   (head-side vector)        y = f_t(t, r)   (tail-side vector)
// so the candidates of all queries that share r are ONE projected table  P_r[e] = y(e, r)  (tail candidates) or
// x(e, r) (head candidates), and ranking against it is exactly the TransE tile kernel (rank_lp_tile.cuh).  Per call:
//   1  queries are grouped by relation on the device (histogram / scan / scatter: no host synchronisation),
//   2  relations are taken a chunk at a time (the chunk's projected tables fit a fixed workspace budget):
//        k_proj_table    one warp per (relation, candidate): the MODEL'S OWN forward function writes the row
//                        (relations no query asks for are skipped),
//        k_proj_queries  one warp per query: its vector by the same function, and the threshold S(true) from
//                        the same function applied to the true entity — bit-identical to the table's row,
//        k_rank_lp_tile_g  the 128 x 128 register-tiled distance kernel per (relation, query tile, entity tile),
//        k_rank_lp_filter_g  known positives taken back out.
// Every S is accumulated by one thread in ascending dimension with lp_acc: counts are exact for this arithmetic,
// and the true entity never beats itself.  Round 1 scored these models one warp per (query, candidate) with a
// full projection per pair (TransD 233 K, TransR 29 K queries/s).
#pragma once
#include <stdlib.h>
#include "host_common.cuh"
#include "rank_lp_tile.cuh"
#include "rank_mat_kernels.cuh"

namespace kge {

struct ProjParams {
  ModelDev md;
  int side;                          // candidate side (KGE_SIDE_H / KGE_SIDE_T)
  const void* queries; int id64; int64_t Q;
  const uint32_t* goff;              // [n_rel + 1] first sorted-query position of every relation
  const uint32_t* perm;              // [Q] sorted position -> query index
  const uint32_t* toff;              // [n_rel + 1] first 128-query tile of every relation (sum <= Q / 128 + n_rel)
  const uint32_t* tile_rel;          // [tiles] tile -> relation
  int rel0, n_rel_chunk;             // relations [rel0, rel0 + n_rel_chunk) are processed by this launch
  int64_t ent_lo, ent_hi; int64_t span_pad;      // candidate shard; rows per projected table (span rounded up to 128)
  float* P; int ldp; int dout;       // [n_rel_chunk][span_pad][ldp]
  float* Qv; float* Tv; float* thr;  // per SORTED query: its vector, the true candidate's row (scratch), clipped S(true)
  const int64_t* f_off; const int64_t* f_end; const int64_t* f_ids;
  int32_t* count;
  float p;
};

// ---- grouping of the queries by relation -----------------------------------------------------------------
static __global__ void __launch_bounds__(256) k_qrel_hist(const void* queries, int id64, int64_t Q, uint32_t* cnt) {
  const int64_t q = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (q < Q) atomicAdd(cnt + load_id(queries, 3 * q + 1, id64), 1u);
}
// exclusive scan of cnt[0..R) -> goff[0..R], one CTA; also clears cnt for the scatter's fill counters.
// TILES: scans ceil(cnt / 128) instead (the relation's query tiles) and leaves cnt alone.
template <bool TILES>
static __global__ void __launch_bounds__(1024) k_qrel_scan(uint32_t* cnt, uint32_t* goff, int64_t R) {
  __shared__ uint32_t s_w[32];
  __shared__ uint32_t s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < R; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const uint32_t v = i < R ? (TILES ? (cnt[i] + LP_T - 1) / LP_T : cnt[i]) : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      const uint32_t w = s_w[lane];
      uint32_t winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, winc, o); if (lane >= o) winc += t; }
      s_w[lane] = winc - w;
    }
    __syncthreads();
    const uint32_t ex = s_carry + s_w[warp] + inc - v;
    if (i < R) { goff[i] = ex; if (!TILES) cnt[i] = 0u; }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = ex + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) goff[R] = s_carry;
}
static __global__ void __launch_bounds__(256) k_qrel_scatter(const void* queries, int id64, int64_t Q, const uint32_t* goff,
                                                             uint32_t* fill, uint32_t* perm) {
  const int64_t q = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  const int64_t r = load_id(queries, 3 * q + 1, id64);
  perm[goff[r] + atomicAdd(fill + r, 1u)] = (uint32_t) q;      // the order inside a relation's group does not matter
}

// tile -> relation list: one thread per relation writes its (few) tiles
static __global__ void __launch_bounds__(256) k_qrel_tiles(const uint32_t* toff, int64_t R, uint32_t* tile_rel) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  for (uint32_t t = toff[r]; t < toff[r + 1]; ++t) tile_rel[t] = (uint32_t) r;
}

// ---- rows by the model's own forward function ---------------------------------------------------------------
// vector models (TransH, TransD): forward(H = e, R, T = e) yields x(e, r) and y(e, r) at once
template <class M, int NV>
__device__ __forceinline__ void proj_vec(const ModelDev& md, int64_t r, int64_t e, bool want_x, int lane, Row<NV>& out) {
  typename M::template Ent<NV> E;
  typename M::template Rel<NV> R;
  E.load(md, e, lane);
  R.load(md, r, lane);
  Row<NV> x, y;
  FwdSaved<NV> sv;
  M::template forward<NV>(md, E, R, E, lane, x, y, sv);
  out = want_x ? x : y;
}
template <class M, int NV>
__global__ void __launch_bounds__(256) k_proj_table(const ProjParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t span = p.ent_hi - p.ent_lo;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (int64_t w = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < span * p.n_rel_chunk; w += nw) {
    const int rl = (int) (w / span);
    const int64_t el = w - (int64_t) rl * span;
    const int64_t r = p.rel0 + rl;
    if (p.goff[r + 1] == p.goff[r]) continue;                  // no query asks for this relation
    Row<NV> row;
    proj_vec<M, NV>(p.md, r, p.ent_lo + el, p.side == KGE_SIDE_H, lane, row);
    store_row(row, p.P + ((int64_t) rl * p.span_pad) * p.ldp, el, p.ldp, p.dout, lane);
  }
}
// S(a, b) over d dims by one thread, the tile kernel's arithmetic
template <int KIND>
__device__ __forceinline__ float lp_row_sum(const float* a, const float* b, int d, float pw) {
  float s = 0.f;
  for (int j = 0; j < d; ++j) s = lp_acc<KIND>(s, a[j], b[j], pw);
  return lp_clip(s, KIND);
}
template <class M, int NV, int KIND>
__global__ void __launch_bounds__(256) k_proj_queries(const ProjParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t lo = p.goff[p.rel0], hi = p.goff[p.rel0 + p.n_rel_chunk];
  const int64_t i = lo + (((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5);       // sorted position
  if (i >= hi) return;
  const int64_t q = p.perm[i];
  const int64_t h = load_id(p.queries, 3 * q + 0, p.id64), r = load_id(p.queries, 3 * q + 1, p.id64),
                t = load_id(p.queries, 3 * q + 2, p.id64);
  const bool head = p.side == KGE_SIDE_H;
  Row<NV> qv, tv;
  proj_vec<M, NV>(p.md, r, head ? t : h, !head, lane, qv);      // head candidates: query = y(t); tail: x(h)
  proj_vec<M, NV>(p.md, r, head ? h : t, head, lane, tv);       // the true candidate's own table row
  store_row(qv, p.Qv, i, p.ldp, p.dout, lane);
  store_row(tv, p.Tv, i, p.ldp, p.dout, lane);
  __syncwarp();
  // head candidates: S = sum |x_e - y_q| (candidate first);  tail: sum |x_q - y_e| (query first): lp_acc(acc, x, y)
  if (lane == 0) {
    const float* a = p.Qv + i * (int64_t) p.ldp;
    const float* b = p.Tv + i * (int64_t) p.ldp;
    p.thr[i] = head ? lp_row_sum<KIND>(b, a, p.dout, p.p) : lp_row_sum<KIND>(a, b, p.dout, p.p);
  }
}

// matrix models (TransR, SE): proj_fwd with the relation's matrix read from global memory (L1 / L2 resident:
// every warp of the launch that works on relation r reads the same 4 d^2 bytes)
template <class M, int NV>
__device__ __forceinline__ void proj_mat(const ModelDev& md, int64_t r, int64_t e, bool want_x, float* scratch, int lane,
                                         Row<NV>& out) {
  MatCtx cx;
  mat_ctx_global<M>(md, r, cx);
  cx.scratch = scratch;
  Row<NV> E;
  sload(E, md.t[KGE_T_ENT].w + e * (int64_t) md.t[KGE_T_ENT].ld, M::d_in(md), lane);
  M::template proj_fwd<NV>(md, cx, want_x ? 0 : 1, E, out, lane);
  if (want_x && M::HAS_REL_VEC) {
    Row<NV> Rv;
    sload(Rv, md.t[KGE_T_REL_A].w + r * (int64_t) md.t[KGE_T_REL_A].ld, M::d_out(md), lane);
#pragma unroll
    for (int u = 0; u < NV; ++u) out.v[u] = out.v[u] + Rv.v[u];
  }
}
template <class M, int NV>
__global__ void __launch_bounds__(256) k_proj_table_mat(const ProjParams p, int dpad) {
  extern __shared__ float4 smem4[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* scratch = reinterpret_cast<float*>(smem4) + warp * dpad;
  const int64_t span = p.ent_hi - p.ent_lo;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  for (int64_t w = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < span * p.n_rel_chunk; w += nw) {
    const int rl = (int) (w / span);
    const int64_t el = w - (int64_t) rl * span;
    const int64_t r = p.rel0 + rl;
    if (p.goff[r + 1] == p.goff[r]) continue;
    Row<NV> row;
    proj_mat<M, NV>(p.md, r, p.ent_lo + el, p.side == KGE_SIDE_H, scratch, lane, row);
    sstore(row, p.P + ((int64_t) rl * p.span_pad + el) * p.ldp, p.dout, lane);
  }
}
// The same rows as an fp32 register-tiled GEMM per relation: C[e][n] = sum_k E[e][k] B_r[k][n], one thread per
// element, k ascending from acc = 0 with fmaf — the evaluation order of matvec_rowin / matvec_colin, so every raw
// element equals proj_fwd's bit for bit (k_proj_queries_mat keeps computing the queries' rows by proj_fwd itself).
// blockIdx = (entity tile, relation of the chunk, 128-column block).  TransR's clip and + r follow in k_proj_post.
template <class M>
__global__ void __launch_bounds__(LP_THREADS) k_proj_gemm(const ProjParams p) {
  __shared__ __align__(16) float sA[LP_DK][LP_T];
  __shared__ __align__(16) float sB[LP_DK][LP_T];
  const int rl = blockIdx.y;
  const int64_t r = p.rel0 + rl;
  if (p.goff[r + 1] == p.goff[r]) return;
  const ModelDev& md = p.md;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t span = p.ent_hi - p.ent_lo, e0 = (int64_t) blockIdx.x * LP_T;
  const int n0 = blockIdx.z * LP_T;
  const int din = M::d_in(md), dout = M::d_out(md);
  MatCtx cx;
  mat_ctx_global<M>(md, r, cx);
  const float* Mat = (M::TWO_MATS && p.side != KGE_SIDE_H) ? cx.M1 : cx.M0;
  const int ldm = cx.ld0;
  constexpr bool TR = M::KIND == KGE_SE;                 // SE: out[n] = sum_k M[n][k] e[k];  TransR: sum_k e[k] M[k][n]
  const float* Et = md.t[KGE_T_ENT].w + p.ent_lo * (int64_t) md.t[KGE_T_ENT].ld;
  const int lde = md.t[KGE_T_ENT].ld;
  float acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
  for (int k0 = 0; k0 < din; k0 += LP_DK) {
    __syncthreads();
    for (int it = tid; it < LP_T * (LP_DK / 4); it += LP_THREADS) {
      const int row = it % LP_T, c4 = it / LP_T;
      const int k = k0 + 4 * c4;
      float4 v = f4zero();
      if (e0 + row < span && k < din) v = *reinterpret_cast<const float4*>(Et + (e0 + row) * (int64_t) lde + k);   // rows are zero padded to lde
      sA[4 * c4 + 0][row] = v.x; sA[4 * c4 + 1][row] = v.y; sA[4 * c4 + 2][row] = v.z; sA[4 * c4 + 3][row] = v.w;
    }
    for (int it = tid; it < LP_T * LP_DK; it += LP_THREADS) {
      const int n = it % LP_T, kk = it / LP_T;
      const int k = k0 + kk, col = n0 + n;
      float v = 0.f;
      if (k < din && col < dout) v = TR ? Mat[(int64_t) col * ldm + k] : Mat[(int64_t) k * ldm + col];
      sB[kk][n] = v;
    }
    __syncthreads();
    const int kn = (din - k0 < LP_DK) ? din - k0 : LP_DK;
    for (int j = 0; j < kn; ++j) {
      const float4 aa = *reinterpret_cast<const float4*>(&sA[j][ty * 4]), ab = *reinterpret_cast<const float4*>(&sA[j][64 + ty * 4]);
      const float4 ba = *reinterpret_cast<const float4*>(&sB[j][tx * 4]), bb = *reinterpret_cast<const float4*>(&sB[j][64 + tx * 4]);
      const float av[8] = {aa.x, aa.y, aa.z, aa.w, ab.x, ab.y, ab.z, ab.w};
      const float bv[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
    }
  }
  float* P = p.P + ((int64_t) rl * p.span_pad) * p.ldp;
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    const int64_t el = e0 + ((a < 4) ? ty * 4 + a : 64 + ty * 4 + (a - 4));
    if (el >= span) continue;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int col = n0 + h * 64 + tx * 4;
      if (col < p.ldp)                                   // columns dout .. ldp are zeros (B is zero there)
        *reinterpret_cast<float4*>(P + el * p.ldp + col) = make_float4(acc[a][4 * h], acc[a][4 * h + 1], acc[a][4 * h + 2], acc[a][4 * h + 3]);
    }
  }
}
// TransR: c = clip1(raw), x = c + r — by proj_fwd's own clip on the raw row, warp per (relation, candidate)
template <class M, int NV>
__global__ void __launch_bounds__(256) k_proj_post(const ProjParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t span = p.ent_hi - p.ent_lo;
  const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
  const bool want_x = p.side == KGE_SIDE_H;
  for (int64_t w = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < span * p.n_rel_chunk; w += nw) {
    const int rl = (int) (w / span);
    const int64_t el = w - (int64_t) rl * span;
    const int64_t r = p.rel0 + rl;
    if (p.goff[r + 1] == p.goff[r]) continue;
    float* row = p.P + ((int64_t) rl * p.span_pad + el) * p.ldp;
    Row<NV> c;
    sload(c, row, p.dout, lane);
    clip1(c, p.md.constraint != 0);
    if (want_x) {
      Row<NV> Rv;
      sload(Rv, p.md.t[KGE_T_REL_A].w + r * (int64_t) p.md.t[KGE_T_REL_A].ld, p.dout, lane);
#pragma unroll
      for (int u = 0; u < NV; ++u) c.v[u] = c.v[u] + Rv.v[u];
    }
    sstore(c, row, p.dout, lane);
  }
}

template <class M, int NV, int KIND>
__global__ void __launch_bounds__(256) k_proj_queries_mat(const ProjParams p, int dpad) {
  extern __shared__ float4 smem4[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* scratch = reinterpret_cast<float*>(smem4) + warp * dpad;
  const int64_t lo = p.goff[p.rel0], hi = p.goff[p.rel0 + p.n_rel_chunk];
  const int64_t i = lo + (((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (i >= hi) return;
  const int64_t q = p.perm[i];
  const int64_t h = load_id(p.queries, 3 * q + 0, p.id64), r = load_id(p.queries, 3 * q + 1, p.id64),
                t = load_id(p.queries, 3 * q + 2, p.id64);
  const bool head = p.side == KGE_SIDE_H;
  Row<NV> qv, tv;
  proj_mat<M, NV>(p.md, r, head ? t : h, !head, scratch, lane, qv);
  proj_mat<M, NV>(p.md, r, head ? h : t, head, scratch, lane, tv);
  sstore(qv, p.Qv + i * (int64_t) p.ldp, p.dout, lane);
  sstore(tv, p.Tv + i * (int64_t) p.ldp, p.dout, lane);
  __syncwarp();
  if (lane == 0) {
    const float* a = p.Qv + i * (int64_t) p.ldp;
    const float* b = p.Tv + i * (int64_t) p.ldp;
    p.thr[i] = head ? lp_row_sum<KIND>(b, a, p.dout, p.p) : lp_row_sum<KIND>(a, b, p.dout, p.p);
  }
}

// ---- the tile kernel of rank_lp_tile.cuh over one relation's queries and projected table -------------------
// blockIdx = (entity tile, query tile of the chunk: tile_rel gives its relation).  HEAD: the candidate's row
// is the first operand of the difference (x_e - y_q), else the query's (x_q - y_e): the order prep used.
template <int KIND, bool HEAD>
__global__ void __launch_bounds__(LP_THREADS) k_rank_lp_tile_g(const ProjParams p) {
  __shared__ __align__(16) float sQ[LP_DK][LP_T];
  __shared__ __align__(16) float sE[LP_DK][LP_T];
  __shared__ int s_cnt[LP_T];
  const uint32_t tile = p.toff[p.rel0] + blockIdx.y;
  if (tile >= p.toff[p.rel0 + p.n_rel_chunk]) return;
  const int64_t r = p.tile_rel[tile];
  const int64_t qhi = p.goff[r + 1];
  const int64_t q0 = p.goff[r] + (int64_t) (tile - p.toff[r]) * LP_T;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t span = p.ent_hi - p.ent_lo, e0 = (int64_t) blockIdx.x * LP_T;        // shard-relative rows
  const float* E = p.P + ((r - p.rel0) * p.span_pad) * p.ldp;
  float acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
  if (tid < LP_T) s_cnt[tid] = 0;
  for (int j0 = 0; j0 < p.dout; j0 += LP_DK) {
    __syncthreads();
    for (int it = tid; it < LP_T * (LP_DK / 4); it += LP_THREADS) {
      const int row = it % LP_T, c4 = it / LP_T;
      const int j = j0 + 4 * c4;
      float4 qv = f4zero(), ev = f4zero();
      if (j < p.dout) {                                  // rows are zero padded to a multiple of 4
        if (q0 + row < qhi) qv = *reinterpret_cast<const float4*>(p.Qv + (q0 + row) * (int64_t) p.ldp + j);
        if (e0 + row < span) ev = *reinterpret_cast<const float4*>(E + (e0 + row) * (int64_t) p.ldp + j);
      }
      sQ[4 * c4 + 0][row] = qv.x; sQ[4 * c4 + 1][row] = qv.y; sQ[4 * c4 + 2][row] = qv.z; sQ[4 * c4 + 3][row] = qv.w;
      sE[4 * c4 + 0][row] = ev.x; sE[4 * c4 + 1][row] = ev.y; sE[4 * c4 + 2][row] = ev.z; sE[4 * c4 + 3][row] = ev.w;
    }
    __syncthreads();
    const int jn = (p.dout - j0 < LP_DK) ? p.dout - j0 : LP_DK;
    for (int j = 0; j < jn; ++j) {
      const float4 qa = *reinterpret_cast<const float4*>(&sQ[j][ty * 4]), qb = *reinterpret_cast<const float4*>(&sQ[j][64 + ty * 4]);
      const float4 ea = *reinterpret_cast<const float4*>(&sE[j][tx * 4]), eb = *reinterpret_cast<const float4*>(&sE[j][64 + tx * 4]);
      const float qq[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
      const float ee[8] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y, eb.z, eb.w};
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = HEAD ? lp_acc<KIND>(acc[a][b], ee[b], qq[a], p.p) : lp_acc<KIND>(acc[a][b], qq[a], ee[b], p.p);
    }
  }
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    const int ql = (a < 4) ? ty * 4 + a : 64 + ty * 4 + (a - 4);
    const int64_t q = q0 + ql;
    if (q >= qhi) continue;
    const float thr = p.thr[q];
    int c = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int el = (b < 4) ? tx * 4 + b : 64 + tx * 4 + (b - 4);
      c += (e0 + el < span && lp_clip(acc[a][b], KIND) < thr) ? 1 : 0;
    }
    if (c) atomicAdd(&s_cnt[ql], c);
  }
  __syncthreads();
  if (tid < LP_T && q0 + tid < qhi && s_cnt[tid]) atomicAdd(p.count + p.perm[q0 + tid], s_cnt[tid]);
}
// known positives of the chunk's queries: one CTA per sorted query
template <int KIND, bool HEAD>
__global__ void __launch_bounds__(256) k_rank_lp_filter_g(const ProjParams p) {
  const int64_t lo = p.goff[p.rel0], hi = p.goff[p.rel0 + p.n_rel_chunk];
  const int64_t i = lo + blockIdx.x;
  if (i >= hi) return;
  const int64_t q = p.perm[i];
  const int64_t r = load_id(p.queries, 3 * q + 1, p.id64);
  const float* E = p.P + ((r - p.rel0) * p.span_pad) * p.ldp;
  const float thr = p.thr[i];
  const float* qv = p.Qv + i * (int64_t) p.ldp;
  int c = 0;
  for (int64_t f = p.f_off[q] + threadIdx.x; f < p.f_end[q]; f += blockDim.x) {
    const int64_t e = p.f_ids[f];
    if (e < p.ent_lo || e >= p.ent_hi) continue;
    const float* er = E + (e - p.ent_lo) * (int64_t) p.ldp;
    float s = 0.f;
    for (int j = 0; j < p.dout; ++j) s = HEAD ? lp_acc<KIND>(s, er[j], qv[j], p.p) : lp_acc<KIND>(s, qv[j], er[j], p.p);
    c += (lp_clip(s, KIND) < thr) ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicSub(p.count + q, c);
}

// ---- host side ----------------------------------------------------------------------------------------------
inline bool rank_proj_eligible(const kge_cfg* c) {
  return (c->model == KGE_TRANSH || c->model == KGE_TRANSD || c->model == KGE_TRANSR || c->model == KGE_SE) &&
         (c->score == KGE_SCORE_LP || c->score == KGE_SCORE_LP_POW);
}
inline int proj_dout(const kge_cfg* c) { return (c->model == KGE_TRANSR || c->model == KGE_TRANSD) ? c->d_rel : c->d_ent; }
constexpr size_t PROJ_TABLE_BUDGET = (size_t) 256 << 20;
struct ProjWs { size_t off_cnt, off_goff, off_toff, off_tiles, off_perm, off_q, off_t, off_thr, off_p, total, total_max; int ldp, rel_chunk; int64_t span_pad; };
inline void layout_proj_ws(const kge_cfg* c, int64_t Q, int64_t span, ProjWs& L) {
  L.ldp = (proj_dout(c) + 3) / 4 * 4;
  L.span_pad = (span + LP_T - 1) / LP_T * LP_T;
  const size_t per_rel = sizeof(float) * (size_t) L.span_pad * L.ldp;
  size_t budget = PROJ_TABLE_BUDGET;
  if (const char* e = getenv("KGE_PROJ_BUDGET_KB")) budget = (size_t) atoll(e) << 10;      // tests: force several relation chunks
  int64_t rc = (int64_t) (budget / per_rel);
  rc = rc < 1 ? 1 : (rc > c->n_rel ? c->n_rel : rc);
  const int64_t qt = Q / LP_T + 1;                       // tiles of one chunk <= Q / 128 + rel_chunk <= 65535 (gridDim.y)
  if (rc > 65535 - qt) rc = 65535 - qt > 1 ? 65535 - qt : 1;
  L.rel_chunk = (int) rc;
  size_t o = 0;
  auto take = [&](size_t b) { size_t at = o; o += (b + 255) / 256 * 256; return at; };
  L.off_cnt = take(sizeof(uint32_t) * (size_t) (c->n_rel + 1));
  L.off_goff = take(sizeof(uint32_t) * (size_t) (c->n_rel + 1));
  L.off_toff = take(sizeof(uint32_t) * (size_t) (c->n_rel + 1));
  L.off_tiles = take(sizeof(uint32_t) * (size_t) (Q / LP_T + c->n_rel + 1));
  L.off_perm = take(sizeof(uint32_t) * (size_t) Q);
  L.off_q = take(sizeof(float) * (size_t) Q * L.ldp);
  L.off_t = take(sizeof(float) * (size_t) Q * L.ldp);
  L.off_thr = take(sizeof(float) * (size_t) Q);
  L.off_p = take(per_rel * (size_t) rc);
  L.total = o;
  // a narrower shard fits more relations per chunk: what any shard of this span or less can ask for
  const size_t all = per_rel * (size_t) c->n_rel;
  const size_t cap = all < budget ? all : budget;
  L.total_max = L.off_p + ((per_rel > cap ? per_rel : cap) + 255) / 256 * 256;
}

template <class M, int NV, int KIND, bool MAT>
inline int rank_proj_chunk(const ProjParams& p, int64_t span, int sm_count, size_t msmem, int dpad, cudaStream_t st) {
  const int64_t warps = span * p.n_rel_chunk;
  int64_t blocks = (warps + 7) / 8;
  if (blocks > (int64_t) sm_count * 32) blocks = (int64_t) sm_count * 32;
  const unsigned qblocks = (unsigned) ((p.Q + 7) / 8);
  if constexpr (MAT) {
    if (getenv("KGE_PROJ_MATVEC")) {                     // A/B switch: the warp-per-row table of the first version
      k_proj_table_mat<M, NV><<<(unsigned) blocks, 256, msmem, st>>>(p, dpad);
    } else {
      const dim3 gg((unsigned) (p.span_pad / LP_T), (unsigned) p.n_rel_chunk, (unsigned) ((p.dout + LP_T - 1) / LP_T));
      k_proj_gemm<M><<<gg, LP_THREADS, 0, st>>>(p);
      if (M::KIND == KGE_TRANSR) k_proj_post<M, NV><<<(unsigned) blocks, 256, 0, st>>>(p);
    }
    k_proj_queries_mat<M, NV, KIND><<<qblocks, 256, msmem, st>>>(p, dpad);
  } else {
    k_proj_table<M, NV><<<(unsigned) blocks, 256, 0, st>>>(p);
    k_proj_queries<M, NV, KIND><<<qblocks, 256, 0, st>>>(p);
  }
  const dim3 grid((unsigned) (p.span_pad / LP_T), (unsigned) (p.Q / LP_T + p.n_rel_chunk));   // upper bound on the chunk's tiles
  if (p.side == KGE_SIDE_H) {
    k_rank_lp_tile_g<KIND, true><<<grid, LP_THREADS, 0, st>>>(p);
    if (p.f_off) k_rank_lp_filter_g<KIND, true><<<(unsigned) p.Q, 256, 0, st>>>(p);
  } else {
    k_rank_lp_tile_g<KIND, false><<<grid, LP_THREADS, 0, st>>>(p);
    if (p.f_off) k_rank_lp_filter_g<KIND, false><<<(unsigned) p.Q, 256, 0, st>>>(p);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("projected-table ranking launch failed: %s", cudaGetErrorString(e)); return KGE_ECUDA; }
  return KGE_OK;
}
template <class M, int NV, bool MAT>
inline int rank_proj_kind(const ProjParams& p, int kind, int64_t span, int sm_count, size_t msmem, int dpad, cudaStream_t st) {
  switch (kind) {
    case LPK_L1: return rank_proj_chunk<M, NV, LPK_L1, MAT>(p, span, sm_count, msmem, dpad, st);
    case LPK_L2: return rank_proj_chunk<M, NV, LPK_L2, MAT>(p, span, sm_count, msmem, dpad, st);
    case LPK_LINF: return rank_proj_chunk<M, NV, LPK_LINF, MAT>(p, span, sm_count, msmem, dpad, st);
    default: return rank_proj_chunk<M, NV, LPK_GEN, MAT>(p, span, sm_count, msmem, dpad, st);
  }
}
// one per model translation unit (model_transh.cu, ...): the chunk launcher for that model
#define KGE_DECL_PROJ(name) int launch_rank_proj_##name(const ProjParams&, int nv, int kind, int64_t span, int sm_count, cudaStream_t);
KGE_DECL_PROJ(transh)
KGE_DECL_PROJ(transd)
KGE_DECL_PROJ(transr)
KGE_DECL_PROJ(se)
#undef KGE_DECL_PROJ
template <class M>
inline int launch_rank_proj_vec(const ProjParams& p, int nv, int kind, int64_t span, int sm_count, cudaStream_t st) {
  switch (nv) {
    case 1: return rank_proj_kind<M, 1, false>(p, kind, span, sm_count, 0, 0, st);
    case 2: return rank_proj_kind<M, 2, false>(p, kind, span, sm_count, 0, 0, st);
    case 4: return rank_proj_kind<M, 4, false>(p, kind, span, sm_count, 0, 0, st);
    default: set_error("projected-table ranking: rows up to 512 floats (nv=%d)", nv); return KGE_EUNSUPPORTED;
  }
}
template <class M>
inline int launch_rank_proj_mat(const ProjParams& p, int nv, int kind, int64_t span, int sm_count, cudaStream_t st) {
  const int dpad = 128 * nv;
  const size_t msmem = sizeof(float) * 8 * (size_t) dpad;
  switch (nv) {
    case 1: return rank_proj_kind<M, 1, true>(p, kind, span, sm_count, msmem, dpad, st);
    case 2: return rank_proj_kind<M, 2, true>(p, kind, span, sm_count, msmem, dpad, st);
    default: set_error("matrix models support embedding sizes up to 256 (nv=%d)", nv); return KGE_EUNSUPPORTED;
  }
}

inline int rank_proj_tiled(kge_ctx_impl* ctx, const kge_cfg* cfg, const kge_tables* tables, const void* queries, int64_t Q,
                           int side, const kge_filter* filter, int64_t ent_lo, int64_t ent_hi, int32_t* count_out,
                           void* ws, size_t ws_bytes, cudaStream_t st) {
  ProjWs L;
  const int64_t span = ent_hi - ent_lo;
  layout_proj_ws(cfg, Q, span, L);
  KGE_REQUIRE(ws && ws_bytes >= L.total, KGE_EWORKSPACE, "rank workspace %zu < required %zu bytes", ws_bytes, L.total);
  KGE_REQUIRE(((uintptr_t) ws & 255) == 0, KGE_EINVAL, "workspace must be 256-byte aligned");
  KGE_REQUIRE(Q / LP_T + 2 < 65535, KGE_EUNSUPPORTED, "projected-table ranking takes up to 8 M queries per call");
  char* wsb = reinterpret_cast<char*>(ws);
  ProjParams p;
  fill_model_dev(cfg, tables, p.md);
  p.side = side; p.queries = queries; p.id64 = cfg->id_dtype; p.Q = Q;
  uint32_t* cnt = reinterpret_cast<uint32_t*>(wsb + L.off_cnt);
  uint32_t* goff = reinterpret_cast<uint32_t*>(wsb + L.off_goff);
  uint32_t* perm = reinterpret_cast<uint32_t*>(wsb + L.off_perm);
  uint32_t* toff = reinterpret_cast<uint32_t*>(wsb + L.off_toff);
  uint32_t* tile_rel = reinterpret_cast<uint32_t*>(wsb + L.off_tiles);
  p.goff = goff; p.perm = perm; p.toff = toff; p.tile_rel = tile_rel;
  p.ent_lo = ent_lo; p.ent_hi = ent_hi; p.span_pad = L.span_pad;
  p.P = reinterpret_cast<float*>(wsb + L.off_p); p.ldp = L.ldp; p.dout = proj_dout(cfg);
  p.Qv = reinterpret_cast<float*>(wsb + L.off_q); p.Tv = reinterpret_cast<float*>(wsb + L.off_t);
  p.thr = reinterpret_cast<float*>(wsb + L.off_thr);
  p.f_off = (filter && filter->off) ? filter->off : nullptr;
  p.f_end = (filter && filter->off) ? (filter->end ? filter->end : filter->off + 1) : nullptr;
  p.f_ids = filter ? filter->ids : nullptr;
  p.count = count_out;
  p.p = cfg->score_p;
  // 1. queries grouped by relation
  KGE_CUDA_OK(cudaMemsetAsync(cnt, 0, sizeof(uint32_t) * (size_t) (cfg->n_rel + 1), st));
  k_qrel_hist<<<(unsigned) ((Q + 255) / 256), 256, 0, st>>>(queries, cfg->id_dtype, Q, cnt);
  k_qrel_scan<true><<<1, 1024, 0, st>>>(cnt, toff, cfg->n_rel);
  k_qrel_scan<false><<<1, 1024, 0, st>>>(cnt, goff, cfg->n_rel);
  k_qrel_tiles<<<(unsigned) ((cfg->n_rel + 255) / 256), 256, 0, st>>>(toff, cfg->n_rel, tile_rel);
  k_qrel_scatter<<<(unsigned) ((Q + 255) / 256), 256, 0, st>>>(queries, cfg->id_dtype, Q, goff, cnt, perm);
  KGE_CUDA_OK(cudaGetLastError());
  // zero padding of the staged rows (columns dout .. ldp, rows beyond the span) is read by the tile kernel's float4 loads
  KGE_CUDA_OK(cudaMemsetAsync(p.Qv, 0, sizeof(float) * (size_t) Q * L.ldp, st));
  const int kind = isinf(cfg->score_p) ? LPK_LINF : (cfg->score_p == 1.f ? LPK_L1 : (cfg->score_p == 2.f ? LPK_L2 : LPK_GEN));
  const bool mat = is_matrix_model(cfg->model);
  const int nv = mat ? pick_nv_strided(cfg) : pick_nv(cfg);
  KGE_REQUIRE(nv > 0, KGE_EUNSUPPORTED, "embedding size too large for the rank kernel");
  for (int64_t r0 = 0; r0 < cfg->n_rel; r0 += L.rel_chunk) {
    p.rel0 = (int) r0;
    p.n_rel_chunk = (int) ((cfg->n_rel - r0 < L.rel_chunk) ? cfg->n_rel - r0 : L.rel_chunk);
    if (L.ldp != p.dout || L.span_pad != span)
      KGE_CUDA_OK(cudaMemsetAsync(p.P, 0, sizeof(float) * (size_t) p.n_rel_chunk * L.span_pad * L.ldp, st));
    int rc;
    switch (cfg->model) {
      case KGE_TRANSH: rc = launch_rank_proj_transh(p, nv, kind, span, ctx->sm_count, st); break;
      case KGE_TRANSD: rc = launch_rank_proj_transd(p, nv, kind, span, ctx->sm_count, st); break;
      case KGE_TRANSR: rc = launch_rank_proj_transr(p, nv, kind, span, ctx->sm_count, st); break;
      default: rc = launch_rank_proj_se(p, nv, kind, span, ctx->sm_count, st);
    }
    if (rc) return rc;
  }
  return KGE_OK;
}

}  // namespace kge
