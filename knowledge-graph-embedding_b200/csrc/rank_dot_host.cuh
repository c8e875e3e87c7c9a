// rank_dot_host.cuh — host side of the tensor-core ranking path (included by rank.cu).
#pragma once
#include <stdlib.h>
#include <cudaTypedefs.h>
#include "host_common.cuh"
#include "rank_dot_tc.cuh"

namespace kge {

// Dot-family with a plain entity row on the candidate side
inline bool rank_dot_eligible(const kge_cfg* c) {
  if (c->model == KGE_DISTMULT || c->model == KGE_RESCAL) return true;
  return c->score == KGE_SCORE_DOT && (c->model == KGE_UM || c->model == KGE_TRANSE);
}

struct DotWs {
  int kpad, words_per_q, n_etiles;
  int64_t qblk;                    // queries per launch (multiple of 128)
  size_t off_qm, off_qhi, off_qlo, off_ehi, off_elo, off_nlo, off_nq, off_pos, off_bias, off_hi, off_lo, off_maxn, off_wins, off_ucnt, off_ulist, off_filt, total;
  int ldE;
};
inline void layout_dot_ws(const kge_cfg* c, int64_t Q, int64_t span, DotWs& L) {
  L.kpad = (c->d_ent + TC_BK - 1) / TC_BK * TC_BK;
  L.n_etiles = (int) ((span + TC_BN - 1) / TC_BN);
  L.words_per_q = L.n_etiles * (TC_BN / 32);
  int64_t cap = ((int64_t) 256 << 20) / ((int64_t) L.words_per_q * 4);
  cap = cap / TC_BM * TC_BM;
  if (cap < TC_BM) cap = TC_BM;
  if (cap > 16384) cap = 16384;
  const int64_t qpad = (Q + TC_BM - 1) / TC_BM * TC_BM;
  L.qblk = qpad < cap ? qpad : cap;
  size_t o = 0;
  auto take = [&](size_t b) { size_t at = o; o += align_up(b); return at; };
  L.off_maxn = take(256);
  L.off_qm = take(sizeof(float) * (size_t) L.qblk * L.kpad);
  L.off_qhi = take(sizeof(float) * (size_t) L.qblk * L.kpad);
  L.off_qlo = take(sizeof(float) * (size_t) L.qblk * L.kpad);
  L.ldE = (c->d_ent + 3) / 4 * 4;
  L.off_ehi = take(sizeof(float) * (size_t) span * L.ldE);
  L.off_elo = take(sizeof(float) * (size_t) span * L.ldE);
  L.off_nlo = take(sizeof(float) * (size_t) L.n_etiles * TC_BN);       // ||e_lo|| per candidate, whole tiles
  L.off_nq = take(sizeof(float) * (size_t) L.qblk);
  L.off_pos = take(sizeof(float) * (size_t) L.qblk);
  L.off_bias = take(sizeof(float) * (size_t) L.qblk);
  L.off_hi = take(sizeof(float) * (size_t) L.qblk);
  L.off_lo = take(sizeof(float) * (size_t) L.qblk);
  L.off_wins = take(sizeof(int32_t) * (size_t) L.qblk);
  L.off_ucnt = take(256);                                         // directly behind wins: one memset clears both
  L.off_ulist = take(sizeof(uint2) * (size_t) L.qblk * TC_UCAP);
  L.off_filt = take(sizeof(uint32_t) * (size_t) L.qblk * L.words_per_q);
  L.total = o;
}

inline PFN_cuTensorMapEncodeTiled_v12000 tensor_map_encoder() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}
// fp32 row-major [rows x cols] at a stride of ld floats; box = 32 floats x box_rows, 128B swizzle, OOB -> 0
inline int make_map_2d(CUtensorMap* m, const float* base, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows) {
  auto enc = tensor_map_encoder();
  KGE_REQUIRE(enc, KGE_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t) TC_BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  KGE_REQUIRE(r == CUDA_SUCCESS, KGE_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int) r);
  return KGE_OK;
}

// count_out[q] += #{e in [ent_lo, ent_hi): score(e) > pos(q), e not filtered}
inline int rank_dot_tensor(kge_ctx_impl* ctx, const kge_cfg* cfg, const kge_tables* tables, const void* queries,
                           int64_t Q, int side, const kge_filter* filter, int64_t ent_lo, int64_t ent_hi,
                           int32_t* count_out, void* ws, size_t ws_bytes, cudaStream_t st, PhaseTimer* pt) {
  DotWs L;
  layout_dot_ws(cfg, Q, ent_hi - ent_lo, L);
  KGE_REQUIRE(ws && ws_bytes >= L.total, KGE_EWORKSPACE, "rank workspace %zu < required %zu bytes", ws_bytes, L.total);
  KGE_REQUIRE(((uintptr_t) ws & 255) == 0, KGE_EINVAL, "workspace must be 256-byte aligned");
  char* wsb = reinterpret_cast<char*>(ws);
  const kge_table& E = tables->t[KGE_T_ENT];
  static bool attr_set = false;
  if (!attr_set) {
    KGE_CUDA_OK(cudaFuncSetAttribute(k_rank_dot_tc<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<3>::SMEM_BYTES));
    KGE_CUDA_OK(cudaFuncSetAttribute(k_rank_dot_tc<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<2>::SMEM_BYTES));
    attr_set = true;
  }
  const int passes = ctx->opt_tc_passes == 2 ? 2 : 3;
  // Two chains in front of the GEMM: the caller's stream takes the query side (max norm, query vectors and their
  // hi/lo split, thresholds), the context's side stream the candidate side (hi/lo split of the table, once per call,
  // and each block's known-positive bitmask); they meet in front of the GEMM.  KGE_RANK_SERIAL=1: one stream.
  static const bool serial = [] { const char* e = getenv("KGE_RANK_SERIAL"); return e && atoi(e) != 0; }();
  const bool fork = ctx->aux != nullptr && !serial;
  cudaStream_t ts = fork ? ctx->aux : st;
  if (fork) {                                  // the side stream starts behind everything already enqueued on st
    KGE_CUDA_OK(cudaEventRecord(ctx->ev_fork, st));
    KGE_CUDA_OK(cudaStreamWaitEvent(ts, ctx->ev_fork, 0));
  }
  unsigned* maxn = reinterpret_cast<unsigned*>(wsb + L.off_maxn);
  KGE_CUDA_OK(cudaMemsetAsync(maxn, 0, 256, st));
  {
    int64_t blocks = (ent_hi - ent_lo + 7) / 8;
    if (blocks > (int64_t) ctx->sm_count * 16) blocks = (int64_t) ctx->sm_count * 16;
    k_rank_maxnorm<<<(unsigned) blocks, 256, 0, st>>>(E.w, E.ld, E.cols, ent_lo, ent_hi, maxn);
    KGE_CUDA_OK(cudaGetLastError());
  }
  // exact TF32 split of the candidate shard (E = E_hi + E_lo)
  const int64_t span = ent_hi - ent_lo;
  float* Ehi = reinterpret_cast<float*>(wsb + L.off_ehi);
  float* Elo = passes == 3 ? reinterpret_cast<float*>(wsb + L.off_elo) : nullptr;
  float* nlo = reinterpret_cast<float*>(wsb + L.off_nlo);
  KGE_CUDA_OK(cudaMemsetAsync(nlo, 0, sizeof(float) * (size_t) L.n_etiles * TC_BN, ts));
  KGE_REQUIRE(E.ld == L.ldE, KGE_EUNSUPPORTED, "tensor-core ranking needs ld == round4(cols) (got %d)", E.ld);
  {
    int64_t blocks = (span + 7) / 8;
    if (blocks > (int64_t) ctx->sm_count * 32) blocks = (int64_t) ctx->sm_count * 32;
    k_rank_split_table<<<(unsigned) blocks, 256, 0, ts>>>(E.w, E.ld, E.cols, ent_lo, ent_hi, Ehi, Elo, nlo);
    KGE_CUDA_OK(cudaGetLastError());
  }
  CUtensorMap map_ah, map_al, map_bh, map_bl;
  float* Qm = reinterpret_cast<float*>(wsb + L.off_qm);
  float* Qhi = reinterpret_cast<float*>(wsb + L.off_qhi);
  float* Qlo = reinterpret_cast<float*>(wsb + L.off_qlo);
  if (int rc = make_map_2d(&map_ah, Qhi, (uint64_t) L.kpad, (uint64_t) L.qblk, (uint64_t) L.kpad, TC_BM)) return rc;
  if (int rc = make_map_2d(&map_al, Qlo, (uint64_t) L.kpad, (uint64_t) L.qblk, (uint64_t) L.kpad, TC_BM)) return rc;
  if (int rc = make_map_2d(&map_bh, Ehi, (uint64_t) E.cols, (uint64_t) span, (uint64_t) E.ld, TC_BN)) return rc;
  if (int rc = make_map_2d(&map_bl, passes == 3 ? Elo : Ehi, (uint64_t) E.cols, (uint64_t) span, (uint64_t) E.ld, TC_BN)) return rc;
  const size_t idb = cfg->id_dtype ? 8 : 4;
  const bool filtered = filter && filter->off;
  for (int64_t q0 = 0; q0 < Q; q0 += L.qblk) {
    const int64_t qn = (Q - q0 < L.qblk) ? Q - q0 : L.qblk;
    KGE_CUDA_OK(cudaMemsetAsync(wsb + L.off_qm, 0, 3 * align_up(sizeof(float) * (size_t) L.qblk * L.kpad), st));
    DotPrepParams pp;
    fill_model_dev(cfg, tables, pp.md);
    pp.queries = (const char*) queries + (size_t) q0 * 3 * idb; pp.id64 = cfg->id_dtype; pp.Q = qn; pp.side = side;
    pp.Qm = Qm; pp.Qhi = Qhi; pp.Qlo = Qlo; pp.kpad = L.kpad;
    pp.pos = reinterpret_cast<float*>(wsb + L.off_pos); pp.bias = reinterpret_cast<float*>(wsb + L.off_bias);
    pp.thr_hi = reinterpret_cast<float*>(wsb + L.off_hi); pp.thr_lo = reinterpret_cast<float*>(wsb + L.off_lo);
    pp.nq = reinterpret_cast<float*>(wsb + L.off_nq);
    pp.maxnorm_bits = maxn;
    // apart from q.e_lo (bounded per candidate by ||q|| ||e_lo||): truncation of q_lo to TF32 inside the MMA
    // (2^-21) + fp32 accumulation of 2*kpad products, with a wide safety factor
    pp.eps = 4.0e-6f + 3.6e-7f * (float) L.kpad;
    k_rank_dot_prep<<<(unsigned) ((qn + 7) / 8), 256, 0, st>>>(pp);
    KGE_CUDA_OK(cudaGetLastError());
    uint32_t* filt = nullptr;
    if (filtered) {
      filt = reinterpret_cast<uint32_t*>(wsb + L.off_filt);
      KGE_CUDA_OK(cudaMemsetAsync(filt, 0, sizeof(uint32_t) * (size_t) qn * L.words_per_q, ts));
      k_rank_filter_bits<<<(unsigned) ((qn + 7) / 8), 256, 0, ts>>>(filter->off + q0, (filter->end ? filter->end : filter->off + 1) + q0,
                                                                    filter->ids, qn, ent_lo, ent_hi, filt, L.words_per_q);
      KGE_CUDA_OK(cudaGetLastError());
    }

    DotTcParams tp;
    tp.n_kc = L.kpad / TC_BK; tp.n_k8 = (cfg->d_ent + 7) / 8; tp.Q = qn; tp.n_qblocks = (int) ((qn + TC_BM - 1) / TC_BM);
    tp.ent_lo = ent_lo; tp.ent_hi = ent_hi; tp.n_etiles = L.n_etiles;
    int split = 1;
    while (tp.n_qblocks * split < 2 * ctx->sm_count && split * 2 <= L.n_etiles) split *= 2;
    tp.n_split = split;
    tp.thr_hi = pp.thr_hi; tp.thr_lo = pp.thr_lo; tp.nq = pp.nq; tp.nlo = nlo;
    tp.wins = reinterpret_cast<int32_t*>(wsb + L.off_wins); tp.ucnt = reinterpret_cast<uint32_t*>(wsb + L.off_ucnt);
    tp.ulist = reinterpret_cast<uint2*>(wsb + L.off_ulist); tp.ucap = (uint32_t) ((size_t) L.qblk * TC_UCAP);
    tp.words_per_q = L.words_per_q;
    KGE_CUDA_OK(cudaMemsetAsync(wsb + L.off_wins, 0, (L.off_ulist - L.off_wins), st));
    tp.filt = filt;
    if (fork) {                                // candidate side ready
      KGE_CUDA_OK(cudaEventRecord(ctx->ev_join, ts));
      KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join, 0));
    }
    int grid = tp.n_qblocks * split;
    if (grid > ctx->sm_count) grid = ctx->sm_count;
    const bool timed = (q0 == 0);           // phase slots hold one interval per call: time the first query block
    if (timed) pt->begin(6);
    if (passes == 3) k_rank_dot_tc<3><<<grid, TC_THREADS, TcCfg<3>::SMEM_BYTES, st>>>(map_ah, map_al, map_bh, map_bl, tp);
    else k_rank_dot_tc<2><<<grid, TC_THREADS, TcCfg<2>::SMEM_BYTES, st>>>(map_ah, map_al, map_bh, map_bl, tp);
    KGE_CUDA_OK(cudaGetLastError());
    if (timed) { pt->end(6); pt->begin(7); }

    DotFixParams fp;
    fp.Qm = Qm; fp.kpad = L.kpad; fp.E = E.w; fp.ldE = E.ld; fp.cols = E.cols;
    fp.pos = pp.pos; fp.bias = pp.bias; fp.wins = tp.wins; fp.ucnt = tp.ucnt; fp.ulist = tp.ulist; fp.ucap = tp.ucap;
    fp.filt = filt; fp.words_per_q = L.words_per_q;
    fp.Q = qn; fp.ent_lo = ent_lo; fp.ent_hi = ent_hi; fp.count = count_out + q0;
    k_rank_dot_fix<<<(unsigned) ctx->sm_count * 8, 256, 0, st>>>(fp);
    KGE_CUDA_OK(cudaGetLastError());
    k_rank_dot_final<<<(unsigned) ((qn + FIX_WARPS * 32 - 1) / (FIX_WARPS * 32)), FIX_WARPS * 32, 0, st>>>(fp);
    KGE_CUDA_OK(cudaGetLastError());
    if (timed) pt->end(7);
    if (fork && q0 + L.qblk < Q) {             // the next block's bitmask may overwrite this one's only now
      KGE_CUDA_OK(cudaEventRecord(ctx->ev_fork, st));
      KGE_CUDA_OK(cudaStreamWaitEvent(ts, ctx->ev_fork, 0));
    }
  }
  return KGE_OK;
}

}  // namespace kge
