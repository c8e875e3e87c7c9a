// rank.cu — host side of kge_score and kge_rank.
#include "host_common.cuh"
#include "rank_kernels.cuh"
#include "rank_mat_kernels.cuh"
#include "rank_dot_host.cuh"
#include "rank_lp_tile.cuh"
#include "rank_proj_tile.cuh"
#include <stdlib.h>

namespace kge {
static bool force_generic_rank() {
  const char* e = getenv("KGE_RANK_GENERIC");     // A/B switch for tests and profiling
  return e && e[0] == '1';
}
size_t rank_workspace_bytes(const kge_cfg* cfg, int64_t Q) {
  if (rank_dot_eligible(cfg)) {
    DotWs L;
    layout_dot_ws(cfg, Q, cfg->n_ent, L);         // sized for the full entity range (largest shard)
    return L.total;
  }
  if (rank_lp_eligible(cfg)) {
    LpWs L;
    layout_lp_ws(cfg, Q, L);
    return L.total;
  }
  if (rank_proj_eligible(cfg)) {
    ProjWs L;
    layout_proj_ws(cfg, Q, cfg->n_ent, L);
    return L.total_max;
  }
  return 256;
}
}  // namespace kge
using namespace kge;

extern "C" int kge_score(kge_ctx* ctx_, const kge_cfg* cfg, const kge_tables* tables, const void* h, const void* r,
                         const void* t, int64_t n, float* out, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  if (int rc = check_tables(cfg, tables, false)) return rc;
  KGE_REQUIRE(h && r && t && out && n > 0, KGE_EINVAL, "h / r / t / out / n invalid");
  ScoreParams p;
  fill_model_dev(cfg, tables, p.md);
  p.h = h; p.r = r; p.t = t; p.id64 = cfg->id_dtype; p.n = n; p.out = out;
  AuxLaunch L;
  const bool mat = is_matrix_model(cfg->model);
  L.nv = mat ? pick_nv_strided(cfg) : pick_nv(cfg);
  int64_t blocks = (n + 7) / 8;
  if (blocks > (int64_t) ctx->sm_count * 16) blocks = (int64_t) ctx->sm_count * 16;
  L.grid = dim3((unsigned) blocks);
  L.stream = (cudaStream_t) stream;
  if (mat) {
    KGE_REQUIRE(L.nv > 0, KGE_EUNSUPPORTED, "matrix models support embedding sizes up to 256");
    MatAux a;
    a.dpad = 128 * L.nv; a.stage = 0;
    const size_t smem = sizeof(float) * 8 * (size_t) a.dpad;
    switch (cfg->model) {
      case KGE_SE: return launch_score_se(p, L, a, smem);
      case KGE_TRANSR: return launch_score_transr(p, L, a, smem);
      default: return launch_score_rescal(p, L, a, smem);
    }
  }
  switch (cfg->model) {
    case KGE_UM: return launch_score_um(p, L);
    case KGE_TRANSE: return launch_score_transe(p, L);
    case KGE_TRANSH: return launch_score_transh(p, L);
    case KGE_TRANSD: return launch_score_transd(p, L);
    case KGE_ROTATE: return launch_score_rotate(p, L);
    case KGE_DISTMULT: return launch_score_distmult(p, L);
  }
  set_error("model %d has no score kernel", cfg->model);
  return KGE_EUNSUPPORTED;
}

extern "C" int kge_rank(kge_ctx* ctx_, const kge_cfg* cfg, const kge_tables* tables, const void* queries, int64_t Q,
                        int side, const kge_filter* filter, int64_t ent_lo, int64_t ent_hi, int32_t* count_out,
                        void* ws, size_t ws_bytes, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  if (int rc = check_tables(cfg, tables, false)) return rc;
  KGE_REQUIRE(queries && count_out && Q > 0, KGE_EINVAL, "queries / count_out / Q invalid");
  KGE_REQUIRE(side == KGE_SIDE_H || side == KGE_SIDE_T, KGE_EINVAL, "side must be 'h' or 't' (BaseModel.py:639-644)");
  KGE_REQUIRE(0 <= ent_lo && ent_lo < ent_hi && ent_hi <= cfg->n_ent, KGE_EINVAL, "bad entity shard [%lld, %lld)",
              (long long) ent_lo, (long long) ent_hi);
  KGE_REQUIRE(!filter || !filter->off || filter->ids, KGE_EINVAL, "filter.off given without filter.ids");
  if (rank_dot_eligible(cfg) && !force_generic_rank()) {
    PhaseTimer ptd(ctx, (cudaStream_t) stream);
    ptd.begin(5);
    if (int rc = rank_dot_tensor(ctx, cfg, tables, queries, Q, side, filter, ent_lo, ent_hi, count_out, ws, ws_bytes,
                                 (cudaStream_t) stream, &ptd)) return rc;
    ptd.end(5);
    return KGE_OK;
  }
  if (rank_lp_eligible(cfg) && !force_generic_rank()) {
    PhaseTimer ptl(ctx, (cudaStream_t) stream);
    ptl.begin(5);
    if (int rc = rank_lp_tiled(cfg, tables, queries, Q, side, filter, ent_lo, ent_hi, count_out, ws, ws_bytes,
                               (cudaStream_t) stream)) return rc;
    ptl.end(5);
    return KGE_OK;
  }
  if (rank_proj_eligible(cfg) && !force_generic_rank()) {
    PhaseTimer ptp(ctx, (cudaStream_t) stream);
    ptp.begin(5);
    if (int rc = rank_proj_tiled(ctx, cfg, tables, queries, Q, side, filter, ent_lo, ent_hi, count_out, ws, ws_bytes,
                                 (cudaStream_t) stream)) return rc;
    ptp.end(5);
    return KGE_OK;
  }
  const bool mat = is_matrix_model(cfg->model);
  RankParams p;
  fill_model_dev(cfg, tables, p.md);
  p.queries = queries; p.id64 = cfg->id_dtype; p.Q = Q; p.side = side;
  p.f_off = filter ? filter->off : nullptr;
  p.f_end = (filter && filter->off) ? (filter->end ? filter->end : filter->off + 1) : nullptr;
  p.f_ids = filter ? filter->ids : nullptr;
  p.ent_lo = ent_lo; p.ent_hi = ent_hi;
  p.count = count_out;
  AuxLaunch L;
  L.nv = mat ? pick_nv_strided(cfg) : pick_nv(cfg);
  KGE_REQUIRE(L.nv > 0, KGE_EUNSUPPORTED, "embedding size too large for the rank kernel");
  L.stream = (cudaStream_t) stream;
  MatAux ma;
  size_t msmem = 0;
  if (mat) {
    ma.dpad = 128 * L.nv;
    const int mr = cfg->d_ent, mc = cfg->model == KGE_TRANSR ? cfg->d_rel : cfg->d_ent;
    const size_t base = sizeof(float) * RANK_WARPS * (size_t) ma.dpad;
    const size_t staged = base + sizeof(float) * (size_t) (cfg->model == KGE_SE ? 2 : 1) * mr * (mc | 1);
    ma.stage = staged <= (size_t) ctx->smem_optin;
    msmem = ma.stage ? staged : base;
  }
  const int64_t span = ent_hi - ent_lo;
  // enough CTAs per query to fill the machine, at least 64 candidates per warp pass
  int chunk = 2048;
  while (chunk > 256 && Q * ((span + chunk - 1) / chunk) < (int64_t) ctx->sm_count * 8) chunk >>= 1;
  p.chunk = chunk;
  PhaseTimer pt(ctx, (cudaStream_t) stream);
  pt.begin(5);
  const int64_t QMAX = 65535;    // gridDim.y limit
  for (int64_t q0 = 0; q0 < Q; q0 += QMAX) {
    const int64_t qn = (Q - q0 < QMAX) ? Q - q0 : QMAX;
    RankParams pp = p;
    pp.queries = (const char*) queries + (size_t) q0 * 3 * (cfg->id_dtype ? 8 : 4);
    pp.count = count_out + q0;
    if (pp.f_off) { pp.f_off = p.f_off + q0; pp.f_end = p.f_end + q0; }
    pp.Q = qn;
    L.grid = dim3((unsigned) ((span + chunk - 1) / chunk), (unsigned) qn);
    int rc;
    switch (cfg->model) {
      case KGE_SE: rc = launch_rank_se(pp, L, ma, msmem); break;
      case KGE_TRANSR: rc = launch_rank_transr(pp, L, ma, msmem); break;
      case KGE_RESCAL: rc = launch_rank_rescal(pp, L, ma, msmem); break;
      case KGE_UM: rc = launch_rank_um(pp, L); break;
      case KGE_TRANSE: rc = launch_rank_transe(pp, L); break;
      case KGE_TRANSH: rc = launch_rank_transh(pp, L); break;
      case KGE_TRANSD: rc = launch_rank_transd(pp, L); break;
      case KGE_ROTATE: rc = launch_rank_rotate(pp, L); break;
      case KGE_DISTMULT: rc = launch_rank_distmult(pp, L); break;
      default: set_error("model %d has no rank kernel", cfg->model); rc = KGE_EUNSUPPORTED;
    }
    if (rc) return rc;
  }
  pt.end(5);
  return KGE_OK;
}
