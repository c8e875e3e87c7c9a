// host_common.cuh — host-side validation and translation of the C-ABI structs into
// the device parameter structs.
#pragma once
#include <math.h>
#include "models.cuh"
#include "loss.cuh"

namespace kge {

inline bool is_matrix_model(int m) { return m == KGE_SE || m == KGE_TRANSR || m == KGE_RESCAL; }
inline bool uses_score_fn(int m) { return m != KGE_DISTMULT && m != KGE_RESCAL; }

// expected logical cols of each table slot, 0 = slot unused
inline void expected_cols(const kge_cfg* c, int cols[KGE_MAX_TABLES]) {
  const int de = c->d_ent, dr = c->d_rel;
  cols[0] = cols[1] = cols[2] = cols[3] = 0;
  switch (c->model) {
    case KGE_UM: cols[0] = de; break;
    case KGE_SE: cols[0] = de; cols[2] = de * de; cols[3] = de * de; break;
    case KGE_TRANSE: cols[0] = de; cols[2] = de; break;
    case KGE_TRANSH: cols[0] = de; cols[2] = de; cols[3] = de; break;
    case KGE_TRANSR: cols[0] = de; cols[2] = dr; cols[3] = de * dr; break;
    case KGE_TRANSD: cols[0] = de; cols[1] = de; cols[2] = dr; cols[3] = dr; break;
    case KGE_ROTATE: cols[0] = 2 * de; cols[2] = de; break;
    case KGE_DISTMULT: cols[0] = de; cols[2] = de; break;
    case KGE_RESCAL: cols[0] = de; cols[2] = de * de; break;
  }
}

inline int check_cfg(const kge_cfg* c) {
  KGE_REQUIRE(c != nullptr, KGE_EINVAL, "cfg is NULL");
  KGE_REQUIRE(c->abi_version == KGE_ABI_VERSION, KGE_EINVAL, "cfg.abi_version %d != library %d",
              c->abi_version, KGE_ABI_VERSION);
  KGE_REQUIRE(c->model >= KGE_UM && c->model <= KGE_RESCAL, KGE_EINVAL, "bad model id %d", c->model);
  KGE_REQUIRE(c->d_ent > 0 && c->d_rel > 0, KGE_EINVAL, "embedding sizes must be positive");
  KGE_REQUIRE(c->n_ent > 0 && c->n_rel > 0, KGE_EINVAL, "n_ent / n_rel must be positive");
  KGE_REQUIRE((uint64_t) c->n_ent + (uint64_t) c->n_rel < 0xffffffffull, KGE_EUNSUPPORTED,
              "n_ent + n_rel must fit 32-bit sort keys");
  KGE_REQUIRE(c->id_dtype == 0 || c->id_dtype == 1, KGE_EINVAL, "id_dtype must be 0 (int32) or 1 (int64)");
  KGE_REQUIRE(c->optimizer >= KGE_OPTIM_ADAM && c->optimizer <= KGE_OPTIM_ADAGRAD, KGE_EINVAL, "bad optimizer id %d", c->optimizer);
  if (uses_score_fn(c->model)) {
    KGE_REQUIRE(c->score >= KGE_SCORE_LP && c->score <= KGE_SCORE_DOT, KGE_EINVAL, "bad score id %d", c->score);
    if (c->score != KGE_SCORE_DOT)
      KGE_REQUIRE(c->score_p > 0.f, KGE_EINVAL, "LpDistance p must be > 0 (got %f)", c->score_p);
    KGE_REQUIRE(!(c->model == KGE_ROTATE && c->score == KGE_SCORE_DOT), KGE_EUNSUPPORTED,
                "RotatE with Dot is not a real-valued score (tests/test_integration.py:47 excludes it)");
  }
  return KGE_OK;
}

inline int check_tables(const kge_cfg* c, const kge_tables* t, bool need_opt, int64_t ent_rows = -1) {
  KGE_REQUIRE(t != nullptr, KGE_EINVAL, "tables is NULL");
  int cols[KGE_MAX_TABLES];
  expected_cols(c, cols);
  for (int s = 0; s < KGE_MAX_TABLES; ++s) {
    const kge_table& T = t->t[s];
    if (cols[s] == 0) continue;
    KGE_REQUIRE(T.w != nullptr, KGE_EINVAL, "table slot %d: weights missing", s);
    KGE_REQUIRE(T.cols == cols[s], KGE_EINVAL, "table slot %d: cols %d, expected %d", s, T.cols, cols[s]);
    KGE_REQUIRE(T.rows == (s < 2 ? (ent_rows >= 0 ? ent_rows : c->n_ent) : c->n_rel), KGE_EINVAL, "table slot %d: rows %lld mismatch", s,
                (long long) T.rows);
    KGE_REQUIRE(T.ld >= T.cols && T.ld % 4 == 0, KGE_EINVAL, "table slot %d: ld %d must be >= cols and %% 4 == 0", s, T.ld);
    KGE_REQUIRE(((uintptr_t) T.w & 15) == 0, KGE_EINVAL, "table slot %d: base not 16-byte aligned", s);
    if (need_opt) {
      KGE_REQUIRE(T.m != nullptr && T.v != nullptr, KGE_EINVAL, "table slot %d: Adam slots missing", s);
      KGE_REQUIRE((((uintptr_t) T.m | (uintptr_t) T.v) & 15) == 0, KGE_EINVAL, "table slot %d: m/v not aligned", s);
    }
  }
  return KGE_OK;
}

inline void fill_model_dev(const kge_cfg* c, const kge_tables* t, ModelDev& md) {
  md.model = c->model;
  md.constraint = c->constraint;
  md.d_ent = c->d_ent;
  md.d_rel = c->d_rel;
  md.cw = c->constraint_weight;
  md.rotate_scale = c->rotate_scale;
  md.sc.pow2 = (c->score == KGE_SCORE_LP_POW);
  md.sc.p = c->score_p;
  if (c->score == KGE_SCORE_DOT) md.sc.kind = SC_DOT;
  else if (isinf(c->score_p)) md.sc.kind = SC_LINF;
  else if (c->score_p == 1.0f) md.sc.kind = SC_L1;
  else if (c->score_p == 2.0f) md.sc.kind = SC_L2;
  else md.sc.kind = SC_LPGEN;
  md.shard_rows = 0;
  for (int s = 0; s < KGE_MAX_TABLES; ++s) {
    md.t[s].w = t->t[s].w;
    md.t[s].rows = t->t[s].rows;
    md.t[s].cols = t->t[s].cols;
    md.t[s].ld = t->t[s].ld;
  }
}

// float4 chunks per lane for the widest vector row of the model; 0 = too wide
inline int pick_nv(const kge_cfg* c) {
  int cols[KGE_MAX_TABLES];
  expected_cols(c, cols);
  int mx = 0;
  for (int s = 0; s < KGE_MAX_TABLES; ++s) mx = cols[s] > mx ? cols[s] : mx;
  const int chunks = (mx + 3) / 4;
  if (chunks <= 32) return 1;
  if (chunks <= 64) return 2;
  if (chunks <= 128) return 4;
  if (chunks <= 256) return 8;
  return 0;
}

// strided layout of the matrix models: 128 floats per NV
inline int pick_nv_strided(const kge_cfg* c) {
  const int mx = c->d_ent > c->d_rel ? c->d_ent : c->d_rel;
  return mx <= 128 ? 1 : (mx <= 256 ? 2 : 0);
}

inline int k_eff_of(const kge_cfg* c) {
  return c->corrupt_side == KGE_SIDE_HT ? 2 * (c->negative_ratio / 2) : c->negative_ratio;
}

inline void fill_sampler_dev(const kge_cfg* c, const kge_sampler* s, uint64_t seed, uint64_t step, SamplerDev& sd) {
  sd.kind = c->sampler;
  sd.side_mode = c->corrupt_side;
  sd.k_eff = k_eff_of(c);
  sd.k_side = c->corrupt_side == KGE_SIDE_HT ? c->negative_ratio / 2 : c->negative_ratio;
  sd.seed = seed;
  sd.step = step;
  sd.step_dyn = nullptr;
  sd.n_ent = c->n_ent;
  sd.i_base = 0;
  if (s) sd.s = *s; else { sd.s = kge_sampler{}; }
}

inline int check_sampler(const kge_cfg* c, const kge_sampler* s) {
  if (c->sampler == KGE_NS_TYPED) {
    KGE_REQUIRE(s && s->ent_type && s->type_off && s->type_members && s->ent_pos, KGE_EINVAL,
                "typed sampler needs ent_type / type_off / type_members / ent_pos");
  } else {
    KGE_REQUIRE(c->sampler == KGE_NS_UNIFORM, KGE_EINVAL, "bad sampler id %d", c->sampler);
    KGE_REQUIRE(!(s && s->pool) || s->pool_len > 0, KGE_EINVAL, "uniform sampler pool is empty");
    KGE_REQUIRE(!(s && s->pool) || (uint64_t) s->pool_len < 0xffffffffull, KGE_EUNSUPPORTED, "uniform sampler pool too long");
  }
  return KGE_OK;
}

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

}  // namespace kge
