// train_step.cu — host side of kge_train_step / kge_sample_negatives /
// kge_constrain_table: validation, workspace carve-up, kernel sequencing.
// Replaces KGEModel.__run_single_batch (BaseModel.py:293-330).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "host_common.cuh"
#include <stdlib.h>
#include "step_launch.cuh"
#include "update_kernels.cuh"
#include "group_kernels.cuh"
#include "update_mat_kernels.cuh"
#include "step_mat_launch.cuh"
#include "lazy_kernels.cuh"
#include "shard_kernels.cuh"

namespace kge {

struct StepWs {
  size_t off_part, off_scal, off_neg, off_keys_in, off_keys_out, off_vals_in, off_vals_out, off_cub,
      off_rec[KGE_MAX_TABLES], off_recP, off_snap[2], off_rowpart, off_matpart, off_seg, off_hdr, off_heavy, off_uniq,
      off_islot, off_recdest, off_facq, off_fcoef, off_long;
  size_t cub_bytes, total, n_seg;
  size_t gdelta;          // the grouping buffers (negatives, keys, record ids, scan, headers, heavy list) exist twice:
                          // set s lives at offset + s * gdelta (the next batch is grouped while this one is applied)
  int64_t n_ent_rec, n_keys, total_rows, n_mat_cta;
  int rld[KGE_MAX_TABLES];
  int ldp;
};

static int sort_bits(const kge_cfg* c) {
  uint64_t mx = (uint64_t) c->n_ent + (uint64_t) c->n_rel;
  int b = 1;
  while ((1ull << b) < mx && b < 32) ++b;
  return b;
}

// which table slots hold per-relation matrices (updated by k_update_mat, not k_update)
static bool is_matrix_slot(const kge_cfg* c, int s) {
  switch (c->model) {
    case KGE_SE: return s == KGE_T_REL_A || s == KGE_T_REL_B;
    case KGE_TRANSR: return s == KGE_T_REL_B;
    case KGE_RESCAL: return s == KGE_T_REL_A;
    default: return false;
  }
}
// dense regulariser on every row: TransH.py:204-211, RESCAL.py:191-198
static bool has_dense_reg(const kge_cfg* c) {
  return c->constraint && (c->model == KGE_TRANSH || c->model == KGE_RESCAL);
}
// the shadow table of kge_tables.wn_valid: entity table of the models whose start-of-batch
// constraint is "normalise every entity row" (UM.py:153, TransE.py:171, DistMult.py:162)
static bool shadow_capable(const kge_ctx_impl* ctx, const kge_cfg* c, const kge_tables* t) {
  return !ctx->opt_no_shadow && !ctx->opt_radix_sort && c->constraint && t->t[KGE_T_ENT].wn != nullptr &&
         t->t[KGE_T_ENT].last == nullptr &&
         (c->model == KGE_UM || c->model == KGE_TRANSE || c->model == KGE_DISTMULT);
}
static int mat_rows(const kge_cfg* c) { return c->d_ent; }
static int mat_cols(const kge_cfg* c) { return c->model == KGE_TRANSR ? c->d_rel : c->d_ent; }
static int mat_tiles(const kge_cfg* c) {
  const int tc = mu_tile_cols(mat_cols(c));
  return ((mat_rows(c) + MU_TILE - 1) / MU_TILE) * ((mat_cols(c) + tc - 1) / tc);
}

int layout_step_ws(const kge_cfg* c, int64_t B, StepWs& L) {
  const int k = k_eff_of(c);
  int cols[KGE_MAX_TABLES];
  expected_cols(c, cols);
  L.n_ent_rec = 2 * B + B * (int64_t) k;
  L.n_keys = L.n_ent_rec + B;
  KGE_REQUIRE(L.n_keys < 0x7fffffffll, KGE_EUNSUPPORTED, "batch too large: %lld gradient records", (long long) L.n_keys);
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t at = o; o += align_up(bytes); return at; };
  L.off_part = take(sizeof(float) * (N_PART * (size_t) B + N_PART_MAX * RED_CTAS));      // partials | k_reduce's per-CTA sums
  L.off_scal = take(sizeof(float) * 16);
  L.n_seg = (size_t) c->n_ent + (size_t) c->n_rel;
  const size_t g0 = o;
  L.off_neg = take(sizeof(int64_t) * (size_t) B * (k > 0 ? k : 1));
  L.off_keys_in = take(sizeof(uint32_t) * (size_t) L.n_keys);
  L.off_vals_in = take(sizeof(uint32_t) * (size_t) L.n_keys);
  L.off_vals_out = take(sizeof(uint32_t) * (size_t) L.n_keys);
  L.off_seg = take(sizeof(uint32_t) * (2 * L.n_seg + 2));   // cnt[K] | off[K+1] | n_heavy   (radix path: seg_lo | seg_hi)
  L.off_hdr = take(sizeof(uint32_t) * 8 * L.n_seg);         // k_canon's per-key headers
  L.off_heavy = take(sizeof(uint32_t) * (size_t) (L.n_keys / UPD_HEAVY + 1));
  L.off_long = take(sizeof(uint32_t) * (size_t) (L.n_keys / 32 + 2));          // k_group_cluster: [count | keys with > 32 records]
  L.gdelta = o - g0;
  o += L.gdelta;                                            // the second set
  L.off_keys_out = take(sizeof(uint32_t) * (size_t) L.n_keys);
  // temp size of the radix-sort mode: queried once per (n_keys, bits), this runs on every call
  static thread_local int64_t memo_n = -1;
  static thread_local int memo_bits = -1;
  static thread_local size_t memo_bytes = 0;
  cudaError_t e = cudaSuccess;
  if (memo_n != L.n_keys || memo_bits != sort_bits(c)) {
    size_t bytes = 0;
    e = cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*) nullptr, (uint32_t*) nullptr,
                                        (const uint32_t*) nullptr, (uint32_t*) nullptr, (int) L.n_keys, 0, sort_bits(c),
                                        (cudaStream_t) 0);
    KGE_REQUIRE(e == cudaSuccess, KGE_ECUDA, "cub temp-size query failed: %s", cudaGetErrorString(e));
    memo_n = L.n_keys; memo_bits = sort_bits(c); memo_bytes = bytes;
  }
  L.cub_bytes = memo_bytes;
  L.off_cub = take(L.cub_bytes);
  L.off_uniq = take(sizeof(uint32_t) * (4 + 3 * (size_t) L.n_keys));   // row-lazy work list (k_heads)
  L.off_islot = take(sizeof(uint32_t) * ((size_t) L.n_keys + 16));      // sharded step: inbox slot per work item | per-owner counters
  L.off_recdest = take(sizeof(uint32_t) * (size_t) L.n_ent_rec);        // sharded step: inbox destination of single-record rows
  size_t scan_bytes = 0;
  if ((int64_t) L.n_seg > SCAN_SMALL_MAX) {
    e = cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const uint32_t*) nullptr, (uint32_t*) nullptr, (int64_t) L.n_seg,
                                      (cudaStream_t) 0);
    KGE_REQUIRE(e == cudaSuccess, KGE_ECUDA, "cub scan temp-size query failed: %s", cudaGetErrorString(e));
    if (scan_bytes > L.cub_bytes) { /* the scan reuses the radix sort's temp region */
      KGE_REQUIRE(false, KGE_EUNSUPPORTED, "scan temp storage exceeds the sort's (%zu > %zu)", scan_bytes, L.cub_bytes);
    }
  }
  L.total_rows = 0;
  for (int s = 0; s < KGE_MAX_TABLES; ++s) {
    L.rld[s] = (cols[s] + 3) / 4 * 4;
    const int64_t nrec = (s < 2) ? L.n_ent_rec : B;
    const bool vec = cols[s] && !is_matrix_slot(c, s);
    L.off_rec[s] = vec ? take(sizeof(float) * (size_t) nrec * L.rld[s]) : 0;
    if (vec) L.total_rows += (s < 2) ? c->n_ent : c->n_rel;
  }
  L.off_facq = L.off_fcoef = 0;
  if (c->model == KGE_DISTMULT) {                            // factored records (StepParams.fac_coef)
    L.off_facq = take(sizeof(float) * 2 * (size_t) B * L.rld[KGE_T_ENT]);
    L.off_fcoef = take(sizeof(float) * 2 * (size_t) B * (k > 0 ? k : 1));   // {coefficient, q-row index} per negative
  }
  L.ldp = 0; L.off_recP = 0; L.n_mat_cta = 0;
  if (is_matrix_model(c->model)) {
    L.ldp = (mat_cols(c) > mat_rows(c) ? mat_cols(c) : mat_rows(c));
    L.ldp = (L.ldp + 3) / 4 * 4;
    L.off_recP = take(sizeof(float) * (size_t) L.n_ent_rec * L.ldp);
    L.n_mat_cta = (int64_t) mat_tiles(c) * c->n_rel * (c->model == KGE_SE ? 2 : 1);
  }
  L.off_snap[0] = L.off_snap[1] = L.off_rowpart = L.off_matpart = 0;
  if (has_dense_reg(c)) {
    if (c->model == KGE_TRANSH) {
      L.off_snap[0] = take(sizeof(float) * (size_t) c->n_rel * L.rld[KGE_T_REL_A]);
      L.off_snap[1] = take(sizeof(float) * (size_t) c->n_rel * L.rld[KGE_T_REL_B]);
    }
    L.off_rowpart = take(sizeof(float) * 2 * (size_t) L.total_rows);
    if (L.n_mat_cta) L.off_matpart = take(sizeof(float) * 2 * (size_t) L.n_mat_cta);
  }
  L.total = o;
  return KGE_OK;
}

// row-lazy replay (kge_tables.lazy_t): requested by `last` on the entity table
static bool lazy_requested(const kge_tables* t) { return t->t[KGE_T_ENT].last != nullptr; }
static int check_lazy(const kge_cfg* c, const kge_tables* t) {
  KGE_REQUIRE(!is_matrix_model(c->model) && !has_dense_reg(c), KGE_EUNSUPPORTED,
              "row-lazy tables: vector models without a dense regulariser only");
  int cols[KGE_MAX_TABLES];
  expected_cols(c, cols);
  for (int s = 0; s < 2; ++s)
    KGE_REQUIRE(!cols[s] || t->t[s].last, KGE_EINVAL, "row-lazy tables: entity slot %d has no `last`", s);
  KGE_REQUIRE(t->lazy_t >= 0 && t->lazy_t < 0xfffffff0ll, KGE_EINVAL, "lazy_t %lld out of range", (long long) t->lazy_t);
  return KGE_OK;
}
// start-of-batch constraint op of an entity slot (the jobs k_lazy_pre takes over from k_constrain)
static int lazy_con_op(const kge_cfg* c, int slot) {
  if (!c->constraint || slot != KGE_T_ENT) return LAZY_CON_NONE;
  switch (c->model) {
    case KGE_UM: case KGE_TRANSE: case KGE_DISTMULT: return CON_NORMALIZE;
    case KGE_TRANSD: return CON_CLIP;
    default: return LAZY_CON_NONE;
  }
}
static void fill_lazy(const kge_cfg* c, const kge_tables* t, uint32_t step_t, LazyParams& lp) {
  int cols[KGE_MAX_TABLES];
  expected_cols(c, cols);
  lp.n_tabs = 0;
  for (int s = 0; s < 2; ++s) {
    if (!cols[s]) continue;
    LazyTable& T = lp.tab[lp.n_tabs++];
    T.w = t->t[s].w; T.m = t->t[s].m; T.v = t->t[s].v; T.last = t->t[s].last;
    T.rec = nullptr; T.rows = t->t[s].rows; T.cols = t->t[s].cols; T.ld = t->t[s].ld; T.rec_ld = 0;
    T.slot = s; T.con_op = lazy_con_op(c, s);
  }
  lp.uniq = nullptr; lp.vals = nullptr; lp.scal = nullptr;
  lp.lr = c->lr; lp.beta1 = c->beta1; lp.beta2 = c->beta2; lp.eps = c->eps;
  lp.opt = c->optimizer;
  lp.t = step_t;
}

// ---- sharded entity tables (kge_shard) ---------------------------------------------------------
static int64_t shard_local_rows(const kge_cfg* c, const kge_shard* sh) {
  const int64_t lo = (int64_t) sh->rank * sh->shard_rows;
  int64_t hi = lo + sh->shard_rows;
  if (hi > c->n_ent) hi = c->n_ent;
  return hi > lo ? hi - lo : 0;
}
static int check_shard(const kge_cfg* c, const kge_tables* t, const kge_shard* sh, int64_t B) {
  KGE_REQUIRE(sh, KGE_EINVAL, "shard is NULL");
  KGE_REQUIRE(sh->world >= 1 && sh->world <= KGE_MAX_PEERS && sh->rank >= 0 && sh->rank < sh->world, KGE_EINVAL,
              "bad kge_shard (world %d rank %d)", sh->world, sh->rank);
  KGE_REQUIRE(sh->shard_rows > 0 && sh->shard_rows < 0xffffffffll && sh->shard_rows * sh->world >= c->n_ent, KGE_EINVAL,
              "shard_rows %lld x world %d does not cover n_ent %lld", (long long) sh->shard_rows, sh->world, (long long) c->n_ent);
  KGE_REQUIRE(sh->cap > 0 && sh->cap < (1ll << 29) / (sh->world > 0 ? sh->world : 1) && (B <= 0 || sh->cap >= 2 * B + B * (int64_t) k_eff_of(c)), KGE_EINVAL,
              "shard.cap %lld < looked-up rows of a batch of %lld", (long long) sh->cap, (long long) B);
  int cols[KGE_MAX_TABLES];
  expected_cols(c, cols);
  for (int j = 0; j < sh->world; ++j) {
    KGE_REQUIRE(sh->ids_peer[j] && sh->cnt_peer[j], KGE_EINVAL, "shard inbox pointers of rank %d missing", j);
    for (int s2 = 0; s2 < 2; ++s2)
      if (cols[s2]) KGE_REQUIRE(sh->w_peer[j][s2] && sh->val_peer[j][s2] && (((uintptr_t) sh->w_peer[j][s2] | (uintptr_t) sh->val_peer[j][s2]) & 15) == 0,
                                KGE_EINVAL, "shard.w_peer / val_peer [%d][%d] missing or not 16-byte aligned", j, s2);
  }
  KGE_REQUIRE(sh->stamp && sh->idx_of && sh->work && sh->contrib, KGE_EINVAL, "shard owner scratch missing");
  for (int s2 = 0; s2 < 2; ++s2)
    if (cols[s2]) {
      KGE_REQUIRE(t->t[s2].w == sh->w_peer[sh->rank][s2], KGE_EINVAL, "tables slot %d is not this rank's shard", s2);
      KGE_REQUIRE(t->t[s2].ld == (cols[s2] + 3) / 4 * 4, KGE_EINVAL, "sharded tables: ld of slot %d must be cols rounded up to 4", s2);
    }
  return KGE_OK;
}
static void fill_shard_dev(const kge_shard* sh, ShardDev& sd) {
  sd.world = sh->world; sd.rank = sh->rank; sd.S = (uint32_t) sh->shard_rows; sd.cap = (uint32_t) sh->cap;
  for (int j = 0; j < KGE_MAX_PEERS; ++j) {
    sd.ids_peer[j] = sh->ids_peer[j]; sd.cnt_peer[j] = sh->cnt_peer[j];
    sd.val_peer[j][0] = sh->val_peer[j][0]; sd.val_peer[j][1] = sh->val_peer[j][1];
  }
  sd.stamp = sh->stamp; sd.idx_of = sh->idx_of; sd.work = sh->work; sd.contrib = sh->contrib;
}

// constraint jobs of _constraint_loss per model; skip_ent: the entity slots are
// row-lazy (k_lazy_pre applies their constraint to the looked-up rows)
static int constraint_jobs(const kge_cfg* c, const kge_tables* t, ConstrainParams& cp, bool skip_ent = false) {
  cp.n_jobs = 0;
  cp.total_rows = 0;
  if (!c->constraint) return 0;
  auto add = [&](int slot, int op) {
    if (skip_ent && slot < 2) return;
    ConstrainJob& J = cp.job[cp.n_jobs++];
    J.w = t->t[slot].w; J.rows = t->t[slot].rows; J.cols = t->t[slot].cols; J.ld = t->t[slot].ld; J.op = op;
    cp.total_rows += J.rows;
  };
  switch (c->model) {
    case KGE_UM: case KGE_SE: case KGE_TRANSE: case KGE_DISTMULT: add(KGE_T_ENT, CON_NORMALIZE); break;
    case KGE_TRANSH: add(KGE_T_REL_B, CON_NORMALIZE); break;
    case KGE_TRANSR: case KGE_TRANSD: add(KGE_T_ENT, CON_CLIP); add(KGE_T_REL_A, CON_CLIP); break;
    default: break;
  }
  return cp.n_jobs;
}

static int launch_constrain(const kge_ctx_impl* ctx, const ConstrainParams& cp, cudaStream_t st) {
  if (cp.n_jobs == 0) return KGE_OK;
  if (launch_constrain_kernel(cp, ctx->sm_count, st) != KGE_OK) { set_error("k_constrain launch failed"); return KGE_ECUDA; }
  return KGE_OK;
}

// sort keys, record ids and the negatives used, one thread per record slot; runs on the side
// stream concurrently with k_step (the keys depend on the ids only, not on any weight)
__global__ void __launch_bounds__(256) k_keys(const StepParams p, uint32_t* cnt) {
  const int64_t nneg = p.B * (int64_t) p.k_eff, nE = 2 * p.B + nneg;
  for (int64_t s = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; s < nE + p.B; s += (int64_t) gridDim.x * blockDim.x) {
    uint32_t key;
    if (s < 2 * p.B) key = (uint32_t) load_id(p.triples, 3 * (s % p.B) + (s < p.B ? 0 : 2), p.id64);
    else if (s < nE) {
      const int64_t slot = s - 2 * p.B, i = slot / p.k_eff;
      const int c = (int) (slot - i * p.k_eff);
      const int64_t e = p.neg_override ? p.neg_override[slot]
                                       : sample_negative(p.sd, i, c, load_id(p.triples, 3 * i, p.id64),
                                                         load_id(p.triples, 3 * i + 2, p.id64));
      p.neg_ids[slot] = e;
      key = (uint32_t) e;
    } else key = (uint32_t) (p.md.t[KGE_T_ENT].rows + load_id(p.triples, 3 * (s - nE) + 1, p.id64));
    p.keys[s] = key;
    p.vals[s] = (uint32_t) s;
    if (cnt) atomicAdd(cnt + key, 1u);          // histogram of the counting sort (group_kernels.cuh)
  }
}


// the same key function for k_group_cluster (group_kernels.cuh)
struct StepKeyFn {
  StepParams p;
  __device__ __forceinline__ uint32_t operator()(int64_t s) const {
    const int64_t nneg = p.B * (int64_t) p.k_eff, nE = 2 * p.B + nneg;
    if (s < 2 * p.B) return (uint32_t) load_id(p.triples, 3 * (s % p.B) + (s < p.B ? 0 : 2), p.id64);
    if (s < nE) {
      const int64_t slot = s - 2 * p.B, i = slot / p.k_eff;
      const int c = (int) (slot - i * p.k_eff);
      const int64_t e = p.neg_override ? p.neg_override[slot]
                                       : sample_negative(p.sd, i, c, load_id(p.triples, 3 * i, p.id64),
                                                         load_id(p.triples, 3 * i + 2, p.id64));
      p.neg_ids[slot] = e;
      return (uint32_t) e;
    }
    return (uint32_t) (p.md.t[KGE_T_ENT].rows + load_id(p.triples, 3 * (s - nE) + 1, p.id64));
  }
};

__global__ void k_sample(SamplerDev sd, const void* triples, int id64, int64_t B, int64_t* out) {
  const int64_t n = (int64_t) B * sd.k_eff;
  for (int64_t s = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (int64_t) gridDim.x * blockDim.x) {
    const int64_t i = s / sd.k_eff;
    const int c = (int) (s - i * sd.k_eff);
    out[s] = sample_negative(sd, i, c, load_id(triples, 3 * i, id64), load_id(triples, 3 * i + 2, id64));
  }
}

}  // namespace kge

using namespace kge;

extern "C" int kge_sample_negatives(kge_ctx* ctx_, const kge_cfg* cfg, const kge_sampler* smp, const void* triples,
                                    int64_t B, uint64_t seed, uint64_t step, int64_t* neg_out, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  if (int rc = check_sampler(cfg, smp)) return rc;
  KGE_REQUIRE(triples && neg_out && B > 0, KGE_EINVAL, "triples / neg_out / B invalid");
  SamplerDev sd;
  fill_sampler_dev(cfg, smp, seed, step, sd);
  if (sd.k_eff == 0) return KGE_OK;
  const int64_t n = B * sd.k_eff;
  int64_t blocks = (n + 255) / 256;
  if (blocks > (int64_t) ctx->sm_count * 16) blocks = (int64_t) ctx->sm_count * 16;
  k_sample<<<(unsigned) blocks, 256, 0, (cudaStream_t) stream>>>(sd, triples, cfg->id_dtype, B, neg_out);
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}

extern "C" int kge_constrain_table(kge_ctx* ctx_, const kge_tables* t, int slot, int op, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx && t, KGE_EINVAL, "ctx / tables is NULL");
  KGE_REQUIRE(slot >= 0 && slot < KGE_MAX_TABLES && t->t[slot].w, KGE_EINVAL, "bad table slot %d", slot);
  KGE_REQUIRE(op == CON_NORMALIZE || op == CON_CLIP, KGE_EINVAL, "bad constraint op %d", op);
  KGE_REQUIRE(t->t[slot].ld % 4 == 0 && ((uintptr_t) t->t[slot].w & 15) == 0, KGE_EINVAL, "table not 16-byte strided");
  ConstrainParams cp;
  cp.n_jobs = 1;
  cp.job[0] = ConstrainJob{t->t[slot].w, t->t[slot].rows, t->t[slot].cols, t->t[slot].ld, op};
  cp.total_rows = t->t[slot].rows;
  return launch_constrain(ctx, cp, (cudaStream_t) stream);
}

namespace kge {
size_t rank_workspace_bytes(const kge_cfg* cfg, int64_t Q);   // rank.cu

// warps per positive: as few as still fill the machine (every warp repeats the positive's row
// loads and the cross-warp reduction, so fewer warps = less redundant work), never more than
// the negatives can feed.  Measured on C2 (B=1024, k'=64): 8 -> 60.8 us, 4 -> 48.8, 2 -> 43.9.
static int pick_wpp(const kge_cfg* c, int k, int64_t B, int sm_count) {
  const int by_k = k >= 48 ? 8 : (k >= 12 ? 4 : (k >= 3 ? 2 : 1));
  const int64_t want_warps = (int64_t) sm_count * 16 * 3 / 4;
  int need = 1;
  while (need < 8 && B * need < want_warps) need <<= 1;
  // DistMult's fast negative loop needs one corrupted side per warp: 'h+t' wants an even count
  if (c->model == KGE_DISTMULT && c->corrupt_side == KGE_SIDE_HT && need < 2 && by_k >= 2) need = 2;
  return need < by_k ? need : by_k;
}
}  // namespace kge

extern "C" int kge_workspace_bytes(const kge_cfg* cfg, int64_t B, int64_t Q, size_t* out) {
  if (int rc = check_cfg(cfg)) return rc;
  KGE_REQUIRE(out, KGE_EINVAL, "out is NULL");
  size_t need = 256;
  if (B > 0) {
    StepWs L;
    if (int rc = layout_step_ws(cfg, B, L)) return rc;
    need = L.total > need ? L.total : need;
  }
  if (Q > 0) {
    size_t r = rank_workspace_bytes(cfg, Q);
    need = r > need ? r : need;
  }
  *out = need;
  return KGE_OK;
}

namespace kge {
// Captured step graph, look-ahead chain: the next batch is fetched by the graph itself from wherever the launch says
// (DynScal.next_src: pinned host memory read over PCIe, or device memory) into its staging buffer — a copy queued in
// front of the graph launch would sit on the step's critical path (5 us per step on C2).
__global__ void __launch_bounds__(256) k_stage_next(uint32_t* __restrict__ dst, const DynScal* d, size_t words) {
  const uint32_t* src = reinterpret_cast<const uint32_t*>(d->next_src);
  if (src == nullptr) return;
  for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (size_t) gridDim.x * blockDim.x)
    dst[i] = src[i];
}
}  // namespace kge

// dp / grads_out / scal_out non-NULL = the gradient half of a data-parallel step (kge_train_grad)
static int train_impl(kge_ctx* ctx_, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                      const kge_dp* dp, const void* triples, int64_t B, uint64_t seed, uint64_t step, uint64_t adam_t,
                      const int64_t* neg_override, int is_train, kge_grads* grads_out, float* scal_out,
                      float* loss_out, int64_t* neg_out, void* ws, size_t ws_bytes, void* stream,
                      const kge_shard* sh = nullptr, int shard_phase = 0 /* 1 lookup, 2 grad */,
                      const void* next_triples = nullptr, uint64_t next_step = 0, const void* next_host = nullptr) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  cudaStream_t st = (cudaStream_t) stream;
  const bool grad_only = dp != nullptr;
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  if (sh) { if (int rc = check_shard(cfg, tables, sh, B)) return rc; }
  if (int rc = check_tables(cfg, tables, (is_train != 0 && dp == nullptr) || (tables && lazy_requested(tables)),
                            sh ? shard_local_rows(cfg, sh) : -1)) return rc;
  // row-lazy entity tables: a training call works on the looked-up rows only; a validation call
  // materialises the tables first and then takes the dense path
  bool lazy = lazy_requested(tables);
  const bool flush_first = lazy && !is_train;
  if (lazy) {
    if (int rc = check_lazy(cfg, tables)) return rc;
    KGE_REQUIRE(dp == nullptr || sh, KGE_EUNSUPPORTED, "row-lazy tables: the data-parallel halves are not implemented");
    KGE_REQUIRE(!is_train || sh || (int64_t) adam_t == tables->lazy_t + 1, KGE_EINVAL,
                "row-lazy tables: adam_t %llu must be lazy_t + 1 = %lld", (unsigned long long) adam_t,
                (long long) tables->lazy_t + 1);
    if (!is_train) lazy = false;
  }
  KGE_REQUIRE(!sh || (lazy && dp), KGE_EINVAL, "sharded step: entity slots need `last`, and kge_dp must be given");
  if (!neg_override) { if (int rc = check_sampler(cfg, smp)) return rc; }
  KGE_REQUIRE(triples && (loss_out || grad_only) && ws && B > 0, KGE_EINVAL, "triples / loss_out / ws / B invalid");
  KGE_REQUIRE(cfg->loss >= KGE_LOSS_HINGE && cfg->loss <= KGE_LOSS_SQUARE, KGE_EINVAL, "bad loss id %d", cfg->loss);
  KGE_REQUIRE(cfg->corrupt_side >= KGE_SIDE_H && cfg->corrupt_side <= KGE_SIDE_HT, KGE_EINVAL, "bad corrupt_side");
  const int k = k_eff_of(cfg);
  KGE_REQUIRE(k > 0, KGE_EINVAL,
              "zero negatives per positive (corrupt_side='h+t' uses negative_ratio//2 per side, BaseModel.py:354-355)");
  KGE_REQUIRE(!is_train || adam_t >= 1, KGE_EINVAL, "adam_t is 1-based");
  const bool mat = is_matrix_model(cfg->model);
  int64_t B_global = B;
  if (grad_only) {
    static kge_grads no_grads = {};
    if (sh && shard_phase == 1) grads_out = &no_grads;
    KGE_REQUIRE(grads_out, KGE_EINVAL, "grads is NULL");
    KGE_REQUIRE(!sh || grads_out->n_peers == 0, KGE_EUNSUPPORTED, "sharded step: relation gradients use grads.dense");
    KGE_REQUIRE(grads_out->n_peers >= 0 && grads_out->n_peers <= KGE_MAX_PEERS, KGE_EINVAL, "grads.n_peers %d out of range",
                grads_out->n_peers);
    KGE_REQUIRE(grads_out->n_peers > 0 || scal_out || shard_phase == 1, KGE_EINVAL, "scal_out is NULL");
    KGE_REQUIRE(dp->world >= 1 && dp->rank >= 0 && dp->rank < dp->world && dp->i_base >= 0 &&
                    dp->B_global >= dp->i_base + B, KGE_EINVAL, "bad kge_dp (world %d rank %d B_global %lld i_base %lld)",
                dp->world, dp->rank, (long long) dp->B_global, (long long) dp->i_base);
    // matrix models: the relation matrices' gradients travel as dense [n_rel x d d'] tables (grads.dense: all-reduce);
    // the peer / owner exchanges and the dense regulariser of RESCAL (constraint = True) stay single-GPU
    KGE_REQUIRE(!mat || (grads_out->n_peers == 0 && !has_dense_reg(cfg) && !sh), KGE_EUNSUPPORTED,
                "data-parallel SE / TransR / RESCAL: gradients through grads.dense only, no dense regulariser (else single-GPU)");
    int gcols[KGE_MAX_TABLES];
    expected_cols(cfg, gcols);
    for (int s = 0; s < KGE_MAX_TABLES; ++s)
      if (gcols[s] && grads_out->owner_rows[s] != 0)
        KGE_REQUIRE(grads_out->n_peers > 0 && grads_out->owner_rows[s] > 0 &&
                        (int64_t) grads_out->owner_rows[s] * grads_out->n_peers >= (s < 2 ? cfg->n_ent : cfg->n_rel),
                    KGE_EINVAL, "grads.owner_rows[%d] = %d does not cover the table over %d ranks", s,
                    grads_out->owner_rows[s], grads_out->n_peers);
    for (int s = 0; s < KGE_MAX_TABLES; ++s) {
      if (!gcols[s]) continue;
      if (sh && (s < 2 || shard_phase == 1)) continue;
      if (grads_out->n_peers == 0)
        KGE_REQUIRE(grads_out->dense[s] && ((uintptr_t) grads_out->dense[s] & 15) == 0, KGE_EINVAL,
                    "grads.dense[%d] missing or not 16-byte aligned", s);
      for (int j = 0; j < grads_out->n_peers; ++j)
        KGE_REQUIRE(grads_out->peer[j][s] && ((uintptr_t) grads_out->peer[j][s] & 15) == 0 && grads_out->peer_scal[j],
                    KGE_EINVAL, "grads.peer[%d][%d] / peer_scal missing or not 16-byte aligned", j, s);
    }
    B_global = dp->B_global;
  }
  const int nv = mat ? pick_nv_strided(cfg) : pick_nv(cfg);
  KGE_REQUIRE(nv > 0, KGE_EUNSUPPORTED, mat ? "matrix models support embedding sizes up to 256" : "row wider than 1024 floats");

  StepWs L;
  if (int rc = layout_step_ws(cfg, B, L)) return rc;
  KGE_REQUIRE(ws_bytes >= L.total, KGE_EWORKSPACE, "workspace %zu < required %zu bytes", ws_bytes, L.total);
  KGE_REQUIRE(((uintptr_t) ws & 255) == 0, KGE_EINVAL, "workspace must be 256-byte aligned");
  char* wsb = reinterpret_cast<char*>(ws);
  int cols[KGE_MAX_TABLES];
  expected_cols(cfg, cols);

  // Grouping of the NEXT batch under this step (kge_train_step_next): the record keys depend on ids and the
  // Philox stream only, so the previous call may already have grouped THIS batch into one of the two sets.
  const bool can_pre = is_train && !lazy && !sh && !ctx->opt_radix_sort && (int64_t) L.n_seg <= SCAN_SMALL_MAX && !neg_override;
  Prefetched& pre = ctx->pre;
  const bool pre_hit = can_pre && pre.valid && pre.triples == triples && pre.B == B && pre.seed == seed && pre.step == step &&
                       pre.ws == ws && pre.i_base == (dp ? dp->i_base : 0) && pre.k == k && pre.n_ent == cfg->n_ent &&
                       pre.side == cfg->corrupt_side && pre.sampler == cfg->sampler && pre.graph == ctx->capturing;
  // steps that span several calls (sharded: lookup / grad / apply) or never look ahead always use set 0
  const int gs = !can_pre ? 0 : (pre_hit ? pre.set : (pre.valid ? 1 - pre.set : 0));
  if (!pre_hit && pre.valid && is_train && !pre.graph && !ctx->capturing)   // a grouped batch nobody asked for: its chain
    KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_pre_done, 0));               // must finish before its set is reused
  if (is_train) pre.valid = false;
  char* gb = wsb + (size_t) gs * L.gdelta;               // base of this step's grouping set

  // 1. start-of-batch constraint (train AND validation batches, BaseModel.py:319): enqueued by
  //    fork_side() below, once every argument check has passed
  PhaseTimer pt(ctx, st);
  ConstrainParams cp;
  constraint_jobs(cfg, tables, cp, lazy);
  KGE_REQUIRE(!tables->t[KGE_T_ENT].wn || ((uintptr_t) tables->t[KGE_T_ENT].wn & 15) == 0, KGE_EINVAL, "wn not 16-byte aligned");
  const bool shadow_ok = shadow_capable(ctx, cfg, tables);
  // valid shadow + training call: rows are gathered from wn and the full-table pass is skipped
  // (a validation call still assigns the normalised rows to w, as the reference does)
  const bool shadow_use = shadow_ok && tables->wn_valid[KGE_T_ENT] && is_train;
  if (shadow_use) { cp.n_jobs = 0; cp.total_rows = 0; }
  // a validation call rewrites w with its normalised rows: wn = normalise(old w) is then no longer
  // normalise(w) bit for bit (normalising twice moves the last bits), so the shadow is dropped
  if (!is_train && cp.n_jobs > 0) tables->wn_valid[KGE_T_ENT] = 0;

  // 2. fused sample + gather + score + loss + backward
  MatStepParams mq;
  StepParams& p = mq.s;
  fill_model_dev(cfg, tables, p.md);
  ShardDev sdv;
  if (sh) {
    fill_shard_dev(sh, sdv);
    p.md.shard_rows = (uint32_t) sh->shard_rows;
    for (int s2 = 0; s2 < 2; ++s2) {
      p.md.t[s2].rows = cfg->n_ent;                       // key space and ids stay global
      for (int j = 0; j < KGE_MAX_PEERS; ++j) p.md.shard_w[s2][j] = sh->w_peer[j][s2];
    }
  }
  uint32_t* item_slot = reinterpret_cast<uint32_t*>(wsb + L.off_islot);
  uint32_t* rcnt = item_slot + L.n_keys;
  uint32_t* rec_dest = reinterpret_cast<uint32_t*>(wsb + L.off_recdest);
  p.rec_dest = nullptr;
  if (sh && shard_phase == 2 && !ctx->opt_no_fused_push) {
    p.rec_dest = rec_dest;
    for (int s2 = 0; s2 < 2; ++s2)
      for (int j = 0; j < KGE_MAX_PEERS; ++j)
        p.inbox[s2][j] = (j < sh->world && sh->val_peer[j][s2])
                             ? sh->val_peer[j][s2] + (size_t) sh->rank * (size_t) sh->cap * (size_t) tables->t[s2].ld : nullptr;
  }
  if (shadow_use) p.md.t[KGE_T_ENT].w = tables->t[KGE_T_ENT].wn;
  for (int s = 0; s < KGE_MAX_TABLES; ++s) {
    p.md.rld[s] = L.rld[s];
    p.rec[s] = (cols[s] && !is_matrix_slot(cfg, s)) ? reinterpret_cast<float*>(wsb + L.off_rec[s]) : nullptr;
  }
  p.loss.kind = cfg->loss;
  p.loss.margin = cfg->margin;
  p.loss.temperature = cfg->temperature;
  p.loss.inv_n = 1.0f / (float) B_global;                 // loss normalisations use the whole step's batch
  p.loss.inv_nk = 1.0f / ((float) B_global * (float) k);
  fill_sampler_dev(cfg, smp, seed, step, p.sd);
  // captured step graph: Philox step and lr_t are read from the device block the graph's first node writes
  const DynScal* dyn = ctx->capturing ? ctx->dyn : nullptr;
  KGE_REQUIRE(!dyn || (!lazy && !flush_first && !sh && is_train), KGE_EUNSUPPORTED,
              "step graphs cover dense-table training steps (not row-lazy / sharded tables, not validation)");
  if (dyn) p.sd.step_dyn = &dyn->step;
  if (grad_only) p.sd.i_base = dp->i_base;
  p.triples = triples;
  p.id64 = cfg->id_dtype;
  p.B = B;
  p.k_eff = k;
  p.wpp = mat ? STEP_WARPS : pick_wpp(cfg, k, B, ctx->sm_count);
  // DistMult (rows <= 256 floats): CTA size chosen so that ONE balanced wave covers the batch when it can
  // (C2: 1024 positives x 2 warps over 148 SMs -> 14 warps per CTA, 147 CTAs), else the usual 8 warps
  p.nwarps = STEP_WARPS;
  if (!mat && cfg->model == KGE_DISTMULT && nv <= 2) {
    const int64_t per_sm = (B * p.wpp + ctx->sm_count - 1) / ctx->sm_count;
    if (per_sm <= STEP_MAX_WARPS) {
      p.nwarps = (int) ((per_sm + p.wpp - 1) / p.wpp) * p.wpp;
      if (p.nwarps > STEP_MAX_WARPS) p.nwarps = STEP_MAX_WARPS / p.wpp * p.wpp;
    }
  }
  // factored records: dense (not row-lazy / sharded) training step whose warps each corrupt one side
  const bool fac = !mat && cfg->model == KGE_DISTMULT && is_train && !lazy && !sh &&
                   (cfg->corrupt_side != KGE_SIDE_HT || (p.wpp & 1) == 0);
  p.fac_coef = fac ? reinterpret_cast<float*>(wsb + L.off_fcoef) : nullptr;
  p.fac_q = fac ? reinterpret_cast<float*>(wsb + L.off_facq) : nullptr;
  p.neg_override = neg_override;
  p.neg_ids = reinterpret_cast<int64_t*>(gb + L.off_neg);
  p.keys = reinterpret_cast<uint32_t*>(gb + L.off_keys_in);
  p.vals = reinterpret_cast<uint32_t*>(gb + L.off_vals_in);
  p.part = reinterpret_cast<float*>(wsb + L.off_part);
  // training: keys / record ids / negatives come from k_keys on the side stream
  p.write_keys = is_train ? 0 : 1;
  p.tail = (B <= TAIL_REDUCE_MAX_B) ? 1 : 0;
  p.done = ctx->ticket;
  p.rp.part = p.part;
  p.rp.B = B;
  p.rp.n_part = is_train ? N_PART : 1;
  p.rp.clip_norm = cfg->clip_norm;
  p.rp.scal = reinterpret_cast<float*>(wsb + L.off_scal);
  p.rp.loss_out = grad_only ? p.rp.scal + 9 : loss_out;
  p.rp.stage = p.part + N_PART * (size_t) B;
  p.rp.done = ctx->ticket + 1;
  // captured step graph on one GPU: the kernel that makes the loss final also hands it to the host (DynScal.host_loss)
  const bool dense_reg_loss = has_dense_reg(cfg);
  p.rp.dyn = (dyn && !grad_only && !dense_reg_loss) ? dyn : nullptr;

  // side stream: sort keys depend on ids only, so key generation, the stable radix sort and the
  // segment table run concurrently with the step kernel and join before the update kernels
  uint32_t* keys_out = reinterpret_cast<uint32_t*>(wsb + L.off_keys_out);
  uint32_t* vals_out = reinterpret_cast<uint32_t*>(gb + L.off_vals_out);
  uint32_t* seg_lo = reinterpret_cast<uint32_t*>(gb + L.off_seg);
  uint32_t* seg_hi = seg_lo + L.n_seg;
  // what the update kernels read: radix path = the two tables of k_segments; counting sort =
  // the scan itself (first position of key = off[key], one past the last = off[key + 1])
  const uint32_t* use_lo = (ctx->opt_radix_sort != 0) ? seg_lo : seg_lo + L.n_seg;
  const uint32_t* use_hi = (ctx->opt_radix_sort != 0) ? seg_hi : seg_lo + L.n_seg + 1;
  // row-lazy: segment table of the relation keys only (indexed by key - n_ent), work list for the rest
  uint32_t* uniq = reinterpret_cast<uint32_t*>(wsb + L.off_uniq);
  if (lazy) { seg_hi = seg_lo + cfg->n_rel; use_lo = seg_lo; use_hi = seg_hi; }
  LazyParams lp;
  if (lazy || flush_first) fill_lazy(cfg, tables, (uint32_t) (flush_first ? tables->lazy_t : (int64_t) adam_t), lp);
  // counting sort of one batch's record keys into the grouping set at `gbase`: histogram (in k_keys) -> scan ->
  // scatter -> canonical order inside each segment (group_kernels.cuh); kp carries the batch (triples, Philox step)
  auto group_chain = [&](cudaStream_t ax, const StepParams& kp0, char* gbase) -> int {
    StepParams kp = kp0;
    kp.neg_ids = reinterpret_cast<int64_t*>(gbase + L.off_neg);
    kp.keys = reinterpret_cast<uint32_t*>(gbase + L.off_keys_in);
    kp.vals = reinterpret_cast<uint32_t*>(gbase + L.off_vals_in);
    uint32_t* cnt = reinterpret_cast<uint32_t*>(gbase + L.off_seg);
    uint32_t* off = cnt + L.n_seg;
    const int64_t K = (int64_t) L.n_seg;
    if (ctx->opt_cluster_group && K <= (int64_t) GC_CTAS * GC_MAX_KC) {
      // opt-in (KGE_OPT_CLUSTER_GROUP): the whole chain in one kernel on one thread-block cluster (+ the long segments'
      // pass).  Same results; measured SLOWER inside a step (C2: grouping 65 -> 144 us): its 1024-thread CTAs need
      // eight empty SMs at once, which the step and update kernels never leave, so it runs after them instead of beside them.
      GroupOut go;
      go.keys = kp.keys; go.vals_tmp = kp.vals; go.vals = reinterpret_cast<uint32_t*>(gbase + L.off_vals_out);
      go.off = off; go.hdr = reinterpret_cast<uint32_t*>(gbase + L.off_hdr);
      go.heavy_keys = reinterpret_cast<uint32_t*>(gbase + L.off_heavy); go.n_heavy = cnt + 2 * L.n_seg + 1;
      go.n_long = reinterpret_cast<uint32_t*>(gbase + L.off_long); go.long_keys = go.n_long + 1;
      go.K = K; go.n_keys = L.n_keys; go.Kc = (int) ((K + GC_CTAS - 1) / GC_CTAS);
      const size_t gsm = group_cluster_smem(go.Kc);
      static size_t gsm_set = 0;
      if (gsm > 48 * 1024 && gsm > gsm_set) {
        KGE_CUDA_OK(cudaFuncSetAttribute(k_group_cluster<StepKeyFn>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) gsm));
        gsm_set = gsm;
      }
      StepKeyFn kf{kp};
      k_group_cluster<StepKeyFn><<<GC_CTAS, GC_THREADS, gsm, ax>>>(kf, go);
      KGE_CUDA_OK(cudaGetLastError());
      k_canon_long<<<64, CANON_WARPS * 32, 0, ax>>>(off, go.long_keys, go.n_long, kp.vals, go.vals, go.hdr);
      KGE_CUDA_OK(cudaGetLastError());
      return KGE_OK;
    }
    int64_t kb = (L.n_keys + 255) / 256;
    if (kb > (int64_t) ctx->sm_count * 8) kb = (int64_t) ctx->sm_count * 8;
    KGE_CUDA_OK(cudaMemsetAsync(cnt, 0, sizeof(uint32_t) * (2 * L.n_seg + 2), ax));
    k_keys<<<(unsigned) kb, 256, 0, ax>>>(kp, cnt);
    KGE_CUDA_OK(cudaGetLastError());
    if (K <= SCAN_SMALL_MAX) k_scan<<<1, SCAN_THREADS, 0, ax>>>(cnt, off, K);
    else {
      size_t scan_bytes = L.cub_bytes;
      KGE_CUDA_OK(cub::DeviceScan::ExclusiveSum(wsb + L.off_cub, scan_bytes, (const uint32_t*) cnt, off, K, ax));
      k_scan_close<<<1, 1, 0, ax>>>(cnt, off, K);
    }
    KGE_CUDA_OK(cudaGetLastError());
    k_scatter<<<(unsigned) kb, 256, 0, ax>>>(kp.keys, L.n_keys, off, cnt, kp.vals);
    KGE_CUDA_OK(cudaGetLastError());
    int64_t cb = (K + CANON_WARPS - 1) / CANON_WARPS;
    if (cb > (int64_t) ctx->sm_count * 16) cb = (int64_t) ctx->sm_count * 16;
    k_canon<<<(unsigned) cb, CANON_WARPS * 32, 0, ax>>>(off, K, kp.vals, reinterpret_cast<uint32_t*>(gbase + L.off_vals_out),
                                                        reinterpret_cast<uint32_t*>(gbase + L.off_hdr),
                                                        reinterpret_cast<uint32_t*>(gbase + L.off_heavy), cnt + 2 * L.n_seg + 1);
    KGE_CUDA_OK(cudaGetLastError());
    return KGE_OK;
  };
  // the NEXT batch's grouping, on its own stream, behind everything enqueued on `st` so far (the previous step,
  // which last used that set, and whatever produced next_triples)
  auto prefetch_next = [&]() -> int {
    if (!can_pre || next_triples == nullptr) return KGE_OK;
    cudaStream_t px = ctx->pre_stream;
    KGE_CUDA_OK(cudaEventRecord(ctx->ev_pre_fork, st));
    KGE_CUDA_OK(cudaStreamWaitEvent(px, ctx->ev_pre_fork, 0));
    if (next_host)          // the next batch comes from host memory: its copy rides the same stream, ahead of the chain
      KGE_CUDA_OK(cudaMemcpyAsync(const_cast<void*>(next_triples), next_host, (size_t) B * 3 * (cfg->id_dtype ? 8 : 4),
                                  cudaMemcpyHostToDevice, px));
    if (dyn) {              // captured: the graph stages the batch itself (no-op when the launch copied it)
      k_stage_next<<<8, 256, 0, px>>>(reinterpret_cast<uint32_t*>(const_cast<void*>(next_triples)), dyn,
                                      (size_t) B * 3 * (cfg->id_dtype ? 2 : 1));
      KGE_CUDA_OK(cudaGetLastError());
    }
    StepParams kp = p;
    kp.triples = next_triples;
    kp.sd.step = next_step;
    if (dyn) kp.sd.step_dyn = &dyn->step_next;
    pt.begin(3, px);
    if (int rc2 = group_chain(px, kp, wsb + (size_t) (1 - gs) * L.gdelta)) return rc2;
    pt.end(3, px);
    KGE_CUDA_OK(cudaEventRecord(ctx->ev_pre_done, px));
    pre.valid = true; pre.triples = next_triples; pre.B = B; pre.seed = seed; pre.step = next_step; pre.ws = ws;
    pre.i_base = dp ? dp->i_base : 0; pre.k = k; pre.n_ent = cfg->n_ent; pre.side = cfg->corrupt_side;
    pre.sampler = cfg->sampler; pre.set = 1 - gs; pre.graph = ctx->capturing;
    return KGE_OK;
  };
  // Host enqueue order: the caller's stream first (fork point, constraint, step kernel), so the
  // GPU starts on the critical path while the host is still enqueueing the side chain.
  auto fork_side = [&]() -> int {
    if (ctx->opt_pre_early) { if (int rc2 = prefetch_next()) return rc2; }
    if (is_train) KGE_CUDA_OK(cudaEventRecord(ctx->ev_fork, st));
    pt.begin(0);
    if (flush_first) { if (int rc2 = launch_lazy<2>(lp, ctx->sm_count, st)) return rc2; }
    if (shard_phase == 1) return KGE_OK;                  // lookup: ids only, no table is touched
    if (int rc2 = launch_constrain(ctx, cp, st)) return rc2;
    pt.end(0);
    return KGE_OK;
  };
  auto side_chain = [&]() -> int {
    if (!is_train) return KGE_OK;
    if (pre_hit) {          // grouped by the previous call: join its stream (a captured graph is ordered by its launches)
      if (!ctx->capturing) KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_pre_done, 0));
      return KGE_OK;
    }
    cudaStream_t ax = ctx->aux;
    KGE_CUDA_OK(cudaStreamWaitEvent(ax, ctx->ev_fork, 0));
    pt.begin(3, ax);
    int64_t kb = (L.n_keys + 255) / 256;
    if (kb > (int64_t) ctx->sm_count * 8) kb = (int64_t) ctx->sm_count * 8;
    if (lazy) {
      KGE_CUDA_OK(cudaMemsetAsync(seg_lo, 0, sizeof(uint32_t) * 2 * (size_t) cfg->n_rel, ax));
      KGE_CUDA_OK(cudaMemsetAsync(uniq, 0, sizeof(uint32_t) * 4, ax));
      k_keys<<<(unsigned) kb, 256, 0, ax>>>(p, nullptr);
      KGE_CUDA_OK(cudaGetLastError());
      size_t cub_bytes = L.cub_bytes;
      KGE_CUDA_OK(cub::DeviceRadixSort::SortPairs(wsb + L.off_cub, cub_bytes, (const uint32_t*) p.keys, keys_out,
                                                  (const uint32_t*) p.vals, vals_out, (int) L.n_keys, 0, sort_bits(cfg), ax));
      k_heads<<<(unsigned) ((L.n_keys + 255) / 256), 256, 0, ax>>>(keys_out, L.n_keys, (uint32_t) cfg->n_ent, uniq, seg_lo, seg_hi);
      KGE_CUDA_OK(cudaGetLastError());
      pt.end(3, ax);
      KGE_CUDA_OK(cudaEventRecord(ctx->ev_join, ax));
      return KGE_OK;
    }
    if (!ctx->opt_radix_sort) {
      if (int rc2 = group_chain(ax, p, gb)) return rc2;
      pt.end(3, ax);
      KGE_CUDA_OK(cudaEventRecord(ctx->ev_join, ax));
      return KGE_OK;
    }
    KGE_CUDA_OK(cudaMemsetAsync(seg_lo, 0, sizeof(uint32_t) * (2 * L.n_seg + 2), ax));
    if ((ctx->opt_radix_sort != 0)) {
      k_keys<<<(unsigned) kb, 256, 0, ax>>>(p, nullptr);
      KGE_CUDA_OK(cudaGetLastError());
      size_t cub_bytes = L.cub_bytes;
      KGE_CUDA_OK(cub::DeviceRadixSort::SortPairs(wsb + L.off_cub, cub_bytes, (const uint32_t*) p.keys, keys_out,
                                                  (const uint32_t*) p.vals, vals_out, (int) L.n_keys, 0, sort_bits(cfg), ax));
      k_segments<<<(unsigned) ((L.n_keys + 255) / 256), 256, 0, ax>>>(keys_out, L.n_keys, seg_lo, seg_hi);
      KGE_CUDA_OK(cudaGetLastError());
    }
    pt.end(3, ax);
    KGE_CUDA_OK(cudaEventRecord(ctx->ev_join, ax));
    return KGE_OK;
  };

  StepLaunch SL;
  SL.nv = nv;
  SL.train = is_train != 0;
  SL.stream = st;
  int rc;
  if (!mat) {
    int ent_rows = (cfg->model == KGE_TRANSD) ? 2 : 1;
    int rel_rows = (cfg->model == KGE_UM) ? 0 : ((cfg->model == KGE_TRANSH || cfg->model == KGE_TRANSD) ? 2 : 1);
    const int ppc = p.nwarps / p.wpp;
    SL.grid = (B + ppc - 1) / ppc;
    SL.smem = step_smem_bytes(nv, p.wpp, k, ent_rows, rel_rows, p.nwarps);
    // DistMult: per-warp ring of negative rows filled by TMA bulk copies (StepParams.ring_slots) — the deepest
    // of 32 / 16 / 8 / 4 rows that fits the shared memory left, never deeper than a warp's negatives
    p.ring_slots = 0; p.ring_off = 0;
    if (cfg->model == KGE_DISTMULT && nv <= 2 && !sh && (cfg->corrupt_side != KGE_SIDE_HT || (p.wpp & 1) == 0)) {   // the fast loop's condition
      const int n_it_max = (k + p.wpp - 1) / p.wpp;
      const size_t base = (SL.smem + 15) & ~(size_t) 15;
      for (int S = 32; S >= 4; S >>= 1) {
        if (S > ctx->opt_ring_max || (S > 4 && (S >> 1) >= n_it_max)) continue;
        const size_t need = base + (((size_t) p.nwarps * (S >> 2) * 8 + 15) & ~(size_t) 15) + (size_t) p.nwarps * S * L.rld[KGE_T_ENT] * 4;
        if (need > (size_t) ctx->smem_optin) continue;
        p.ring_slots = S; p.ring_off = (uint32_t) base; SL.smem = need;
        break;
      }
    }
    // TransE / UM rows of <= 64 floats, one warp per positive, one corrupted side: two positives per warp
    // (step_narrow_kernels.cuh) — k_step's fast path for these models, bit for bit, in half the warps
    const bool narrow = (cfg->model == KGE_TRANSE || cfg->model == KGE_UM) && nv == 1 && tables->t[KGE_T_ENT].cols <= 64 &&
                        p.wpp == 1 && cfg->corrupt_side != KGE_SIDE_HT && cfg->score != KGE_SCORE_DOT &&
                        cfg->loss != KGE_LOSS_SELF_ADV && !getenv("KGE_STEP_NARROW_OFF");
    // RotatE with LpDistance(1) on wide rows: the streamed kernel (step_rot_kernels.cuh) — same arithmetic, accumulators in
    // shared memory, two CTAs per SM where k_step's 255 registers allow one
    bool rot_stream = false;
    if (cfg->model == KGE_ROTATE && nv >= 4 && !sh && cfg->score == KGE_SCORE_LP && cfg->score_p == 1.0f &&
        (cfg->corrupt_side != KGE_SIDE_HT || (p.wpp & 1) == 0) && !getenv("KGE_STEP_ROT_OLD")) {
      // CTA size: one balanced wave when the batch allows it (as DistMult's step), else 8 warps; fewer if the rings do not fit
      int nwp = STEP_WARPS;
      const int64_t per_sm = (B * p.wpp + ctx->sm_count - 1) / ctx->sm_count;
      if (per_sm <= STEP_MAX_WARPS) {
        nwp = (int) ((per_sm + p.wpp - 1) / p.wpp) * p.wpp;
        if (nwp > STEP_MAX_WARPS) nwp = STEP_MAX_WARPS / p.wpp * p.wpp;
      }
      while (nwp > p.wpp && step_rot_smem_bytes_host(nv, p.wpp, k, nwp, (int) tables->t[KGE_T_ENT].ld) > (size_t) ctx->smem_optin) nwp -= p.wpp;
      const size_t rb = step_rot_smem_bytes_host(nv, p.wpp, k, nwp, (int) tables->t[KGE_T_ENT].ld);
      if (rb <= (size_t) ctx->smem_optin && nwp >= p.wpp) {
        rot_stream = true; SL.smem = rb; p.nwarps = nwp;
        SL.grid = (B + nwp / p.wpp - 1) / (nwp / p.wpp);
      }
    }
    KGE_REQUIRE(SL.smem <= (size_t) ctx->smem_optin, KGE_EUNSUPPORTED, "k_step needs %zu B shared memory", SL.smem);
    if (shard_phase != 2) { if (int rc2 = fork_side()) return rc2; }      // after the last argument check
    else { pt.begin(0); if (int rc2 = launch_constrain(ctx, cp, st)) return rc2; pt.end(0); }
    lp.uniq = uniq; lp.vals = vals_out;
    if (lazy && shard_phase != 2) {
      // the step kernel gathers rows that k_lazy_pre must have replayed: the work list (sort) comes first
      if (int rc2 = side_chain()) return rc2;
      KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join, 0));
      if (shard_phase == 1) {                   // route the ids to their owners; the owners replay (kge_shard_pre)
        KGE_CUDA_OK(cudaMemsetAsync(rcnt, 0, sizeof(uint32_t) * 16, st));
        KGE_CUDA_OK(cudaMemsetAsync(rec_dest, 0xff, sizeof(uint32_t) * (size_t) L.n_ent_rec, st));
        k_route<<<(unsigned) ctx->sm_count * 4, 256, 0, st>>>(uniq, sdv, rcnt, item_slot, vals_out, rec_dest);
        KGE_CUDA_OK(cudaGetLastError());
        k_route_cnt<<<1, 32, 0, st>>>(sdv, rcnt);
        KGE_CUDA_OK(cudaGetLastError());
        return KGE_OK;
      }
      pt.begin(0);
      if (int rc2 = launch_lazy<0>(lp, ctx->sm_count, st)) return rc2;
      pt.end(0);
    }
    pt.begin(1);
    switch (cfg->model) {
      case KGE_UM: rc = narrow ? launch_step_narrow_um(p, SL) : launch_step_um(p, SL); break;
      case KGE_TRANSE: rc = narrow ? launch_step_narrow_transe(p, SL) : launch_step_transe(p, SL); break;
      case KGE_TRANSH: rc = launch_step_transh(p, SL); break;
      case KGE_TRANSD: rc = launch_step_transd(p, SL); break;
      case KGE_ROTATE: rc = rot_stream ? launch_step_rot(p, SL) : launch_step_rotate(p, SL); break;
      case KGE_DISTMULT: rc = launch_step_distmult(p, SL); break;
      default: set_error("model %d has no step kernel", cfg->model); rc = KGE_EUNSUPPORTED;
    }
  } else {
    mq.recP = reinterpret_cast<float*>(wsb + L.off_recP);
    mq.ldp = L.ldp;
    mq.dpad = 128 * nv;
    const int n_mats = cfg->model == KGE_SE ? 2 : 1;
    SL.grid = B;
    SL.smem = step_mat_smem_bytes(nv, k, mq.dpad, SL.train, true, n_mats, mat_rows(cfg), mat_cols(cfg));
    mq.stage = SL.smem <= (size_t) ctx->smem_optin;
    if (!mq.stage) SL.smem = step_mat_smem_bytes(nv, k, mq.dpad, SL.train, false, n_mats, mat_rows(cfg), mat_cols(cfg));
    KGE_REQUIRE(SL.smem <= (size_t) ctx->smem_optin, KGE_EUNSUPPORTED, "k_step_mat needs %zu B shared memory", SL.smem);
    // TransR / SE: four vectors per warp through every mat-vec (k_step_proj) when the positive's rows fit shared memory
    mq.proj = 0;
    if (cfg->model != KGE_RESCAL && k + 2 < PJ_MAX_ROWS && !getenv("KGE_STEP_MAT_OLD")) {
      const size_t pb = sizeof(float) * step_proj_smem_floats(nv, k, mq.dpad, SL.train, n_mats, mat_rows(cfg), mat_cols(cfg));
      if (pb <= (size_t) ctx->smem_optin) { mq.proj = 1; SL.smem = pb; }
    }
    if (int rc2 = fork_side()) return rc2;      // after the last argument check
    pt.begin(1);
    switch (cfg->model) {
      case KGE_SE: rc = launch_step_se(mq, SL); break;
      case KGE_TRANSR: rc = launch_step_transr(mq, SL); break;
      default: rc = launch_step_rescal(mq, SL);
    }
  }
  if (rc) return rc;
  pt.end(1);
  // KGE_OPT_PRE_EARLY = 0: the next batch's grouping starts BEHIND the step kernel.  Measured on C2: the step kernel
  // gains 4 us (32 vs 36) but the chain (43 us beside the update kernels) then outlasts the update: 96 vs 85 us per step
  if (!ctx->opt_pre_early) { if (int rc2 = prefetch_next()) return rc2; }
  if (!lazy) { if (int rc2 = side_chain()) return rc2; }

  // 3. loss + per-table clip scales (sparse-gradient tables)
  const ReduceParams& rp = p.rp;
  if (!p.tail) {
    pt.begin(2);
    k_reduce<<<(unsigned) (B >= 8192 ? RED_CTAS : (B + 1023) / 1024), 1024, 0, st>>>(rp);
    KGE_CUDA_OK(cudaGetLastError());
    pt.end(2);
  }
  if (is_train && shard_phase != 2 && !pre_hit) KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join, 0));     // join the side stream
  if (neg_out)
    KGE_CUDA_OK(cudaMemcpyAsync(neg_out, p.neg_ids, sizeof(int64_t) * (size_t) B * k, cudaMemcpyDeviceToDevice, st));
  if (grad_only) {    // scal[4..8] = norm^2 per slot, loss share
    if (grads_out->n_peers > 0) {
      ScalSlots ss;
      ss.n = grads_out->n_peers;
      for (int j = 0; j < ss.n; ++j) ss.s[j] = grads_out->peer_scal[j];
      k_push_scal<<<1, 32, 0, st>>>(rp.scal + 4, ss);
      KGE_CUDA_OK(cudaGetLastError());
    } else {
      KGE_CUDA_OK(cudaMemcpyAsync(scal_out, rp.scal + 4, sizeof(float) * 5, cudaMemcpyDeviceToDevice, st));
    }
  }
  const bool dense = has_dense_reg(cfg) && !grad_only;    // regularisers are added once, by kge_train_apply
  if (!is_train && !dense) return KGE_OK;

  // 4. (side stream, above) records sorted by (table, row): stable LSD radix sort => occurrence
  //    order inside a row
  // 5. per-row segmented sum + clip + Adam (every row of every table, TF semantics)
  UpdateParams up;
  up.n_tabs = 0;
  up.total_rows = 0;
  ReduceDenseParams rd;
  rd.dyn = nullptr;
  rd.row_begin[0] = 0;
  for (int s = 0; s < KGE_MAX_TABLES; ++s) {
    if (!cols[s] || is_matrix_slot(cfg, s)) continue;
    if (lazy && s < 2) continue;              // k_lazy_update's tables
    UpdateTable& U = up.tab[up.n_tabs];
    U.w = tables->t[s].w; U.m = tables->t[s].m; U.v = tables->t[s].v;
    U.rec = p.rec[s];
    U.rows = tables->t[s].rows; U.cols = tables->t[s].cols; U.ld = tables->t[s].ld;
    U.rec_ld = L.rld[s];
    U.key_base = (s < 2) ? 0u : (uint32_t) cfg->n_ent;
    U.rec_base = (s < 2) ? 0u : (uint32_t) L.n_ent_rec;
    U.slot = s;
    U.fac_coef = nullptr; U.fac_q = nullptr; U.fac_first = 0; U.fac_k = 1; U.fac_ht = 0;
    if (s == KGE_T_ENT && p.fac_coef != nullptr) {
      U.fac_coef = p.fac_coef; U.fac_q = p.fac_q; U.fac_first = (uint32_t) (2 * B); U.fac_k = k;
      U.fac_ht = cfg->corrupt_side == KGE_SIDE_HT;
    }
    U.dense_kind = DK_NONE; U.other_w = nullptr; U.other_ld = 0;
    U.n_peer = 0;
    U.owner_rows = 0; U.row_lo = 0; U.n_out = 0; U.w_mc = nullptr; U.wn_mc = nullptr; U.w_private = 0;
    if (grad_only) {
      if (grads_out->n_peers > 0) {
        U.n_peer = grads_out->n_peers;
        for (int j = 0; j < U.n_peer; ++j) U.peer[j] = grads_out->peer[j][s];
        U.owner_rows = grads_out->owner_rows[s];
      } else { U.n_peer = 1; U.peer[0] = grads_out->dense[s]; }
    }
    U.w_in = (shadow_use && s == KGE_T_ENT) ? tables->t[s].wn : tables->t[s].w;
    U.wn_out = (shadow_ok && s == KGE_T_ENT && !dense) ? tables->t[s].wn : nullptr;
    if (dense) {
      if (cfg->model == KGE_RESCAL) U.dense_kind = DK_RESCAL_ENT;
      else if (s == KGE_T_ENT) U.dense_kind = DK_TRANSH_ENT;
      else {
        U.dense_kind = (s == KGE_T_REL_A) ? DK_TRANSH_REL : DK_TRANSH_HYPER;
        U.other_w = reinterpret_cast<const float*>(wsb + L.off_snap[s == KGE_T_REL_A ? 1 : 0]);
        U.other_ld = L.rld[s == KGE_T_REL_A ? KGE_T_REL_B : KGE_T_REL_A];
      }
    }
    rd.slot[up.n_tabs] = s;
    rd.is_dense[up.n_tabs] = U.dense_kind != DK_NONE;
    up.total_rows += U.rows;
    rd.row_begin[++up.n_tabs] = up.total_rows;
  }
  up.keys = keys_out;
  up.vals = vals_out;
  up.n_keys = is_train ? L.n_keys : 0;
  up.seg_lo = use_lo; up.seg_hi = use_hi;
  up.seg_key0 = lazy ? (uint32_t) cfg->n_ent : 0u;
  up.hdr = reinterpret_cast<const uint32_t*>(gb + L.off_hdr);
  up.heavy_keys = reinterpret_cast<const uint32_t*>(gb + L.off_heavy);
  up.n_heavy = seg_lo + 2 * L.n_seg + 1;
  up.scal = rp.scal;
  up.lr_t = cfg->optimizer == KGE_OPTIM_ADAM ? adam_lr(cfg->lr, cfg->beta1, cfg->beta2, adam_t) : cfg->lr;
  up.lr_dyn = dyn ? &dyn->lr_t : nullptr;
  up.opt = cfg->optimizer;
  up.beta1 = cfg->beta1; up.beta2 = cfg->beta2; up.eps = cfg->eps;
  up.cw = cfg->constraint_weight;
  up.rowpart = dense ? reinterpret_cast<float*>(wsb + L.off_rowpart) : nullptr;
  const int64_t ublocks = (up.total_rows + 7) / 8;
  int max_ch = 1;
  for (int j = 0; j < up.n_tabs; ++j) max_ch = ((up.tab[j].cols + 3) / 4 > max_ch) ? (up.tab[j].cols + 3) / 4 : max_ch;
  const int cpl = max_ch <= 32 ? 1 : (max_ch <= 64 ? 2 : (max_ch <= 128 ? 4 : 8));

  MatUpdParams mu;
  dim3 mgrid(1, 1, 1);
  if (mat) {
    mu.n_tabs = 0;
    auto add = [&](int slot, int sh, int st_, int neg_mode, int er_sh, int er_st, int er_neg) {
      MatUpdTable& T = mu.tab[mu.n_tabs++];
      T.w = tables->t[slot].w; T.m = tables->t[slot].m; T.v = tables->t[slot].v;
      T.rows_m = mat_rows(cfg); T.cols_m = mat_cols(cfg); T.ld = tables->t[slot].ld; T.slot = slot;
      T.use_sh = sh; T.use_st = st_; T.neg_mode = neg_mode;
      T.ent_row_sh = er_sh; T.ent_row_st = er_st; T.ent_row_neg = er_neg;
      T.gout = grad_only ? grads_out->dense[slot] : nullptr;
    };
    if (cfg->model == KGE_TRANSR) add(KGE_T_REL_B, 1, 1, 1, 1, 1, 1);          // G += ent (x) P
    else if (cfg->model == KGE_SE) { add(KGE_T_REL_A, 1, 0, 2, 0, 0, 0); add(KGE_T_REL_B, 0, 1, 3, 0, 0, 0); }   // G += P (x) ent
    else add(KGE_T_REL_A, 1, 1, 0, 1, 0, 0);                                    // h (x) P  +  P (x) t
    mu.n_rel = cfg->n_rel;
    mu.ent_w = tables->t[KGE_T_ENT].w; mu.ent_ld = tables->t[KGE_T_ENT].ld; mu.d_ent = cfg->d_ent;
    mu.recP = mq.recP; mu.ldp = L.ldp; mu.dp = mat_cols(cfg);
    mu.triples = triples; mu.id64 = cfg->id_dtype; mu.neg_ids = p.neg_ids;
    mu.B = B; mu.k = k; mu.side_mode = cfg->corrupt_side;
    mu.keys = keys_out; mu.vals = vals_out; mu.n_keys = is_train ? L.n_keys : 0;
    mu.seg_lo = use_lo; mu.seg_hi = use_hi;
    mu.key_base = (uint32_t) cfg->n_ent; mu.rec_base = (uint32_t) L.n_ent_rec;
    mu.opt = cfg->optimizer;
    mu.scal = rp.scal; mu.lr_t = up.lr_t; mu.lr_dyn = up.lr_dyn; mu.beta1 = cfg->beta1; mu.beta2 = cfg->beta2; mu.eps = cfg->eps;
    mu.reg_coef = (dense && cfg->model == KGE_RESCAL) ? 2.0f * cfg->constraint_weight / (float) cfg->n_rel : 0.f;
    mu.part = dense ? reinterpret_cast<float*>(wsb + L.off_matpart) : nullptr;
    mgrid = dim3((unsigned) mat_tiles(cfg), (unsigned) cfg->n_rel, (unsigned) mu.n_tabs);
    KGE_REQUIRE(cfg->n_rel <= 65535, KGE_EUNSUPPORTED, "matrix models: n_rel must be <= 65535");
  }

  pt.begin(4);
  if (grad_only) {
    if (sh) {           // relation tables -> grads.dense (general kernel on k_heads' segment table); entity rows -> owners
      if (up.n_tabs > 0) {            // beside the push, on the side stream
        KGE_CUDA_OK(cudaEventRecord(ctx->ev_mid, st));
        KGE_CUDA_OK(cudaStreamWaitEvent(ctx->aux, ctx->ev_mid, 0));
        if (int rc2 = launch_update<3>(up, cpl, ublocks, ctx->aux)) return rc2;
        KGE_CUDA_OK(cudaEventRecord(ctx->ev_heavy, ctx->aux));
      }
      for (int j = 0; j < lp.n_tabs; ++j) { lp.tab[j].rec = p.rec[lp.tab[j].slot]; lp.tab[j].rec_ld = L.rld[lp.tab[j].slot]; }
      if (int rc2 = launch_shard_rows<0>(lp, sdv, item_slot, ctx->sm_count, st, p.rec_dest != nullptr)) return rc2;
      if (up.n_tabs > 0) KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_heavy, 0));
    } else if (ctx->opt_radix_sort) {
      if (int rc2 = launch_update<3>(up, cpl, ublocks, st)) return rc2;
    } else {
      KGE_CUDA_OK(cudaEventRecord(ctx->ev_mid, st));
      KGE_CUDA_OK(cudaStreamWaitEvent(ctx->aux, ctx->ev_mid, 0));
      if (int rc2 = launch_update_heavy<3>(up, cpl, L.n_keys / UPD_HEAVY + 1, ctx->aux)) return rc2;
      KGE_CUDA_OK(cudaEventRecord(ctx->ev_heavy, ctx->aux));
      if (int rc2 = launch_update_fast<3>(up, cpl, st)) return rc2;
      KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_heavy, 0));
    }
    if (mat) {          // this rank's share of every relation matrix's gradient, dense (tables untouched)
      if (mu_tile_cols(mat_cols(cfg)) == MU_TILE) k_update_mat<2, MU_TILE><<<mgrid, MU_THREADS, 0, st>>>(mu);
      else k_update_mat<2, 2 * MU_TILE><<<mgrid, MU_THREADS, 0, st>>>(mu);
      KGE_CUDA_OK(cudaGetLastError());
    }
    pt.end(4);
    if (!sh) pt.keep_open();           // kge_train_apply continues this call's profile slot
    if (ctx->capturing && pre.valid) KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_pre_done, 0));   // a graph ends joined
    return KGE_OK;
  }
  if (dense) {
    // densified gradients: norm pass, reduce, then apply  (tf.clip_by_norm on a dense tensor)
    if (cfg->model == KGE_TRANSH)
      for (int j = 0; j < 2; ++j) {
        const kge_table& S = tables->t[j == 0 ? KGE_T_REL_A : KGE_T_REL_B];
        KGE_CUDA_OK(cudaMemcpy2DAsync(wsb + L.off_snap[j], sizeof(float) * L.rld[j == 0 ? KGE_T_REL_A : KGE_T_REL_B], S.w,
                                      sizeof(float) * S.ld, sizeof(float) * L.rld[j == 0 ? KGE_T_REL_A : KGE_T_REL_B],
                                      (size_t) S.rows, cudaMemcpyDeviceToDevice, st));
      }
    if (int rc2 = launch_update<1>(up, cpl, ublocks, st)) return rc2;
    rd.rowpart = up.rowpart; rd.total_rows = up.total_rows; rd.n_tabs = up.n_tabs;
    rd.matpart = nullptr; rd.n_mat = 0; rd.mat_slot = KGE_T_REL_A; rd.mat_loss_coef = 0.f;
    if (mat) {
      if (mu_tile_cols(mat_cols(cfg)) == MU_TILE) k_update_mat<1, MU_TILE><<<mgrid, MU_THREADS, 0, st>>>(mu);
      else k_update_mat<1, 2 * MU_TILE><<<mgrid, MU_THREADS, 0, st>>>(mu);
      KGE_CUDA_OK(cudaGetLastError());
      rd.matpart = mu.part; rd.n_mat = L.n_mat_cta; rd.mat_loss_coef = cfg->constraint_weight / (float) cfg->n_rel;
    }
    rd.clip_norm = cfg->clip_norm; rd.write_scales = is_train != 0; rd.scal = rp.scal; rd.loss_out = loss_out;
    rd.dyn = (dyn && !grad_only) ? dyn : nullptr;
    k_reduce_dense<<<1, 1024, 0, st>>>(rd);
    KGE_CUDA_OK(cudaGetLastError());
    if (!is_train) { pt.end(4); return KGE_OK; }
  }
  if (mat) {      // reads the pre-update entity rows: must precede k_update
    if (mu_tile_cols(mat_cols(cfg)) == MU_TILE) k_update_mat<0, MU_TILE><<<mgrid, MU_THREADS, 0, st>>>(mu);
    else k_update_mat<0, 2 * MU_TILE><<<mgrid, MU_THREADS, 0, st>>>(mu);
    KGE_CUDA_OK(cudaGetLastError());
  }
  if (lazy) {
    // looked-up entity rows on the caller's stream; the small dense tables (relations) beside them
    KGE_CUDA_OK(cudaEventRecord(ctx->ev_mid, st));
    if (up.n_tabs > 0) {
      KGE_CUDA_OK(cudaStreamWaitEvent(ctx->aux, ctx->ev_mid, 0));
      if (int rc2 = launch_update<0>(up, cpl, ublocks, ctx->aux)) return rc2;
      KGE_CUDA_OK(cudaEventRecord(ctx->ev_heavy, ctx->aux));
    }
    lp.scal = rp.scal;
    for (int j = 0; j < lp.n_tabs; ++j) { lp.tab[j].rec = p.rec[lp.tab[j].slot]; lp.tab[j].rec_ld = L.rld[lp.tab[j].slot]; }
    if (int rc2 = launch_lazy<1>(lp, ctx->sm_count, st)) return rc2;
    if (up.n_tabs > 0) KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_heavy, 0));
    tables->wn_valid[KGE_T_ENT] = 0;
    tables->lazy_t = (int64_t) adam_t;
  } else if (dense) {
    if (int rc2 = launch_update<2>(up, cpl, ublocks, st)) return rc2;
    tables->wn_valid[KGE_T_ENT] = 0;
  } else if (ctx->opt_radix_sort) {
    if (int rc2 = launch_update<0>(up, cpl, ublocks, st)) return rc2;
    tables->wn_valid[KGE_T_ENT] = 0;
  } else {
    // light rows: independent warps on the caller's stream; the few rows with many records are
    // summed by whole CTAs on the side stream at the same time (disjoint rows)
    KGE_CUDA_OK(cudaEventRecord(ctx->ev_mid, st));
    KGE_CUDA_OK(cudaStreamWaitEvent(ctx->aux, ctx->ev_mid, 0));
    if (int rc2 = launch_update_heavy<0>(up, cpl, L.n_keys / UPD_HEAVY + 1, ctx->aux)) return rc2;
    KGE_CUDA_OK(cudaEventRecord(ctx->ev_heavy, ctx->aux));
    if (int rc2 = launch_update_fast<0>(up, cpl, st)) return rc2;
    KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_heavy, 0));
    tables->wn_valid[KGE_T_ENT] = shadow_ok ? 1 : 0;        // wn now holds normalise(new w)
  }
  pt.end(4);
  if (ctx->capturing && pre.valid) KGE_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_pre_done, 0));     // a graph ends joined
  return KGE_OK;
}

extern "C" int kge_train_step(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                              const void* triples, int64_t B, uint64_t seed, uint64_t step, uint64_t adam_t,
                              const int64_t* neg_override, int is_train, float* loss_out, int64_t* neg_out,
                              void* ws, size_t ws_bytes, void* stream) {
  return train_impl(ctx, cfg, tables, smp, nullptr, triples, B, seed, step, adam_t, neg_override, is_train, nullptr,
                    nullptr, loss_out, neg_out, ws, ws_bytes, stream);
}

extern "C" int kge_train_step_host(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                                   const void* host_triples, void* dev_triples, int64_t B, uint64_t seed, uint64_t step,
                                   uint64_t adam_t, int is_train, float* dev_loss, float* host_loss_out, void* ws,
                                   size_t ws_bytes, void* stream) {
  KGE_REQUIRE(cfg && host_triples && dev_triples && dev_loss && host_loss_out && B > 0, KGE_EINVAL,
              "cfg / host_triples / dev_triples / dev_loss / host_loss_out / B invalid");
  cudaStream_t st = (cudaStream_t) stream;
  KGE_CUDA_OK(cudaMemcpyAsync(dev_triples, host_triples, (size_t) B * 3 * (cfg->id_dtype ? 8 : 4), cudaMemcpyHostToDevice, st));
  if (int rc = train_impl(ctx, cfg, tables, smp, nullptr, dev_triples, B, seed, step, adam_t, nullptr, is_train, nullptr,
                          nullptr, dev_loss, nullptr, ws, ws_bytes, stream)) return rc;
  KGE_CUDA_OK(cudaMemcpyAsync(host_loss_out, dev_loss, sizeof(float), cudaMemcpyDeviceToHost, st));
  return KGE_OK;
}

extern "C" int kge_train_step_next(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                                   const void* triples, int64_t B, uint64_t seed, uint64_t step, uint64_t adam_t,
                                   float* loss_out, const void* next_triples, uint64_t next_step, void* ws, size_t ws_bytes,
                                   void* stream) {
  return train_impl(ctx, cfg, tables, smp, nullptr, triples, B, seed, step, adam_t, nullptr, 1, nullptr, nullptr, loss_out,
                    nullptr, ws, ws_bytes, stream, nullptr, 0, next_triples, next_step);
}

extern "C" int kge_train_step_host_next(kge_ctx* ctx_, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                                        const void* host_triples, void* dev_triples, int64_t B, uint64_t seed, uint64_t step,
                                        uint64_t adam_t, float* dev_loss, float* host_loss_out, const void* host_next,
                                        void* dev_next, uint64_t next_step, void* ws, size_t ws_bytes, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx && cfg && host_triples && dev_triples && dev_loss && host_loss_out && B > 0, KGE_EINVAL,
              "ctx / cfg / host_triples / dev_triples / dev_loss / host_loss_out / B invalid");
  KGE_REQUIRE(!host_next || (dev_next && dev_next != dev_triples), KGE_EINVAL, "host_next needs its own staging buffer dev_next");
  cudaStream_t st = (cudaStream_t) stream;
  // the previous call copied (and grouped) this batch as its `next`: nothing to copy
  const Prefetched& pre = ctx->pre;
  const bool staged = pre.valid && !pre.graph && pre.triples == dev_triples && pre.B == B && pre.seed == seed && pre.step == step &&
                      pre.ws == ws;
  if (!staged)
    KGE_CUDA_OK(cudaMemcpyAsync(dev_triples, host_triples, (size_t) B * 3 * (cfg->id_dtype ? 8 : 4), cudaMemcpyHostToDevice, st));
  if (int rc = train_impl(ctx_, cfg, tables, smp, nullptr, dev_triples, B, seed, step, adam_t, nullptr, 1, nullptr, nullptr,
                          dev_loss, nullptr, ws, ws_bytes, stream, nullptr, 0, host_next ? dev_next : nullptr, next_step, host_next))
    return rc;
  KGE_CUDA_OK(cudaMemcpyAsync(host_loss_out, dev_loss, sizeof(float), cudaMemcpyDeviceToHost, st));
  return KGE_OK;
}

extern "C" int kge_train_grad(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                              const kge_dp* dp, const void* triples, int64_t B, uint64_t seed, uint64_t step,
                              const int64_t* neg_override, kge_grads* out, float* scal_out, int64_t* neg_out, void* ws,
                              size_t ws_bytes, void* stream) {
  KGE_REQUIRE(dp, KGE_EINVAL, "dp is NULL");
  return train_impl(ctx, cfg, tables, smp, dp, triples, B, seed, step, 1, neg_override, 1, out, scal_out, nullptr,
                    neg_out, ws, ws_bytes, stream);
}

// Second half of a data-parallel step: the rank-summed gradients replace the record sums.
extern "C" int kge_train_apply(kge_ctx* ctx_, const kge_cfg* cfg, kge_tables* tables, const kge_grads* g,
                               const float* scal_summed, int64_t B, uint64_t adam_t, float* loss_out, void* ws,
                               size_t ws_bytes, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  cudaStream_t st = (cudaStream_t) stream;
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  if (int rc = check_tables(cfg, tables, true)) return rc;
  KGE_REQUIRE(g && loss_out && ws && B > 0 && adam_t >= 1, KGE_EINVAL, "bad kge_train_apply arguments");
  KGE_REQUIRE(!lazy_requested(tables), KGE_EUNSUPPORTED, "row-lazy tables: the data-parallel halves are not implemented");
  KGE_REQUIRE(g->n_peers >= 0 && g->n_peers <= KGE_MAX_PEERS, KGE_EINVAL, "grads.n_peers %d out of range", g->n_peers);
  const bool mat = is_matrix_model(cfg->model);
  KGE_REQUIRE(!mat || (g->n_peers == 0 && !has_dense_reg(cfg)), KGE_EUNSUPPORTED,
              "data-parallel SE / TransR / RESCAL: gradients through grads.dense only, no dense regulariser (else single-GPU)");
  StepWs L;
  if (int rc = layout_step_ws(cfg, B, L)) return rc;
  KGE_REQUIRE(ws_bytes >= L.total, KGE_EWORKSPACE, "workspace %zu < required %zu bytes", ws_bytes, L.total);
  KGE_REQUIRE(((uintptr_t) ws & 255) == 0, KGE_EINVAL, "workspace must be 256-byte aligned");
  char* wsb = reinterpret_cast<char*>(ws);
  int cols[KGE_MAX_TABLES];
  expected_cols(cfg, cols);
  float* scal = reinterpret_cast<float*>(wsb + L.off_scal);
  PhaseTimer pt(ctx, st, true);
  pt.begin(8);
  ScalSlots ss;
  if (g->n_peers > 0) {
    ss.n = g->n_peers;
    for (int j = 0; j < ss.n; ++j) { KGE_REQUIRE(g->peer_scal[j], KGE_EINVAL, "grads.peer_scal[%d] is NULL", j); ss.s[j] = g->peer_scal[j]; }
  } else { KGE_REQUIRE(scal_summed, KGE_EINVAL, "scal_summed is NULL"); ss.n = 1; ss.s[0] = scal_summed; }
  const bool dense = has_dense_reg(cfg);
  // captured step graph: the kernel that makes the loss final hands it to the host (DynScal.host_loss)
  const DynScal* dyn_pub = ctx->capturing ? ctx->dyn : nullptr;
  k_dp_scales<<<1, 32, 0, st>>>(ss, cfg->clip_norm, scal, loss_out, g->abort_word, dense ? nullptr : dyn_pub);
  KGE_CUDA_OK(cudaGetLastError());

  const bool shadow_ok = shadow_capable(ctx, cfg, tables);
  UpdateParams up;
  up.n_tabs = 0;
  up.total_rows = 0;
  ReduceDenseParams rd;
  rd.dyn = dense ? dyn_pub : nullptr;
  rd.row_begin[0] = 0;
  int max_ch = 1;
  for (int s = 0; s < KGE_MAX_TABLES; ++s) {
    if (!cols[s] || is_matrix_slot(cfg, s)) continue;      // matrix tables: k_mat_apply below
    UpdateTable& U = up.tab[up.n_tabs];
    U.w = tables->t[s].w; U.m = tables->t[s].m; U.v = tables->t[s].v;
    U.rec = nullptr;
    U.rows = tables->t[s].rows; U.cols = tables->t[s].cols; U.ld = tables->t[s].ld;
    U.rec_ld = 0; U.key_base = 0; U.rec_base = 0;
    U.slot = s;
    U.fac_coef = nullptr; U.fac_q = nullptr; U.fac_first = 0; U.fac_k = 1; U.fac_ht = 0;
    U.dense_kind = DK_NONE; U.other_w = nullptr; U.other_ld = 0;
    U.owner_rows = 0; U.row_lo = 0; U.n_out = 0; U.w_mc = nullptr; U.wn_mc = nullptr; U.w_private = 0;
    if (g->n_peers > 0) {
      U.n_peer = g->n_peers;
      for (int j = 0; j < U.n_peer; ++j) {
        KGE_REQUIRE(g->peer[j][s] && ((uintptr_t) g->peer[j][s] & 15) == 0, KGE_EINVAL,
                    "grads.peer[%d][%d] missing or not 16-byte aligned", j, s);
        U.peer[j] = g->peer[j][s];
      }
    } else {
      KGE_REQUIRE(g->dense[s] && ((uintptr_t) g->dense[s] & 15) == 0, KGE_EINVAL, "grads.dense[%d] missing or not 16-byte aligned", s);
      U.n_peer = 1; U.peer[0] = g->dense[s];
    }
    U.w_in = (shadow_ok && tables->wn_valid[KGE_T_ENT] && s == KGE_T_ENT) ? tables->t[s].wn : tables->t[s].w;
    U.wn_out = (shadow_ok && s == KGE_T_ENT && !dense) ? tables->t[s].wn : nullptr;
    if (g->n_peers > 0 && g->owner_rows[s] != 0) {        // owner-partitioned: my rows only, new rows to every rank
      KGE_REQUIRE(!dense, KGE_EUNSUPPORTED, "owner-partitioned exchange: no dense regulariser");
      KGE_REQUIRE(g->owner_rows[s] > 0 && g->owner_rank >= 0 && g->owner_rank < g->n_peers &&
                      (int64_t) g->owner_rows[s] * g->n_peers >= U.rows, KGE_EINVAL, "bad grads.owner_rows / owner_rank");
      U.owner_rows = g->owner_rows[s];
      U.row_lo = (int64_t) g->owner_rank * g->owner_rows[s];
      int64_t hi = U.row_lo + g->owner_rows[s];
      if (hi > U.rows) hi = U.rows;
      U.rows = hi > U.row_lo ? hi - U.row_lo : 0;
      U.n_out = g->n_peers;
      U.w_private = g->w_private;
      U.w_mc = g->w_mc[s];
      U.wn_mc = (U.wn_out && g->w_mc[s]) ? g->wn_mc : nullptr;
      KGE_REQUIRE(!(U.wn_out && U.w_mc) || U.wn_mc, KGE_EINVAL, "grads.w_mc is given but grads.wn_mc is missing");
      KGE_REQUIRE((((uintptr_t) U.w_mc | (uintptr_t) U.wn_mc) & 15) == 0, KGE_EINVAL, "grads.w_mc / wn_mc not 16-byte aligned");
      for (int j = 0; j < U.n_out; ++j) {
        KGE_REQUIRE(g->w_peer[j][s] && ((uintptr_t) g->w_peer[j][s] & 15) == 0, KGE_EINVAL, "grads.w_peer[%d][%d] missing", j, s);
        U.w_outs[j] = g->w_peer[j][s];
        U.wn_outs[j] = nullptr;
        if (U.wn_out) {
          KGE_REQUIRE(g->wn_peer[j] && ((uintptr_t) g->wn_peer[j] & 15) == 0, KGE_EINVAL, "grads.wn_peer[%d] missing", j);
          U.wn_outs[j] = g->wn_peer[j];
        }
      }
      KGE_REQUIRE(g->w_peer[g->owner_rank][s] == tables->t[s].w, KGE_EINVAL, "grads.w_peer[owner_rank][%d] is not this rank's table", s);
    }
    if (dense) {            // TransH only (RESCAL is a matrix model)
      if (s == KGE_T_ENT) U.dense_kind = DK_TRANSH_ENT;
      else {
        U.dense_kind = (s == KGE_T_REL_A) ? DK_TRANSH_REL : DK_TRANSH_HYPER;
        U.other_w = reinterpret_cast<const float*>(wsb + L.off_snap[s == KGE_T_REL_A ? 1 : 0]);
        U.other_ld = L.rld[s == KGE_T_REL_A ? KGE_T_REL_B : KGE_T_REL_A];
      }
    }
    rd.slot[up.n_tabs] = s;
    rd.is_dense[up.n_tabs] = U.dense_kind != DK_NONE;
    up.total_rows += U.rows;
    rd.row_begin[++up.n_tabs] = up.total_rows;
    max_ch = ((U.cols + 3) / 4 > max_ch) ? (U.cols + 3) / 4 : max_ch;
  }
  up.keys = nullptr; up.vals = nullptr; up.n_keys = 0; up.seg_lo = nullptr; up.seg_hi = nullptr; up.seg_key0 = 0;
  up.hdr = nullptr; up.heavy_keys = nullptr; up.n_heavy = nullptr;
  up.scal = scal;
  up.lr_t = cfg->optimizer == KGE_OPTIM_ADAM ? adam_lr(cfg->lr, cfg->beta1, cfg->beta2, adam_t) : cfg->lr;
  up.lr_dyn = ctx->capturing ? &ctx->dyn->lr_t : nullptr;
  up.opt = cfg->optimizer;
  up.beta1 = cfg->beta1; up.beta2 = cfg->beta2; up.eps = cfg->eps;
  up.cw = cfg->constraint_weight;
  up.rowpart = dense ? reinterpret_cast<float*>(wsb + L.off_rowpart) : nullptr;
  const int64_t ublocks = (up.total_rows + 7) / 8;
  const int cpl = max_ch <= 32 ? 1 : (max_ch <= 64 ? 2 : (max_ch <= 128 ? 4 : 8));
  if (dense) {
    for (int j = 0; j < 2; ++j) {
      const int sl = j == 0 ? KGE_T_REL_A : KGE_T_REL_B;
      const kge_table& S = tables->t[sl];
      KGE_CUDA_OK(cudaMemcpy2DAsync(wsb + L.off_snap[j], sizeof(float) * L.rld[sl], S.w, sizeof(float) * S.ld,
                                    sizeof(float) * L.rld[sl], (size_t) S.rows, cudaMemcpyDeviceToDevice, st));
    }
    if (int rc = launch_update<1>(up, cpl, ublocks, st)) return rc;
    rd.rowpart = up.rowpart; rd.total_rows = up.total_rows; rd.n_tabs = up.n_tabs;
    rd.matpart = nullptr; rd.n_mat = 0; rd.mat_slot = KGE_T_REL_A; rd.mat_loss_coef = 0.f;
    rd.clip_norm = cfg->clip_norm; rd.write_scales = 1; rd.scal = scal; rd.loss_out = loss_out;
    k_reduce_dense<<<1, 1024, 0, st>>>(rd);
    KGE_CUDA_OK(cudaGetLastError());
    if (int rc = launch_update<2>(up, cpl, ublocks, st)) return rc;
  } else {
    if (int rc = launch_update_fast<4>(up, cpl, st)) return rc;      // streaming pass: gradient = gsum row
  }
  if (mat) {
    for (int s = 0; s < KGE_MAX_TABLES; ++s) {
      if (!cols[s] || !is_matrix_slot(cfg, s)) continue;
      KGE_REQUIRE(g->dense[s] && ((uintptr_t) g->dense[s] & 15) == 0, KGE_EINVAL, "grads.dense[%d] missing or not 16-byte aligned", s);
      MatApplyParams ma;
      ma.w = tables->t[s].w; ma.m = tables->t[s].m; ma.v = tables->t[s].v; ma.gsum = g->dense[s];
      ma.n_rel = cfg->n_rel; ma.elems = cols[s]; ma.ld = tables->t[s].ld; ma.slot = s;
      ma.scal = scal; ma.lr_t = up.lr_t; ma.lr_dyn = up.lr_dyn; ma.beta1 = cfg->beta1; ma.beta2 = cfg->beta2; ma.eps = cfg->eps;
      ma.opt = cfg->optimizer;
      const int64_t total = cfg->n_rel * (int64_t) cols[s];
      int64_t blocks = (total + 255) / 256;
      if (blocks > (int64_t) ctx->sm_count * 16) blocks = (int64_t) ctx->sm_count * 16;
      k_mat_apply<<<(unsigned) blocks, 256, 0, st>>>(ma);
      KGE_CUDA_OK(cudaGetLastError());
    }
  }
  tables->wn_valid[KGE_T_ENT] = (shadow_ok && !dense) ? 1 : 0;
  pt.end(8);
  return KGE_OK;
}

// Materialise the row-lazy tables: every row replayed to tables->lazy_t (kge_abi.h).
extern "C" int kge_flush_rows(kge_ctx* ctx_, const kge_cfg* cfg, kge_tables* tables, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  KGE_REQUIRE(tables, KGE_EINVAL, "tables is NULL");
  if (!lazy_requested(tables)) return KGE_OK;
  if (int rc = check_tables(cfg, tables, true, tables->t[KGE_T_ENT].rows)) return rc;   // a shard holds its own rows only
  if (int rc = check_lazy(cfg, tables)) return rc;
  LazyParams lp;
  fill_lazy(cfg, tables, (uint32_t) tables->lazy_t, lp);
  return launch_lazy<2>(lp, ctx->sm_count, (cudaStream_t) stream);
}

// ---- sharded entity tables: the four calls of one step (kge_abi.h, shard_kernels.cuh) ---------
extern "C" int kge_shard_lookup(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                                const kge_shard* sh, const kge_dp* dp, const void* triples, int64_t B, uint64_t seed,
                                uint64_t step, const int64_t* neg_override, void* ws, size_t ws_bytes, void* stream) {
  KGE_REQUIRE(sh && dp, KGE_EINVAL, "shard / dp is NULL");
  return train_impl(ctx, cfg, tables, smp, dp, triples, B, seed, step, 1, neg_override, 1, nullptr, nullptr, nullptr,
                    nullptr, ws, ws_bytes, stream, sh, 1);
}

extern "C" int kge_shard_pre(kge_ctx* ctx_, const kge_cfg* cfg, kge_tables* tables, const kge_shard* sh, uint64_t adam_t,
                             void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  cudaStream_t st = (cudaStream_t) stream;
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  KGE_REQUIRE(tables && lazy_requested(tables), KGE_EINVAL, "sharded step: entity slots need `last`");
  if (int rc = check_shard(cfg, tables, sh, 0)) return rc;
  if (int rc = check_tables(cfg, tables, true, shard_local_rows(cfg, sh))) return rc;
  if (int rc = check_lazy(cfg, tables)) return rc;
  KGE_REQUIRE((int64_t) adam_t == tables->lazy_t + 1, KGE_EINVAL, "sharded step: adam_t %llu must be lazy_t + 1 = %lld",
              (unsigned long long) adam_t, (long long) tables->lazy_t + 1);
  ShardDev sdv;
  fill_shard_dev(sh, sdv);
  KGE_CUDA_OK(cudaMemsetAsync(sh->work, 0, sizeof(uint32_t) * 4, st));
  int64_t cb = (sh->cap + 255) / 256;
  if (cb > (int64_t) ctx->sm_count * 2) cb = (int64_t) ctx->sm_count * 2;
  const dim3 grid((unsigned) cb, (unsigned) sh->world, 1);
  k_claim<<<grid, 256, 0, st>>>(sdv, (uint32_t) adam_t);
  KGE_CUDA_OK(cudaGetLastError());
  k_contrib<<<grid, 256, 0, st>>>(sdv);
  KGE_CUDA_OK(cudaGetLastError());
  LazyParams lp;
  fill_lazy(cfg, tables, (uint32_t) adam_t, lp);
  lp.uniq = sh->work;
  return launch_lazy<0>(lp, ctx->sm_count, st);
}

extern "C" int kge_shard_grad(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                              const kge_shard* sh, const kge_dp* dp, const void* triples, int64_t B, uint64_t seed,
                              uint64_t step, const int64_t* neg_override, kge_grads* out, float* scal_out,
                              int64_t* neg_out, void* ws, size_t ws_bytes, void* stream) {
  KGE_REQUIRE(sh && dp && out && scal_out, KGE_EINVAL, "shard / dp / grads / scal_out is NULL");
  return train_impl(ctx, cfg, tables, smp, dp, triples, B, seed, step, 1, neg_override, 1, out, scal_out, nullptr,
                    neg_out, ws, ws_bytes, stream, sh, 2);
}

extern "C" int kge_shard_apply(kge_ctx* ctx_, const kge_cfg* cfg, kge_tables* tables, const kge_shard* sh,
                               const kge_grads* g, const float* scal_summed, int64_t B, uint64_t adam_t,
                               float* loss_out, void* ws, size_t ws_bytes, void* stream) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  cudaStream_t st = (cudaStream_t) stream;
  KGE_REQUIRE(ctx, KGE_EINVAL, "ctx is NULL");
  if (int rc = check_cfg(cfg)) return rc;
  KGE_REQUIRE(tables && lazy_requested(tables), KGE_EINVAL, "sharded step: entity slots need `last`");
  if (int rc = check_shard(cfg, tables, sh, B)) return rc;
  if (int rc = check_tables(cfg, tables, true, shard_local_rows(cfg, sh))) return rc;
  if (int rc = check_lazy(cfg, tables)) return rc;
  KGE_REQUIRE(g && scal_summed && loss_out && ws && B > 0, KGE_EINVAL, "bad kge_shard_apply arguments");
  KGE_REQUIRE((int64_t) adam_t == tables->lazy_t + 1, KGE_EINVAL, "sharded step: adam_t %llu must be lazy_t + 1 = %lld",
              (unsigned long long) adam_t, (long long) tables->lazy_t + 1);
  StepWs L;
  if (int rc = layout_step_ws(cfg, B, L)) return rc;
  KGE_REQUIRE(ws_bytes >= L.total, KGE_EWORKSPACE, "workspace %zu < required %zu bytes", ws_bytes, L.total);
  KGE_REQUIRE(((uintptr_t) ws & 255) == 0, KGE_EINVAL, "workspace must be 256-byte aligned");
  char* wsb = reinterpret_cast<char*>(ws);
  int cols[KGE_MAX_TABLES];
  expected_cols(cfg, cols);
  float* scal = reinterpret_cast<float*>(wsb + L.off_scal);
  PhaseTimer pt(ctx, st);
  pt.begin(4);
  ScalSlots ss;
  ss.n = 1; ss.s[0] = scal_summed;
  k_dp_scales<<<1, 32, 0, st>>>(ss, cfg->clip_norm, scal, loss_out, nullptr);
  KGE_CUDA_OK(cudaGetLastError());
  // replicated relation tables: the data-parallel streaming apply
  UpdateParams up;
  up.n_tabs = 0; up.total_rows = 0;
  int max_ch = 1;
  for (int s = 2; s < KGE_MAX_TABLES; ++s) {
    if (!cols[s]) continue;
    KGE_REQUIRE(g->dense[s] && ((uintptr_t) g->dense[s] & 15) == 0, KGE_EINVAL, "grads.dense[%d] missing or not 16-byte aligned", s);
    UpdateTable& U = up.tab[up.n_tabs++];
    U.w = tables->t[s].w; U.m = tables->t[s].m; U.v = tables->t[s].v;
    U.rec = nullptr; U.rows = tables->t[s].rows; U.cols = tables->t[s].cols; U.ld = tables->t[s].ld;
    U.rec_ld = 0; U.key_base = 0; U.rec_base = 0; U.slot = s;
    U.fac_coef = nullptr; U.fac_q = nullptr; U.fac_first = 0; U.fac_k = 1; U.fac_ht = 0;
    U.dense_kind = DK_NONE; U.other_w = nullptr; U.other_ld = 0;
    U.n_peer = 1; U.peer[0] = g->dense[s];
    U.owner_rows = 0; U.row_lo = 0; U.n_out = 0; U.w_mc = nullptr; U.wn_mc = nullptr; U.w_private = 0;
    U.w_in = tables->t[s].w; U.wn_out = nullptr;
    up.total_rows += U.rows;
    max_ch = ((U.cols + 3) / 4 > max_ch) ? (U.cols + 3) / 4 : max_ch;
  }
  up.keys = nullptr; up.vals = nullptr; up.n_keys = 0; up.seg_lo = nullptr; up.seg_hi = nullptr; up.seg_key0 = 0;
  up.hdr = nullptr; up.heavy_keys = nullptr; up.n_heavy = nullptr;
  up.scal = scal;
  up.lr_t = cfg->optimizer == KGE_OPTIM_ADAM ? adam_lr(cfg->lr, cfg->beta1, cfg->beta2, adam_t) : cfg->lr;
  up.lr_dyn = nullptr;
  up.opt = cfg->optimizer;
  up.beta1 = cfg->beta1; up.beta2 = cfg->beta2; up.eps = cfg->eps;
  up.cw = cfg->constraint_weight; up.rowpart = nullptr;
  const int cpl = max_ch <= 32 ? 1 : (max_ch <= 64 ? 2 : (max_ch <= 128 ? 4 : 8));
  if (up.n_tabs > 0) { if (int rc = launch_update_fast<4>(up, cpl, st)) return rc; }
  // this rank's entity rows
  ShardDev sdv;
  fill_shard_dev(sh, sdv);
  LazyParams lp;
  fill_lazy(cfg, tables, (uint32_t) adam_t, lp);
  lp.uniq = sh->work; lp.scal = scal;
  if (int rc = launch_shard_rows<1>(lp, sdv, nullptr, ctx->sm_count, st)) return rc;
  tables->wn_valid[KGE_T_ENT] = 0;
  tables->lazy_t = (int64_t) adam_t;
  pt.end(4);
  return KGE_OK;
}

// ---- cross-GPU barrier over peer memory + the one-call data-parallel step -----------------------
namespace kge {
struct PeerFlagsDev { int world, rank; uint32_t* flags[KGE_MAX_PEERS]; };
__global__ void k_peer_barrier(PeerFlagsDev f, uint32_t epoch, const uint32_t* epoch_dyn) {
  const int j = threadIdx.x;
  if (j >= f.world) return;
  if (epoch_dyn) epoch += *epoch_dyn;                            // captured step graph: base epoch of this launch (DynScal)
  __threadfence_system();                                        // this stream's earlier stores first
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.flags[j] + f.rank), "r"(epoch) : "memory");
  const uint32_t* mine = f.flags[f.rank] + j;
  const long long t0 = clock64();
  uint32_t v;
  do {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    if (v >= epoch) break;
    if (clock64() - t0 > 40000000000ll) {                        // ~20 s: a peer is gone; do not hang the GPU
      f.flags[f.rank][KGE_MAX_PEERS] = KGE_PEER_DEAD;            // sticky error word: the apply half skips, the host raises
      break;
    }
  } while (true);
  __threadfence_system();
}
}  // namespace kge

extern "C" int kge_peer_barrier(kge_ctx* ctx_, const kge_peer_flags* pf, uint32_t epoch, void* stream) {
  KGE_REQUIRE(ctx_ && pf, KGE_EINVAL, "ctx / flags is NULL");
  KGE_REQUIRE(pf->world >= 1 && pf->world <= KGE_MAX_PEERS && pf->rank >= 0 && pf->rank < pf->world, KGE_EINVAL,
              "bad kge_peer_flags (world %d rank %d)", pf->world, pf->rank);
  PeerFlagsDev f;
  f.world = pf->world; f.rank = pf->rank;
  for (int j = 0; j < KGE_MAX_PEERS; ++j) {
    f.flags[j] = pf->flags[j];
    KGE_REQUIRE(j >= pf->world || pf->flags[j], KGE_EINVAL, "kge_peer_flags.flags[%d] is NULL", j);
  }
  auto* c = reinterpret_cast<kge_ctx_impl*>(ctx_);
  k_peer_barrier<<<1, 32, 0, (cudaStream_t) stream>>>(f, epoch, c->capturing ? &c->dyn->epoch : nullptr);
  KGE_CUDA_OK(cudaGetLastError());
  return KGE_OK;
}

extern "C" int kge_train_step_dp_next(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                                      const kge_dp* dp, const void* triples, int64_t B, uint64_t seed, uint64_t step,
                                      uint64_t adam_t, kge_grads* push, const kge_grads* slots, const kge_peer_flags* pf,
                                      uint32_t epoch, float* loss_out, const void* next_triples, uint64_t next_step, void* ws,
                                      size_t ws_bytes, void* stream) {
  KGE_REQUIRE(dp && push && slots && push->n_peers > 0 && slots->n_peers > 0, KGE_EINVAL,
              "kge_train_step_dp needs kge_dp and the peer-staging form of kge_grads");
  KGE_REQUIRE(pf && pf->rank >= 0 && pf->rank < KGE_MAX_PEERS && pf->flags[pf->rank], KGE_EINVAL, "bad kge_peer_flags");
  auto* c = reinterpret_cast<kge_ctx_impl*>(ctx);
  if (int rc = train_impl(ctx, cfg, tables, smp, dp, triples, B, seed, step, 1, nullptr, 1, push, nullptr, nullptr, nullptr, ws,
                          ws_bytes, stream, nullptr, 0, next_triples, next_step)) return rc;
  const int slot = (c && !c->capturing) ? c->prof_open : -1;           // the step's profile slot (the halves share it)
  auto mark = [&](int idx) {
    if (slot >= 0) cudaEventRecord(c->ev[slot * 2 * KGE_N_PHASES + idx], (cudaStream_t) stream);
    else if (c && c->capturing && c->gev) { cudaEventRecordWithFlags(c->gev[idx], (cudaStream_t) stream, cudaEventRecordExternal); c->gev_mask |= 1u << 9; }
  };
  mark(2 * 9);
  if (int rc = kge_peer_barrier(ctx, pf, epoch, stream)) return rc;
  mark(2 * 9 + 1);
  if (slot >= 0) c->ev_mask[slot] |= 1u << 9;
  kge_grads sl = *slots;
  sl.abort_word = pf->flags[pf->rank] + KGE_MAX_PEERS;
  if (int rc = kge_train_apply(ctx, cfg, tables, &sl, nullptr, B, adam_t, loss_out, ws, ws_bytes, stream)) return rc;
  return kge_peer_barrier(ctx, pf, epoch + 1, stream);
}
extern "C" int kge_train_step_dp(kge_ctx* ctx, const kge_cfg* cfg, kge_tables* tables, const kge_sampler* smp,
                                 const kge_dp* dp, const void* triples, int64_t B, uint64_t seed, uint64_t step,
                                 uint64_t adam_t, kge_grads* push, const kge_grads* slots, const kge_peer_flags* pf,
                                 uint32_t epoch, float* loss_out, void* ws, size_t ws_bytes, void* stream) {
  return kge_train_step_dp_next(ctx, cfg, tables, smp, dp, triples, B, seed, step, adam_t, push, slots, pf, epoch, loss_out,
                                nullptr, 0, ws, ws_bytes, stream);
}

// ---- a training step captured once as a CUDA graph and replayed (kge_abi.h: kge_step_graph) -----------
namespace kge {
// first node of every captured step: the launch's scalars, patched per launch with
// cudaGraphExecKernelNodeSetParams, land in the device block the other kernels read
__global__ void k_set_dyn(DynScal* d, unsigned long long step, unsigned long long step_next, float lr_t, uint32_t epoch,
                          const void* next_src, float* host_loss, unsigned long long* host_seq, unsigned long long seq) {
  d->step = step; d->step_next = step_next; d->lr_t = lr_t; d->epoch = epoch;
  d->next_src = next_src; d->host_loss = host_loss; d->host_seq = host_seq; d->seq = seq;
}
struct GraphVariant {
  unsigned ev_mask = 0;          // phases whose events this variant records (KGE_OPT_GRAPH_PROFILE)
  cudaGraphExec_t exec = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphNode_t set_node = nullptr;
  int wn_valid_after = 0;
};
}  // namespace kge
struct kge_step_graph {
  kge_ctx_impl* ctx;
  kge_cfg cfg;
  kge_tables tables;         // pointers as of creation (wn_valid / lazy_t are taken from the caller's struct per launch)
  kge_sampler smp; bool has_smp;
  void* stage[2];            // device staging buffers of the batch (two: this batch / the look-ahead batch)
  int64_t B; uint64_t seed;
  float* dev_loss; void* ws; size_t ws_bytes;
  cudaStream_t cap;
  cudaEvent_t gev[2 * KGE_N_PHASES + 2];   // phase boundaries recorded BY the graph (+ whole step), timing enabled; NULL = off
  bool prof;
  unsigned last_mask;
  cudaEvent_t ev[6];         // the context's fork / join events are replaced by these while capturing (an event recorded
                             // inside a capture cannot be waited on by ordinary calls afterwards)
  // variant key: bit 0 shadow table valid, bit 1 this batch was grouped by the previous launch, bit 2 this launch groups
  // the next batch, bit 3 staging parity of this batch
  kge::GraphVariant var[32];
  // data-parallel form (kge_step_graph_create_dp): the captured step is kge_train_step_dp_next
  bool dp_mode;
  kge_dp dp;
  kge_grads push[2], slots[2];   // by staging parity of the exchange (bit 4 of the variant key)
  kge_peer_flags pf;
  // the host's end of the loss (kge_step_graph_wait_loss): a mapped word the graph writes `seq` to behind the loss
  unsigned long long* host_seq; unsigned long long* host_seq_dev; unsigned long long seq;
  bool loss_direct;          // the last launch wrote the loss into the caller's pinned buffer itself
  cudaStream_t last_stream;
  int parity;                // staging buffer the NEXT ungrouped batch goes to
  bool next_grouped;         // the previous launch grouped a batch (into stage[parity])
  uint64_t next_step;
};

static void P_fix_ibase(Prefetched& P, int64_t i_base) { if (P.valid) P.i_base = i_base; }
static int graph_capture(kge_step_graph* g, int key) {
  kge_ctx_impl* ctx = g->ctx;
  kge::GraphVariant& V = g->var[key];
  const int par = (key >> 3) & 1;
  kge_tables T = g->tables;
  T.wn_valid[KGE_T_ENT] = key & 1;
  const Prefetched saved = ctx->pre;
  ctx->pre = Prefetched{};
  if (key & 2) {             // pretend the previous launch grouped this batch into the set this parity uses
    Prefetched& P = ctx->pre;
    P.valid = true; P.triples = g->stage[par]; P.ws = g->ws; P.B = g->B; P.i_base = 0; P.n_ent = g->cfg.n_ent;
    P.seed = g->seed; P.step = 0; P.k = k_eff_of(&g->cfg); P.side = g->cfg.corrupt_side; P.sampler = g->cfg.sampler;
    P.set = par; P.graph = 1;
  }
  cudaEvent_t* slots[6] = {&ctx->ev_fork, &ctx->ev_join, &ctx->ev_mid, &ctx->ev_heavy, &ctx->ev_pre_fork, &ctx->ev_pre_done};
  cudaEvent_t keep[6];
  for (int j = 0; j < 6; ++j) { keep[j] = *slots[j]; *slots[j] = g->ev[j]; }
  struct Restore {
    cudaEvent_t** slots; cudaEvent_t* keep; kge_ctx_impl* c; const Prefetched* saved;
    ~Restore() { for (int j = 0; j < 6; ++j) *slots[j] = keep[j]; c->capturing = 0; c->pre = *saved; }
  } restore{slots, keep, ctx, &saved};
  KGE_CUDA_OK(cudaStreamBeginCapture(g->cap, cudaStreamCaptureModeRelaxed));
  ctx->capturing = 1;
  ctx->gev = g->prof ? g->gev : nullptr;
  ctx->gev_mask = 0;
  k_set_dyn<<<1, 1, 0, g->cap>>>(ctx->dyn, 0ull, 0ull, 0.f, 0u, nullptr, nullptr, nullptr, 0ull);
  if (g->prof) cudaEventRecordWithFlags(g->gev[2 * KGE_N_PHASES], g->cap, cudaEventRecordExternal);
  int rc;
  if (g->dp_mode) {
    const int sp = (key >> 4) & 1;
    P_fix_ibase(ctx->pre, g->dp.i_base);
    rc = kge_train_step_dp_next(reinterpret_cast<kge_ctx*>(ctx), &g->cfg, &T, g->has_smp ? &g->smp : nullptr, &g->dp, g->stage[par],
                                g->B, g->seed, 0, 1, &g->push[sp], &g->slots[sp], &g->pf, 0, g->dev_loss,
                                (key & 4) ? g->stage[1 - par] : nullptr, 0, g->ws, g->ws_bytes, g->cap);
  } else
    rc = train_impl(reinterpret_cast<kge_ctx*>(ctx), &g->cfg, &T, g->has_smp ? &g->smp : nullptr, nullptr, g->stage[par], g->B,
                    g->seed, 0, 1, nullptr, 1, nullptr, nullptr, g->dev_loss, nullptr, g->ws, g->ws_bytes, g->cap, nullptr, 0,
                    (key & 4) ? g->stage[1 - par] : nullptr, 0);
  if (g->prof) cudaEventRecordWithFlags(g->gev[2 * KGE_N_PHASES + 1], g->cap, cudaEventRecordExternal);
  ctx->capturing = 0;
  ctx->gev = nullptr;
  V.ev_mask = ctx->gev_mask;
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(g->cap, &graph);
  if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
  KGE_REQUIRE(e == cudaSuccess && graph, KGE_ECUDA, "step graph capture failed: %s", cudaGetErrorString(e));
  size_t n_root = 0;
  KGE_CUDA_OK(cudaGraphGetRootNodes(graph, nullptr, &n_root));
  KGE_REQUIRE(n_root == 1, KGE_ECUDA, "step graph has %zu root nodes, expected the scalar node alone", n_root);
  KGE_CUDA_OK(cudaGraphGetRootNodes(graph, &V.set_node, &n_root));
  KGE_CUDA_OK(cudaGraphInstantiate(&V.exec, graph, 0));
  V.graph = graph;
  V.wn_valid_after = T.wn_valid[KGE_T_ENT];
  return KGE_OK;
}

extern "C" int kge_step_graph_create(kge_ctx* ctx_, const kge_cfg* cfg, const kge_tables* tables, const kge_sampler* smp,
                                     void* dev_triples, void* dev_triples_next, int64_t B, uint64_t seed, float* dev_loss,
                                     void* ws, size_t ws_bytes, kge_step_graph** out) {
  auto* ctx = reinterpret_cast<kge_ctx_impl*>(ctx_);
  KGE_REQUIRE(ctx && cfg && tables && dev_triples && dev_loss && ws && out && B > 0, KGE_EINVAL, "bad kge_step_graph_create arguments");
  KGE_REQUIRE(dev_triples_next != dev_triples, KGE_EINVAL, "the look-ahead staging buffer must differ from dev_triples");
  if (int rc = check_cfg(cfg)) return rc;
  if (int rc = check_tables(cfg, tables, true)) return rc;
  KGE_REQUIRE(!lazy_requested(tables), KGE_EUNSUPPORTED, "step graphs cover dense tables (not row-lazy / sharded)");
  auto* g = new kge_step_graph();
  g->ctx = ctx; g->cfg = *cfg; g->tables = *tables; g->has_smp = smp != nullptr;
  if (smp) g->smp = *smp;
  g->stage[0] = dev_triples; g->stage[1] = dev_triples_next;
  g->B = B; g->seed = seed; g->dev_loss = dev_loss; g->ws = ws; g->ws_bytes = ws_bytes;
  g->parity = 0; g->next_grouped = false; g->next_step = 0; g->dp_mode = false;
  g->host_seq = nullptr; g->host_seq_dev = nullptr; g->seq = 0; g->loss_direct = false; g->last_stream = nullptr;
  if (cudaHostAlloc(reinterpret_cast<void**>(&g->host_seq), 64, cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer(reinterpret_cast<void**>(&g->host_seq_dev), g->host_seq, 0) != cudaSuccess) {
    delete g; set_error("pinned word for the loss hand-off could not be allocated"); return KGE_ECUDA;
  }
  *g->host_seq = 0;
  if (cudaStreamCreateWithFlags(&g->cap, cudaStreamNonBlocking) != cudaSuccess) { delete g; set_error("stream create failed"); return KGE_ECUDA; }
  for (int j = 0; j < 6; ++j)
    if (cudaEventCreateWithFlags(&g->ev[j], cudaEventDisableTiming) != cudaSuccess) { set_error("event create failed"); return KGE_ECUDA; }
  g->prof = ctx->opt_graph_profile != 0; g->last_mask = 0;
  for (int j = 0; j < 2 * KGE_N_PHASES + 2; ++j) {
    g->gev[j] = nullptr;
    if (g->prof && cudaEventCreate(&g->gev[j]) != cudaSuccess) { set_error("event create failed"); return KGE_ECUDA; }
  }
  *out = g;
  return KGE_OK;
}

extern "C" int kge_step_graph_create_dp(kge_ctx* ctx_, const kge_cfg* cfg, const kge_tables* tables, const kge_sampler* smp,
                                        const kge_dp* dp, const kge_grads* push2, const kge_grads* slots2, const kge_peer_flags* pf,
                                        void* dev_triples, void* dev_triples_next, int64_t B, uint64_t seed, float* dev_loss,
                                        void* ws, size_t ws_bytes, kge_step_graph** out) {
  KGE_REQUIRE(dp && push2 && slots2 && pf, KGE_EINVAL, "dp / push / slots / flags is NULL");
  KGE_REQUIRE(push2[0].n_peers > 0 && slots2[0].n_peers > 0, KGE_EINVAL, "the data-parallel step graph needs the peer-staging form of kge_grads");
  if (int rc = kge_step_graph_create(ctx_, cfg, tables, smp, dev_triples, dev_triples_next, B, seed, dev_loss, ws, ws_bytes, out)) return rc;
  kge_step_graph* g = *out;
  g->dp_mode = true; g->dp = *dp; g->pf = *pf;
  for (int j = 0; j < 2; ++j) { g->push[j] = push2[j]; g->slots[j] = slots2[j]; }
  return KGE_OK;
}

static int graph_launch(kge_step_graph* g, kge_tables* tables, const void* triples, uint64_t step, uint64_t adam_t, uint32_t epoch,
                        int staging_parity, const void* next_triples, uint64_t next_step, float* host_loss_out, void* stream);
extern "C" int kge_step_graph_launch(kge_step_graph* g, kge_tables* tables, const void* triples, uint64_t step, uint64_t adam_t,
                                     const void* next_triples, uint64_t next_step, float* host_loss_out, void* stream) {
  KGE_REQUIRE(g && !g->dp_mode, KGE_EINVAL, "kge_step_graph_launch: not a single-GPU step graph");
  return graph_launch(g, tables, triples, step, adam_t, 0, 0, next_triples, next_step, host_loss_out, stream);
}
extern "C" int kge_step_graph_launch_dp(kge_step_graph* g, kge_tables* tables, const void* triples, uint64_t step, uint64_t adam_t,
                                        uint32_t epoch, int staging_parity, const void* next_triples, uint64_t next_step,
                                        float* host_loss_out, void* stream) {
  KGE_REQUIRE(g && g->dp_mode, KGE_EINVAL, "kge_step_graph_launch_dp: not a data-parallel step graph");
  return graph_launch(g, tables, triples, step, adam_t, epoch, staging_parity & 1, next_triples, next_step, host_loss_out, stream);
}
static int graph_launch(kge_step_graph* g, kge_tables* tables, const void* triples, uint64_t step, uint64_t adam_t, uint32_t epoch,
                        int staging_parity, const void* next_triples, uint64_t next_step, float* host_loss_out, void* stream) {
  KGE_REQUIRE(g && tables && adam_t >= 1, KGE_EINVAL, "bad kge_step_graph_launch arguments");
  KGE_REQUIRE(!next_triples || g->stage[1], KGE_EINVAL, "the graph was created without a look-ahead staging buffer");
  cudaStream_t st = (cudaStream_t) stream;
  const size_t bytes = (size_t) g->B * 3 * (g->cfg.id_dtype ? 8 : 4);
  const bool grouped = g->next_grouped && g->next_step == step;       // the previous launch staged + grouped this batch
  if (!grouped) g->parity = 0;                                         // no staged batch (or not taken up): buffer / set 0
  const int par = g->parity;
  if (!grouped) {
    KGE_REQUIRE(triples, KGE_EINVAL, "triples is NULL and the batch was not staged by the previous launch");
    KGE_CUDA_OK(cudaMemcpyAsync(g->stage[par], triples, bytes, cudaMemcpyDefault, st));
  }
  // The look-ahead batch and the loss cross the host boundary INSIDE the graph when the caller's buffers are visible
  // to the device (pinned / registered host memory, or device memory): k_stage_next reads the batch on the side
  // chain, the kernel that makes the loss final stores it into the host buffer followed by a sequence word.
  // Pageable host memory keeps the copies queued around the graph.
  auto device_view = [](const void* ptr) -> void* {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) return at.devicePointer;
    return nullptr;
  };
  static const bool direct_off = [] { const char* e = getenv("KGE_GRAPH_DIRECT_IO"); return e && atoi(e) == 0; }();
  const void* next_src = nullptr;
  if (next_triples) {
    next_src = direct_off ? nullptr : device_view(next_triples);
    if (!next_src) KGE_CUDA_OK(cudaMemcpyAsync(g->stage[1 - par], next_triples, bytes, cudaMemcpyDefault, st));
  }
  float* loss_direct = nullptr;
  if (host_loss_out && !direct_off) loss_direct = reinterpret_cast<float*>(device_view(host_loss_out));
  const int key = (tables->wn_valid[KGE_T_ENT] ? 1 : 0) | (grouped ? 2 : 0) | (next_triples ? 4 : 0) | (par << 3) | (staging_parity << 4);
  kge::GraphVariant& V = g->var[key];
  if (!V.exec) { if (int rc = graph_capture(g, key)) return rc; }
  unsigned long long a0 = step, a1 = next_step;
  float lr_t = g->cfg.optimizer == KGE_OPTIM_ADAM ? adam_lr(g->cfg.lr, g->cfg.beta1, g->cfg.beta2, adam_t) : g->cfg.lr;
  DynScal* dptr = g->ctx->dyn;
  unsigned long long seq = loss_direct ? ++g->seq : g->seq;
  unsigned long long* seq_dev = g->host_seq_dev;
  void* args[9] = {&dptr, &a0, &a1, &lr_t, &epoch, &next_src, &loss_direct, &seq_dev, &seq};
  cudaKernelNodeParams np = {};
  np.func = (void*) kge::k_set_dyn;
  np.gridDim = dim3(1, 1, 1); np.blockDim = dim3(1, 1, 1); np.sharedMemBytes = 0; np.kernelParams = args; np.extra = nullptr;
  KGE_CUDA_OK(cudaGraphExecKernelNodeSetParams(V.exec, V.set_node, &np));
  KGE_CUDA_OK(cudaGraphLaunch(V.exec, st));
  if (host_loss_out && !loss_direct) KGE_CUDA_OK(cudaMemcpyAsync(host_loss_out, g->dev_loss, sizeof(float), cudaMemcpyDeviceToHost, st));
  g->loss_direct = loss_direct != nullptr;
  g->last_stream = st;
  tables->wn_valid[KGE_T_ENT] = V.wn_valid_after;
  g->last_mask = V.ev_mask;
  g->next_grouped = next_triples != nullptr;
  g->next_step = next_step;
  if (next_triples) g->parity = 1 - par;         // the grouped batch sits in the other buffer / set
  return KGE_OK;
}

// Returns once the loss of the last launch is in the caller's host buffer.  When the graph stored it there itself
// (pinned buffer, one GPU) that is BEFORE the update kernels have run: the host may read the loss and enqueue the
// next step while they finish (stream order keeps the steps in sequence).  Otherwise: a stream synchronisation.
extern "C" int kge_step_graph_wait_loss(kge_step_graph* g) {
  KGE_REQUIRE(g, KGE_EINVAL, "graph is NULL");
  if (!g->loss_direct) {
    KGE_CUDA_OK(cudaStreamSynchronize(g->last_stream));
    return KGE_OK;
  }
  const unsigned long long want = g->seq;
  volatile unsigned long long* w = g->host_seq;
  for (unsigned long long spin = 0;; ++spin) {
    if (*w >= want) return KGE_OK;
    if ((spin & 0xfffff) == 0xfffff) {           // every ~1M polls: has the stream died or finished without the store?
      cudaError_t e = cudaStreamQuery(g->last_stream);
      if (e == cudaSuccess) { KGE_REQUIRE(*w >= want, KGE_ECUDA, "the step finished without publishing its loss"); return KGE_OK; }
      KGE_REQUIRE(e == cudaErrorNotReady, KGE_ECUDA, "stream error while waiting for the loss: %s", cudaGetErrorString(e));
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
}

extern "C" int kge_step_graph_phases(kge_step_graph* g, float* phase_ms, float* total_ms) {
  KGE_REQUIRE(g && phase_ms && total_ms, KGE_EINVAL, "bad kge_step_graph_phases arguments");
  KGE_REQUIRE(g->prof, KGE_EINVAL, "the graph was created without KGE_OPT_GRAPH_PROFILE");
  KGE_CUDA_OK(cudaDeviceSynchronize());
  for (int p = 0; p < KGE_N_PHASES; ++p) {
    phase_ms[p] = 0.f;
    if (g->last_mask & (1u << p)) KGE_CUDA_OK(cudaEventElapsedTime(&phase_ms[p], g->gev[2 * p], g->gev[2 * p + 1]));
  }
  KGE_CUDA_OK(cudaEventElapsedTime(total_ms, g->gev[2 * KGE_N_PHASES], g->gev[2 * KGE_N_PHASES + 1]));
  return KGE_OK;
}

extern "C" int kge_step_graph_destroy(kge_step_graph* g) {
  if (!g) return KGE_OK;
  for (auto& V : g->var) {
    if (V.exec) cudaGraphExecDestroy(V.exec);
    if (V.graph) cudaGraphDestroy(V.graph);
  }
  if (g->cap) cudaStreamDestroy(g->cap);
  if (g->host_seq) cudaFreeHost(g->host_seq);
  for (int j = 0; j < 6; ++j) if (g->ev[j]) cudaEventDestroy(g->ev[j]);
  for (int j = 0; j < 2 * KGE_N_PHASES + 2; ++j) if (g->gev[j]) cudaEventDestroy(g->gev[j]);
  delete g;
  return KGE_OK;
}
