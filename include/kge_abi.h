/* kge_abi.h — C-ABI of libkge_b200.so: the B200-native batch hot path of
 * melissakou/knowledge-graph-embedding (training step + filtered ranking).
 *
 * The reference has no FFI: its "operator API" is the Python protocol of
 * KGE/models/base_model/BaseModel.py.  Each entry point below names the reference
 * interface it replaces; INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *  - plain C, no CUDA / torch types: device pointers are void* / float*, the stream
 *    is a cudaStream_t passed as void*.
 *  - the caller owns every buffer (tables, workspace, outputs); compute calls never
 *    allocate, free or retain device memory.  The context itself owns a 256-byte
 *    device counter block and one side stream (kge_train_step forks the key sort onto
 *    it and joins before returning control of the caller's stream), so one context
 *    serves one call at a time.
 *  - every call enqueues on the given stream and returns; no host sync.
 *  - return 0 on success, a negative kge_status otherwise; kge_last_error() gives the
 *    thread-local message.  There is no CPU fallback: without a CUDA device the
 *    compute calls return KGE_ECUDA.
 *  - all floating point is fp32 (every reference table is np.float32, TransE.py:96).
 *  - table rows are `cols` logical floats at a stride of `ld` floats; ld % 4 == 0 and
 *    16-byte aligned bases are required (128-bit row access); padding must be zero.
 */
#ifndef KGE_ABI_H
#define KGE_ABI_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KGE_ABI_VERSION 12
#define KGE_MAX_TABLES 4

typedef enum { KGE_OK = 0, KGE_EINVAL = -1, KGE_EUNSUPPORTED = -2, KGE_ECUDA = -3,
               KGE_EWORKSPACE = -4 } kge_status;

/* KGE/models/{translating_based,semantic_based}/ *.py */
typedef enum { KGE_UM = 0, KGE_SE = 1, KGE_TRANSE = 2, KGE_TRANSH = 3, KGE_TRANSR = 4,
               KGE_TRANSD = 5, KGE_ROTATE = 6, KGE_DISTMULT = 7, KGE_RESCAL = 8 } kge_model;
/* KGE/score.py:49 LpDistance, :65 LpDistancePow, :78 Dot */
typedef enum { KGE_SCORE_LP = 0, KGE_SCORE_LP_POW = 1, KGE_SCORE_DOT = 2 } kge_score_kind;
/* KGE/loss.py:49,84,115,145,184 */
typedef enum { KGE_LOSS_HINGE = 0, KGE_LOSS_LOGISTIC = 1, KGE_LOSS_BCE = 2,
               KGE_LOSS_SELF_ADV = 3, KGE_LOSS_SQUARE = 4 } kge_loss_kind;
/* BaseModel.py:243-246 accepts any Keras optimizer.  Row update of the touched rows (TF 2.5 OptimizerV2
 * sparse apply on the de-duplicated, summed rows):
 *   ADAM    m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; w -= lr_t m / (sqrt(v) + eps)   — m, v, w of EVERY row move
 *   SGD     w -= lr g                                  (momentum = 0; slots m, v unused)
 *   ADAGRAD v += g^2; w -= lr g / (sqrt(v) + eps)      (v = accumulator, caller initialises it to
 *                                                       initial_accumulator_value = 0.1; m unused)
 * cfg.lr / cfg.eps carry learning_rate / epsilon.  Every optimizer works on every table form (dense, row-lazy,
 * sharded): a row without a gradient is not moved by SGD / Adagrad, so its skipped steps replay the constraint only. */
typedef enum { KGE_OPTIM_ADAM = 0, KGE_OPTIM_SGD = 1, KGE_OPTIM_ADAGRAD = 2 } kge_opt_kind;
/* BaseModel.py:49 corrupt_side */
typedef enum { KGE_SIDE_H = 0, KGE_SIDE_T = 1, KGE_SIDE_HT = 2 } kge_side;
/* KGE/ns_strategy.py:23 UniformStrategy, :67 TypedStrategy */
typedef enum { KGE_NS_UNIFORM = 0, KGE_NS_TYPED = 1 } kge_sampler_kind;

/* Table slots (model_weights keys):
 *   0 ent_emb (RotatE: cols = 2*d, re/im interleaved, RotatE.py:96)
 *   1 ent_proj (TransD)
 *   2 rel_emb (TransE/H/R/D, RotatE) | rel_inter (DistMult; RESCAL d*d) | rel_proj_h (SE d*d)
 *   3 rel_hyper (TransH) | rel_proj (TransR d_e*d_r row-major; TransD d_r) | rel_proj_t (SE) */
enum { KGE_T_ENT = 0, KGE_T_ENT_PROJ = 1, KGE_T_REL_A = 2, KGE_T_REL_B = 3 };

typedef struct kge_table {
  float* w;          /* weights                                   */
  float* m;          /* Adam first moment  (NULL for score/rank)   */
  float* v;          /* Adam second moment (NULL for score/rank)   */
  float* wn;         /* optional shadow: the row-normalised table the NEXT batch starts from, same
                        layout as w (see kge_tables.wn_valid); NULL = none                        */
  uint32_t* last;    /* optional, [rows]: row-lazy replay (see kge_tables.lazy_t); NULL = dense    */
  int64_t rows;
  int32_t cols;      /* logical floats per row                     */
  int32_t ld;        /* row stride in floats, multiple of 4        */
} kge_table;

/* wn_valid[s] != 0: t[s].wn holds normalise(t[s].w) — what the start-of-batch constraint
 * (TransE.py:171-172, DistMult.py:162, UM.py:153) would assign.  kge_train_step then skips the
 * full-table constraint pass: rows are gathered from wn, and the row update (which has the new row
 * in registers anyway) writes both the reference-visible w and the next batch's wn.  The library
 * sets the flag after a training step that maintained wn; the CALLER clears it whenever it writes
 * w itself (assign / restore / init-time constraint).  Entity table of UM / TransE / DistMult with
 * constraint = True only; ignored elsewhere.  Results are bit-identical with and without wn. */
/* Row-lazy replay (entity tables much larger than a batch).  The reference
 * rewrites EVERY row of every table on every batch: the start-of-batch constraint assigns the whole
 * normalised / clipped table and Keras Adam's sparse apply decays m, v and moves w on all rows
 * (SURVEY.md App. B.2).  A row without a gradient follows a closed recurrence, so when the entity
 * tables carry `last` (uint32 per row, zero-initialised together with m and v), kge_train_step
 * only reads and writes the rows the batch looks up: last[row] = the Adam step the stored
 * w / m / v of that row belong to, and the skipped steps are replayed in registers, with the
 * dense kernels' arithmetic, when the row is next looked up.  lazy_t = the Adam step the tables
 * are logically at (library-maintained like wn_valid; a training call must pass
 * adam_t == lazy_t + 1).  kge_flush_rows replays every row to lazy_t: afterwards w / m / v are
 * bit-identical to the dense run.  The caller flushes before it reads w itself, before
 * kge_score / kge_rank, and before it writes w (then it also sets last[] = lazy_t or resets
 * both to 0 with m, v).  A validation call (is_train = 0) flushes by itself.  Vector models
 * without a dense regulariser, single-GPU step only; every used entity slot must carry `last`. */
typedef struct kge_tables {
  kge_table t[KGE_MAX_TABLES];   /* unused slots: w == NULL */
  int32_t wn_valid[KGE_MAX_TABLES];
  int64_t lazy_t;
} kge_tables;

typedef struct kge_cfg {
  int32_t abi_version;           /* KGE_ABI_VERSION */
  int32_t model;                 /* kge_model */
  int32_t score;                 /* kge_score_kind (ignored by DistMult / RESCAL) */
  int32_t loss;                  /* kge_loss_kind */
  int32_t sampler;               /* kge_sampler_kind */
  int32_t corrupt_side;          /* kge_side */
  int32_t negative_ratio;        /* as given by the user; 'h+t' uses negative_ratio/2 per side */
  int32_t constraint;            /* model's constraint flag */
  int32_t id_dtype;              /* 0 = int32 triples, 1 = int64 (test_data_utils.py:82,90) */
  int32_t d_ent, d_rel;          /* embedding sizes (RotatE: complex dimension) */
  int32_t optimizer;             /* kge_opt_kind (0 = Adam: structs zeroed by older callers keep their meaning) */
  float score_p;                 /* LpDistance.p; INFINITY allowed (score.py:60) */
  float margin, temperature;     /* loss.py:77, 171-172 */
  float constraint_weight;       /* TransH.py:66, DistMult.py:40, RESCAL.py:39 */
  float lr, beta1, beta2, eps;   /* tf.optimizers.Adam(): 1e-3, .9, .999, 1e-7 */
  float clip_norm;               /* 5.0, BaseModel.py:327 */
  float rotate_scale;            /* pi / limit, RotatE.py:160 */
  int64_t n_ent, n_rel;
} kge_cfg;

/* Negative-sampler data (device pointers).  Uniform: pool == NULL means arange(n_ent)
 * (BaseModel.py:257).  Typed: CSR of metadata["type2inds"] (BaseModel.py:259-263). */
typedef struct kge_sampler {
  const int64_t* pool;  int64_t pool_len;
  const int64_t* ent_type;      /* [n_ent] type id of each entity                 */
  const int64_t* type_off;      /* [n_types+1]                                    */
  const int64_t* type_members;  /* members, grouped by type                       */
  const int64_t* ent_pos;       /* [n_ent] position inside own type pool, or -1   */
  int64_t n_types;
} kge_sampler;

/* Known-positive filter for ranking (replaces the two scans of positive_X per query at
 * BaseModel.py:646-650): ids[off[q] .. END(q)) are the candidate entities to exclude for query q, with
 * END(q) = end[q] when `end` is given, else off[q + 1] (a packed CSR).  off == NULL means unfiltered.
 * An entity must not be listed twice for one query.  kge_filter_lookup fills this struct from a
 * kge_filter_index (ranges into the index's own sorted entity array: nothing is copied). */
typedef struct kge_filter {
  const int64_t* off;  const int64_t* ids;
  const int64_t* end;   /* nullable */
} kge_filter;

typedef struct kge_ctx kge_ctx;

int kge_abi_version(void);
const char* kge_last_error(void);
int kge_ctx_create(int device, kge_ctx** out);
int kge_ctx_destroy(kge_ctx* ctx);

/* Bytes of workspace kge_train_step (batch B) / kge_rank (Q queries) need. */
int kge_workspace_bytes(const kge_cfg* cfg, int64_t B, int64_t Q, size_t* out);

/* Replaces NegativeSampler.__call__ + __corrupt_h/__corrupt_t (ns_strategy.py:39,94;
 * BaseModel.py:332-408).  neg_out: int64[B*k_eff] entity ids in the reference's row
 * order (row i*k_eff+c belongs to positive i; 'h+t': even c = head corruption).
 * Stream = counter-based Philox4x32-10, see oracle/philox.py for the spec. */
int kge_sample_negatives(kge_ctx*, const kge_cfg*, const kge_sampler*, const void* triples,
                         int64_t B, uint64_t seed, uint64_t step, int64_t* neg_out, void* stream);

/* Replaces model.score_hrt(h, r, t) on id vectors (BaseModel.py:410 + each model). */
int kge_score(kge_ctx*, const kge_cfg*, const kge_tables*, const void* h, const void* r,
              const void* t, int64_t n, float* out, void* stream);

/* Start-of-batch constraint of _constraint_loss (e.g. TransE.py:171-172) on its own;
 * kge_train_step runs it itself.  Exposed for init-time normalisation
 * (TransE.py:108-109, RESCAL.py:109-115): op 0 = normalise L2 rows to 1, 1 = clip L2
 * rows to <= 1, applied to table slot `slot`. */
int kge_constrain_table(kge_ctx*, const kge_tables*, int slot, int op, void* stream);

/* Replaces KGEModel.__run_single_batch (BaseModel.py:293-330): constraint, negative
 * sampling (or neg_override: int64[B*k_eff] ids as above), scores, loss, and when
 * is_train: analytic backward, per-variable clip_by_norm, TF-semantics Adam step
 * number adam_t (1-based).  loss_out: device float.  neg_out (nullable) receives
 * the negatives used. */
int kge_train_step(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*,
                   const void* triples, int64_t B, uint64_t seed, uint64_t step,
                   uint64_t adam_t, const int64_t* neg_override, int is_train,
                   float* loss_out, int64_t* neg_out, void* ws, size_t ws_bytes, void* stream);

/* kge_train_step for a caller whose batch and loss live in HOST memory (what the reference's
 * loop does per batch, BaseModel.py:139-147: iterator -> step -> loss.numpy()): copies the
 * (B, 3) ids from `host_triples` (pinned for a truly asynchronous copy) into the caller's device
 * staging buffer `dev_triples`, runs the step, and copies the loss to `host_loss_out`; all three
 * are enqueued on `stream`, no host sync — the loss is valid once the stream has been
 * synchronised.  `dev_loss` is a device float the step writes first. */
int kge_train_step_host(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*,
                        const void* host_triples, void* dev_triples, int64_t B, uint64_t seed,
                        uint64_t step, uint64_t adam_t, int is_train, float* dev_loss,
                        float* host_loss_out, void* ws, size_t ws_bytes, void* stream);

/* kge_train_step (training) that also groups the NEXT batch's gradient records while this step runs.  The
 * reference's loop (BaseModel.py:139-147) draws batch t + 1 from its iterator only after step t; here the caller
 * that already holds it (a device-resident epoch, an iterator with one batch of look-ahead) passes it along: the
 * record keys of a batch depend on its ids and the Philox stream (seed, next_step) only, so their counting sort —
 * otherwise the longer of the two concurrent branches of a step — runs on a stream of its own beside this step's
 * kernels, into the second of two grouping sets of the workspace.  The next call that names the same
 * (triples pointer, B, seed, step, workspace) skips its own grouping; any other call groups inline as before.
 * next_triples must stay unchanged until that call.  NULL = no look-ahead.  Results are identical either way.
 * Dense tables (not row-lazy), counting-sort mode; otherwise next_* is ignored.
 * _host_next: kge_train_step_host with the look-ahead batch in host memory: host_next is copied into its own
 * staging buffer dev_next ahead of the grouping, and the call whose dev_triples is that buffer copies nothing. */
int kge_train_step_next(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*, const void* triples, int64_t B,
                        uint64_t seed, uint64_t step, uint64_t adam_t, float* loss_out, const void* next_triples,
                        uint64_t next_step, void* ws, size_t ws_bytes, void* stream);
int kge_train_step_host_next(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*, const void* host_triples,
                             void* dev_triples, int64_t B, uint64_t seed, uint64_t step, uint64_t adam_t,
                             float* dev_loss, float* host_loss_out, const void* host_next, void* dev_next,
                             uint64_t next_step, void* ws, size_t ws_bytes, void* stream);

/* A training step captured ONCE as a CUDA graph and replayed: the ~20 launches / event calls of kge_train_step
 * become one graph launch, and the kernels of a step are released by the GPU's own dependency tracking instead of
 * the host's enqueue order.  What changes from step to step — the Philox step, lr_t of the Adam step, the
 * look-ahead step — lives in a device block written by the graph's first node, whose arguments are patched per
 * launch (cudaGraphExecKernelNodeSetParams).  Per launch the host issues one node update and one graph launch — no
 * host synchronisation.  The look-ahead batch and the loss cross the host boundary INSIDE the graph when the caller's
 * buffers are visible to the device (pinned / registered host memory, or device memory): a staging kernel on the
 * look-ahead chain reads next_triples, and the kernel that makes the loss final stores it into host_loss_out; with
 * pageable host memory the same transfers are copies queued in front of / behind the graph (ABI v12; before, always).
 *   kge_step_graph_create : fixes cfg, table pointers, sampler, batch size B, seed, loss word and workspace.
 *                           dev_triples / dev_triples_next: (B, 3) device staging buffers of cfg->id_dtype the
 *                           graph's kernels read (the second one is for the look-ahead batch, NULL = none).
 *                           Capturing happens lazily at the first launch of each variant (shadow table valid or
 *                           not, batch grouped ahead or not, look-ahead or not, staging parity).
 *   kge_step_graph_launch : one training step on `triples` (host or device pointer; ignored when the previous
 *                           launch passed this batch as next_triples with next_step == step) with Adam step
 *                           adam_t (1-based); next_triples (nullable) = the batch of the following launch, grouped
 *                           under this step as kge_train_step_next does.  `tables` is the caller's struct: its
 *                           wn_valid is read and updated as kge_train_step does.
 * Dense tables, training steps, single GPU; results are bit-identical to kge_train_step.  One graph serves one
 * stream at a time (the step's scalars live in the context's device block). */
typedef struct kge_step_graph kge_step_graph;
int kge_step_graph_create(kge_ctx*, const kge_cfg*, const kge_tables*, const kge_sampler*, void* dev_triples,
                          void* dev_triples_next, int64_t B, uint64_t seed, float* dev_loss, void* ws, size_t ws_bytes,
                          kge_step_graph** out);
int kge_step_graph_launch(kge_step_graph*, kge_tables*, const void* triples, uint64_t step, uint64_t adam_t,
                          const void* next_triples, uint64_t next_step, float* host_loss_out, void* stream);
int kge_step_graph_destroy(kge_step_graph*);
/* Blocks the host until the loss of the LAST launch is in its host_loss_out.  When the graph stored it there itself
 * (pinned buffer, single-GPU graph) this returns as soon as the loss is final — before the update kernels of the
 * step have run — so a loop that reads every step's loss (the reference's loss.numpy() per batch,
 * BaseModel.py:139-145) enqueues the next step under them; stream order keeps the steps in sequence.  Otherwise
 * it is cudaStreamSynchronize on the launch's stream.  (ABI v12) */
int kge_step_graph_wait_loss(kge_step_graph*);
/* Phase times of the LAST launch of a graph created while KGE_OPT_GRAPH_PROFILE was 1 (the graph then records
 * its own phase-boundary events: measuring, not for production): phase_ms[KGE_N_PHASES] as kge_profile_end,
 * *total_ms = first to last node.  Synchronises the device. */
int kge_step_graph_phases(kge_step_graph*, float* phase_ms, float* total_ms);

/* Data-parallel form of kge_train_step (the reference is single-device): every
 * rank holds all tables and B of the step's B_global positives, [i_base, i_base + B).
 *   kge_train_grad  : constraint, sampling on the GLOBAL Philox counters (the negatives equal a
 *                     one-GPU step on the whole batch), scores, loss terms normalised by
 *                     B_global, backward, per-row segmented sum -> this rank's summed gradient
 *                     per table and scal_out[0] = loss share, scal_out[1+s] = sum of squared
 *                     norms of slot s over this rank's lookup occurrences.  Tables untouched.
 *   (caller)        : sum `grads` and scal over ranks (NCCL all-reduce; any order-fixed reduction
 *                     keeps replicas bit-identical).
 *   kge_train_apply : per-variable clip_by_norm from the summed norms, dense regularisers,
 *                     TF-semantics Adam on every row; loss_out = summed loss (+ regulariser).
 * grads.dense[s]: [rows x ld] fp32 with the table's own ld (every row is written by _grad).
 * (Tables where records << rows do not use this form at all: they are sharded, kge_shard below.) */
typedef struct kge_dp {
  int32_t world, rank;
  int64_t B_global;              /* positives of the whole step */
  int64_t i_base;                /* global index of this rank's first positive */
} kge_dp;
#define KGE_MAX_PEERS 8
typedef struct kge_grads {
  float* dense[KGE_MAX_TABLES];
  /* Fused exchange over peer memory (NVLink / NVSwitch), n_peers > 0: replaces `dense` / scal.
   *   kge_train_grad : peer[j][s] / peer_scal[j] = the address, mapped into THIS process, of this
   *                    rank's slot in rank j's staging buffer (j == rank is local memory); the
   *                    gradient kernels store every row sum to all n_peers destinations, so the
   *                    all-gather rides on the kernel's own stores.
   *   (caller)       : one cross-GPU barrier (all ranks' kernels done, stores visible).
   *   kge_train_apply: peer[j][s] / peer_scal[j] = LOCAL address of rank j's slot; the update
   *                    sums the n_peers slots in rank order (identical bits on every rank).
   * The caller alternates two staging buffers between steps (a rank may still be applying
   * step i while a faster peer already pushes step i + 1). */
  int32_t n_peers;
  float* peer[KGE_MAX_PEERS][KGE_MAX_TABLES];
  float* peer_scal[KGE_MAX_PEERS];
  /* Owner-partitioned form of the fused exchange (n_peers > 0 and owner_rows[s] > 0 for every used
   * slot): rank j owns rows [j * owner_rows[s], (j+1) * owner_rows[s]) of table s — a reduce-scatter /
   * all-gather carried by the kernels' own stores instead of an all-reduce.
   *   kge_train_grad : the row sum of row r goes to peer[r / owner_rows[s]][s] ONLY.
   *   (caller)       : cross-GPU barrier.
   *   kge_train_apply: this rank (owner_rank) sums the n_peers slots of ITS rows in rank order,
   *                    applies clip + Adam with its own m, v (m, v of rows it does not own are never
   *                    touched and stay zero) and stores the new row into w_peer[j][s] of EVERY rank
   *                    (and normalise(row) into wn_peer[j] when the shadow table is in use): the tables
   *                    must live in peer-mapped memory.  w_peer[owner_rank][s] == tables.t[s].w.
   *   (caller)       : cross-GPU barrier before the next step gathers.
   * Per rank and step the NVLink traffic is one gradient table out and (n_peers - 1) / n_peers of a
   * weight table out, and the update work is divided by n_peers; replicas are identical by
   * construction.  Vector models without a dense regulariser. */
  int32_t owner_rank;
  int32_t owner_rows[KGE_MAX_TABLES];
  float* w_peer[KGE_MAX_PEERS][KGE_MAX_TABLES];
  float* wn_peer[KGE_MAX_PEERS];
  /* kge_train_apply: this rank's error word of kge_peer_barrier (flags[rank] + KGE_MAX_PEERS), or NULL.
   * When it reads KGE_PEER_DEAD the apply half leaves every table as it is (a peer never arrived: its
   * staging slots are stale) and loss_out = NaN; the host reads the word at its next sync and raises. */
  const uint32_t* abort_word;
  /* Optional NVSwitch multicast (NVLS) addresses of the weight tables / the shadow table, i.e. one address
   * whose stores land in the same place of EVERY rank's replica (CUDA multicast objects; torch symmetric
   * memory's multicast_ptr).  When given, the owner's apply half writes each new row once with multimem.st
   * instead of once per rank: outbound NVLink bytes of the all-gather drop from n_peers - 1 replicas to one.
   * NULL = unicast stores through w_peer / wn_peer. */
  float* w_mc[KGE_MAX_TABLES];
  float* wn_mc;
  /* Owner-partitioned apply with the shadow table in use: != 0 keeps the new w row at its owner and sends only
   * wn = normalise(w) to the replicas — all a training step ever reads from another rank's rows (it gathers
   * from wn, and the owner's update starts from wn as well): half the all-gather bytes.  The replicas' w of
   * rows they do not own then go stale; the caller gathers w from the owners before anything reads it
   * (evaluation, validation batches, checkpoints: KGE._b200.engine.Engine.dp_sync_weights). */
  int32_t w_private;
} kge_grads;
/* Cross-GPU barrier over peer memory, enqueued on `stream` like a kernel (no host involvement):
 * flags[j] = rank j's flag array (uint32[KGE_MAX_PEERS + 1], zeroed once) as mapped HERE.  Every rank calls
 * it with the same strictly increasing `epoch`: the kernel fences its stream's earlier stores at system
 * scope, stores epoch into flags[j][rank] of every peer and spins until flags[rank][j] >= epoch for all
 * j.  The spin is bounded (~20 s of SM clocks) so a lost peer cannot hang the GPU: on expiry the kernel
 * writes KGE_PEER_DEAD into the error word flags[rank][KGE_MAX_PEERS] — a sticky, fatal condition: every
 * later kge_train_apply given the word (kge_grads.abort_word; kge_train_step_dp passes it itself) skips its
 * update, and the caller must read the word whenever it synchronises and stop (KGE._b200.engine raises). */
#define KGE_PEER_DEAD 0xdeadu
typedef struct kge_peer_flags {
  int32_t world, rank;
  uint32_t* flags[KGE_MAX_PEERS];
} kge_peer_flags;
int kge_peer_barrier(kge_ctx*, const kge_peer_flags*, uint32_t epoch, void* stream);

/* One data-parallel step in ONE call (host work of a single-GPU step): kge_train_grad(push) ->
 * kge_peer_barrier(epoch) -> kge_train_apply(slots) -> kge_peer_barrier(epoch + 1).  `push` / `slots`
 * are the kge_grads of the two halves (n_peers > 0: peer staging, store-to-all or owner-partitioned). */
int kge_train_step_dp(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*, const kge_dp*,
                      const void* triples, int64_t B, uint64_t seed, uint64_t step, uint64_t adam_t,
                      kge_grads* push, const kge_grads* slots, const kge_peer_flags*, uint32_t epoch,
                      float* loss_out, void* ws, size_t ws_bytes, void* stream);

/* kge_train_step_dp with the look-ahead batch of kge_train_step_next: next_triples = this rank's share of the
 * NEXT step's positives (same B, same i_base), grouped while this step and its exchange run. */
int kge_train_step_dp_next(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*, const kge_dp*,
                           const void* triples, int64_t B, uint64_t seed, uint64_t step, uint64_t adam_t,
                           kge_grads* push, const kge_grads* slots, const kge_peer_flags*, uint32_t epoch,
                           float* loss_out, const void* next_triples, uint64_t next_step, void* ws, size_t ws_bytes,
                           void* stream);
/* The data-parallel step as a captured graph (kge_step_graph above): push2 / slots2 = the kge_grads of the two
 * staging parities the caller alternates between, kge_dp / flags as for kge_train_step_dp.  Per launch the
 * caller passes the barrier epoch (strictly increasing by 2, the graph's barriers use epoch and epoch + 1) and
 * the staging parity of this step.  Every rank launches the same sequence. */
int kge_step_graph_create_dp(kge_ctx*, const kge_cfg*, const kge_tables*, const kge_sampler*, const kge_dp*,
                             const kge_grads* push2, const kge_grads* slots2, const kge_peer_flags*, void* dev_triples,
                             void* dev_triples_next, int64_t B, uint64_t seed, float* dev_loss, void* ws,
                             size_t ws_bytes, kge_step_graph** out);
int kge_step_graph_launch_dp(kge_step_graph*, kge_tables*, const void* triples, uint64_t step, uint64_t adam_t,
                             uint32_t epoch, int staging_parity, const void* next_triples, uint64_t next_step,
                             float* host_loss_out, void* stream);

/* Matrix models (SE / TransR / RESCAL without its dense regulariser): kge_train_grad / kge_train_apply through
 * grads.dense only (n_peers = 0) — dense[slot] of a matrix table is the [n_rel x ld] sum of this rank's pair outer
 * products (k_update_mat<2>), kge_train_apply steps every element from the rank-summed table (k_mat_apply). */
int kge_train_grad(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*, const kge_dp*,
                   const void* triples, int64_t B, uint64_t seed, uint64_t step,
                   const int64_t* neg_override, kge_grads* out, float* scal_out /* [8] device */,
                   int64_t* neg_out, void* ws, size_t ws_bytes, void* stream);
int kge_train_apply(kge_ctx*, const kge_cfg*, kge_tables*, const kge_grads* summed,
                    const float* scal_summed /* [8] device */, int64_t B, uint64_t adam_t,
                    float* loss_out, void* ws, size_t ws_bytes, void* stream);

/* Training step with the ENTITY TABLES SHARDED over the GPUs of one NVLink / NVSwitch node
 * (configuration C5; the reference is single-device).  Rank g owns entity rows
 * [g * shard_rows, min(n_ent, (g+1) * shard_rows)): kge_tables slots 0 / 1 of the calls below hold
 * THIS RANK'S ROWS ONLY (rows = size of the shard) with m, v and `last` (the owner's update is the
 * row-lazy one, kge_tables.lazy_t); relation tables are replicated.  w of every shard, and the three
 * inboxes, are mapped into every rank (peer memory: CUDA VMM / torch symmetric memory).  One step,
 * all ranks in lock-step, B of the step's B_global positives each (kge_dp as for kge_train_grad):
 *   kge_shard_lookup : negatives on the global Philox counters, unique looked-up rows, their ids
 *                      stored into the owners' id inboxes over NVLink.
 *   (caller)           cross-GPU barrier.
 *   kge_shard_pre    : owner: de-duplicates the ids of all sources, replays the skipped steps of
 *                      those rows and applies the start-of-batch constraint (w of its shard).
 *   (caller)           cross-GPU barrier.
 *   kge_shard_grad   : scores / loss / backward with the rows gathered straight from the owners'
 *                      memory; one summed gradient row per looked-up row is stored into the owner's
 *                      value inbox; relation-table gradients go to grads.dense[2..3] and
 *                      scal_out = [norm^2 x4, loss share] exactly as kge_train_grad does.
 *   (caller)           all-reduce (sum) of grads.dense[2..3] and scal — also the barrier that
 *                      makes every rank's inbox stores visible.
 *   kge_shard_apply  : owner: contributions summed in rank order, global clip, Adam on its rows;
 *                      every rank: the identical update of the replicated relation tables.
 * The same workspace must be passed to lookup / grad / apply of one step.  stamp / idx_of / work /
 * contrib are owner-local scratch, zeroed ONCE by the caller.  Vector models without a dense
 * regulariser.  With world == 1 the step is bit-identical to kge_train_step on row-lazy tables. */
typedef struct kge_shard {
  int32_t world, rank;
  int64_t shard_rows;
  int64_t cap;                               /* inbox rows per (owner, source): >= 2 B + B k_eff of any source */
  const float* w_peer[KGE_MAX_PEERS][2];     /* entity slot s of rank j's shard, as mapped HERE */
  uint32_t* ids_peer[KGE_MAX_PEERS];         /* rank j's id inbox      uint32[world][cap]       */
  uint32_t* cnt_peer[KGE_MAX_PEERS];         /* rank j's count inbox   uint32[world]            */
  float* val_peer[KGE_MAX_PEERS][2];         /* rank j's value inbox of slot s: float[world][cap][ld] */
  uint32_t* stamp;                           /* [shard_rows]                                    */
  uint32_t* idx_of;                          /* [shard_rows]                                    */
  uint32_t* work;                            /* [4 + 3 * min(shard_rows, world * cap)]          */
  uint32_t* contrib;                         /* [min(shard_rows, world * cap) * world]          */
} kge_shard;
int kge_shard_lookup(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*, const kge_shard*, const kge_dp*,
                     const void* triples, int64_t B, uint64_t seed, uint64_t step, const int64_t* neg_override,
                     void* ws, size_t ws_bytes, void* stream);
int kge_shard_pre(kge_ctx*, const kge_cfg*, kge_tables*, const kge_shard*, uint64_t adam_t, void* stream);
int kge_shard_grad(kge_ctx*, const kge_cfg*, kge_tables*, const kge_sampler*, const kge_shard*, const kge_dp*,
                   const void* triples, int64_t B, uint64_t seed, uint64_t step, const int64_t* neg_override,
                   kge_grads* out, float* scal_out /* [8] device */, int64_t* neg_out, void* ws, size_t ws_bytes,
                   void* stream);
int kge_shard_apply(kge_ctx*, const kge_cfg*, kge_tables*, const kge_shard*, const kge_grads* summed,
                    const float* scal_summed /* [8] device */, int64_t B, uint64_t adam_t, float* loss_out, void* ws,
                    size_t ws_bytes, void* stream);

/* Replaces the evaluate()/get_rank() loop (BaseModel.py:602-604, 620-654) for Q
 * queries at once: count_out[q] += #{e in [ent_lo, ent_hi) : score(e) > pos_score(q),
 * e not filtered}; rank = 1 + count (summed over entity shards).  count_out must be
 * zeroed by the caller.  side = KGE_SIDE_H or KGE_SIDE_T. */
int kge_rank(kge_ctx*, const kge_cfg*, const kge_tables*, const void* queries, int64_t Q,
             int side, const kge_filter* filter, int64_t ent_lo, int64_t ent_hi,
             int32_t* count_out, void* ws, size_t ws_bytes, void* stream);

/* kge_rank for callers that cannot allocate or hand over scratch memory (CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments and results as kge_rank minus the workspace; covers the CUDA-core
 * Lp-tile path only (UM / TransE / RotatE with LpDistance / LpDistancePow, else KGE_EUNSUPPORTED).  The
 * scratch (query vectors + thresholds, 4 * (ld + 1) bytes per query rounded up to 128 queries, ld = row
 * floats rounded up to 4) is a `static __device__` array of KGE_RANK_STATIC_BYTES inside the library —
 * every GPU that loads the library holds it until exit — and a call that needs more returns
 * KGE_EWORKSPACE (C1, d = 50: up to 19 712 queries per call).  The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine).  All calls share the array: it keeps the previous call's
 * contents (the part a call uses is cleared first) and calls that overlap on the GPU race — one call at a
 * time per device. */
#define KGE_RANK_STATIC_BYTES (4 << 20)
int kge_rank_lp_static(kge_ctx*, const kge_cfg*, const kge_tables*, const void* queries, int64_t Q,
                       int side, const kge_filter* filter, int64_t ent_lo, int64_t ent_hi,
                       int32_t* count_out, void* stream);

/* Filter index on the device: ONE sort of positive_X per (positive_X, side) instead of two
 * scans of it per query (BaseModel.py:646-650).
 *   kge_filter_index_bytes : bytes of the index and of the build workspace for n_pos triples.
 *   kge_filter_index_build : positive = (n_pos, 3) ids of cfg->id_dtype on the device; sorts the composite
 *                            key (r * n_ent + kept entity) * n_ent + candidate (64-bit radix sort), drops
 *                            duplicate triples and triples with ids outside the tables, and leaves
 *                            [header | pair keys | candidate entities] in `index`.  The workspace is free
 *                            again when the call's work has run; the index is read by the calls below.
 *   kge_filter_lookup      : begin_out[q], end_out[q] = the range of query q's (r, kept entity) pair in the
 *                            index; filter_out (host struct, nullable) is filled for kge_rank.
 * No host synchronisation anywhere: the number of unique triples never leaves the device.
 * n_rel * n_ent^2 must stay below 2^63. */
int kge_filter_index_bytes(const kge_cfg*, int64_t n_pos, size_t* index_bytes, size_t* ws_bytes);
int kge_filter_index_build(kge_ctx*, const kge_cfg*, const void* positive, int64_t n_pos, int side,
                           void* index, size_t index_bytes, void* ws, size_t ws_bytes, void* stream);
int kge_filter_lookup(kge_ctx*, const kge_cfg*, const void* index, int64_t n_pos, const void* queries,
                      int64_t Q, int side, int64_t* begin_out, int64_t* end_out, kge_filter* filter_out,
                      void* stream);

/* Batch pipeline on the device.  Replaces the reference's tf.data iterator
 * (KGE/data_utils.py:176-196: from_tensor_slices / csv -> shuffle -> repeat -> batch, one host -> device copy
 * per step) for training sets that fit the GPU.
 *   kge_epoch_perm_bytes : workspace bytes of kge_epoch_perm for n triples.
 *   kge_epoch_perm       : perm_out[0 .. n) = the order of pass `epoch`: STABLE argsort of the keys
 *                          Philox4x32-10(key = seed; counter = (i, epoch, bit 1)) word 0 — the sampler's
 *                          generator (oracle/philox.py: draw_u32(seed, epoch, 1, n)); stateless, so any pass
 *                          can be regenerated (resume).  n < 2^32.
 *   kge_batch_gather     : out[(B, 3)] = triples[perm[offset .. offset + B)] (perm NULL = identity), ids of
 *                          in_id_dtype in, out_id_dtype out (0 = int32, 1 = int64); offset + B <= n: a batch that
 *                          straddles two passes is two calls. */
int kge_epoch_perm_bytes(int64_t n, size_t* ws_bytes);
int kge_epoch_perm(kge_ctx*, int64_t n, uint64_t seed, uint64_t epoch, uint32_t* perm_out, void* ws,
                   size_t ws_bytes, void* stream);
int kge_batch_gather(kge_ctx*, int in_id_dtype, const void* triples, int64_t n, const uint32_t* perm,
                     int64_t offset, int64_t B, int out_id_dtype, void* out, void* stream);

/* Materialises the row-lazy tables (kge_tables.lazy_t): every row is replayed to step lazy_t and
 * last[] = lazy_t.  No-op when no table carries `last`.  One dense pass over the lazy tables
 * (read + write of w, m, v of the rows that are behind). */
int kge_flush_rows(kge_ctx*, const kge_cfg*, kge_tables*, void* stream);

/* Implementation switches (testing / A-B measurement; results are identical either way).
 * KGE_OPT_RADIX_SORT: 1 = group the step's gradient records with cub's stable radix sort
 * instead of the counting sort of csrc/group_kernels.cuh (default 0).
 * KGE_OPT_NO_SHADOW: 1 = ignore kge_table.wn and run the full-table constraint pass every batch.
 * KGE_OPT_NO_FUSED_PUSH: 1 = sharded step: the step kernel writes every gradient record locally and
 * k_shard_push delivers all rows (default 0: single-record rows are stored into the owner's inbox by
 * the step kernel itself).
 * KGE_OPT_RING_SLOTS: deepest per-warp ring (rows: 32, 16, 8, 4, or 0 = none) of negative rows the DistMult
 * step kernel stages through shared memory with TMA bulk copies (default 32).
 * KGE_OPT_GRAPH_PROFILE: 1 = step graphs created from now on record phase-boundary events (kge_step_graph_phases).
 * KGE_OPT_RANK_TC_PASSES: MMAs per K step of the tensor-core ranking GEMM: 3 (default: 3xTF32) or 2 (candidate side in
 * TF32 only, looser per-candidate bound: fewer tensor operations, more candidates for the exact fix-up).
 * KGE_OPT_CLUSTER_GROUP: 1 = group the gradient records with ONE kernel on one thread-block cluster (k_group_cluster:
 * histogram / scan / scatter through distributed shared memory) where the key space is small enough, instead of the
 * four-kernel chain (default 0: identical results, but measured slower inside a step — its 1024-thread CTAs cannot
 * run beside the step kernel).
 * KGE_OPT_PRE_EARLY: 0 = the look-ahead grouping (kge_train_step_next) starts behind the step kernel instead of beside it
 * (default 1; measured: the chain then outlasts the update kernels). */
enum { KGE_OPT_RADIX_SORT = 1, KGE_OPT_NO_SHADOW = 2, KGE_OPT_NO_FUSED_PUSH = 3, KGE_OPT_RING_SLOTS = 4,
       KGE_OPT_GRAPH_PROFILE = 5, KGE_OPT_RANK_TC_PASSES = 6, KGE_OPT_CLUSTER_GROUP = 7, KGE_OPT_PRE_EARLY = 8 };
int kge_set_option(kge_ctx*, int option, int value);

/* Per-phase device timing (CUDA events recorded on the call's stream, read back only at
 * kge_profile_end: no sync inside the timed region).  Phases: 0 constraint, 1 fused
 * sample+gather+score+loss+backward, 2 partial reduction, 3 record sort, 4 row update,
 * 5 ranking (whole call), 6 ranking: tensor-core GEMM kernel alone, 7 ranking: exact fix-up +
 * filter kernels, 8 data-parallel apply half (kge_train_apply; phase 4 is then the gradient half's
 * row sums), 9 cross-GPU barriers of kge_train_step_dp.  phase_ms[p] = summed milliseconds over the
 * n_calls recorded calls; one data-parallel step (kge_train_grad followed by kge_train_apply, or
 * kge_train_step_dp) is ONE recorded call. */
#define KGE_N_PHASES 10
int kge_profile_begin(kge_ctx*, int max_calls);
int kge_profile_end(kge_ctx*, float* phase_ms /* [KGE_N_PHASES] */, int* n_calls);

#ifdef __cplusplus
}
#endif
#endif /* KGE_ABI_H */
